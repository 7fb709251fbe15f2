"""Counter-based synthetic TPC-H-shaped tables: value(col, row) = f(splitmix64(seed_col + row)).

Any row range of any column can be generated independently, so a GPU shard, the CPU oracle's sample and a test
all see identical bytes. Two back ends produce the same values: numpy (host) and torch (device, int64 arithmetic
with explicit logical shifts) — torch is used only to fill benchmark buffers in HBM, never on a measured path.

Distributions follow TPC-H's: quantity 1..50, extendedprice 901.00..104949.50, discount 0.00..0.10, tax 0.00..0.08,
ship dates 1992-01-02..1998-12-01, returnflag/linestatus derived from the dates (4 groups), DECIMAL(15,2) as
DECIMAL64 scaled integers, DATE as int32 Julian days.
"""
from __future__ import annotations

import numpy as np

from . import capi as K
from . import queries as Q

SEED = 0x5EED0000
MKTSEGMENTS = [b"AUTOMOBILE", b"BUILDING", b"FURNITURE", b"HOUSEHOLD", b"MACHINERY"]
RETURNFLAGS = [b"A", b"N", b"R"]
LINESTATUS = [b"F", b"O"]
REGIONS = [b"AFRICA", b"AMERICA", b"ASIA", b"EUROPE", b"MIDDLE EAST"]  # SSB: 5 regions x 5 nations, 5 manufacturers
SSB_DATE_LO = Q.from_date(1992, 1, 1)  # the dates dimension: 2556 days, 1992-01-01 .. 1998-12-30
SSB_DAYS = 2556

D_SHIP_LO = Q.from_date(1992, 1, 2)
D_SHIP_HI = Q.from_date(1998, 12, 1)
D_CUTOFF = Q.from_date(1995, 6, 17)
D_ORDER_LO = Q.from_date(1992, 1, 1)
D_ORDER_HI = Q.from_date(1998, 8, 2)

_M64 = (1 << 64) - 1


# ---- numpy back end ---------------------------------------------------------------------------------------------
def _mix_np(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = x + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _h_np(col_id: int, row0: int, n: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        rows = np.arange(row0, row0 + n, dtype=np.uint64) + np.uint64((SEED + col_id) * 0x100000001B3 & _M64)
    return _mix_np(rows)


# ---- torch back end (device) -----------------------------------------------------------------------------------------
def _lsr(x, s: int):
    return (x >> s) & ((1 << (64 - s)) - 1)


def _i64(c: int) -> int:
    c &= _M64
    return c - (1 << 64) if c >= (1 << 63) else c


def _mix_t(x):
    z = x + _i64(0x9E3779B97F4A7C15)
    z = (z ^ _lsr(z, 30)) * _i64(0xBF58476D1CE4E5B9)
    z = (z ^ _lsr(z, 27)) * _i64(0x94D049BB133111EB)
    return z ^ _lsr(z, 31)


def _h_t(col_id: int, row0: int, n: int, device):
    import torch
    rows = torch.arange(row0, row0 + n, dtype=torch.int64, device=device) + _i64((SEED + col_id) * 0x100000001B3)
    return _mix_t(rows)


def _umod_t(h, m: int):
    """unsigned 64-bit h mod m for int64-stored h (m < 2^31): split into high/low 32-bit halves"""
    hi = _lsr(h, 32)
    lo = h & 0xFFFFFFFF
    return ((hi % m) * ((1 << 32) % m) + lo % m) % m


class _Np:
    @staticmethod
    def h(col, row0, n):
        return _h_np(col, row0, n)

    @staticmethod
    def arange(row0, n):
        return np.arange(row0, row0 + n, dtype=np.int64)

    @staticmethod
    def umod(h, m):
        return (h % np.uint64(m)).astype(np.int64)

    @staticmethod
    def where(c, a, b):
        return np.where(c, a, b)

    @staticmethod
    def cast(x, dt):
        return x.astype({"i32": np.int32, "i64": np.int64, "i8": np.int8}[dt])


class _Torch:
    def __init__(self, device):
        self.device = device

    def h(self, col, row0, n):
        return _h_t(col, row0, n, self.device)

    def arange(self, row0, n):
        import torch
        return torch.arange(row0, row0 + n, dtype=torch.int64, device=self.device)

    @staticmethod
    def umod(h, m):
        return _umod_t(h, m)

    @staticmethod
    def where(c, a, b):
        import torch
        return torch.where(c, a, b)

    @staticmethod
    def cast(x, dt):
        import torch
        return x.to({"i32": torch.int32, "i64": torch.int64, "i8": torch.int8}[dt])


def _lineitem_cols(B, row0: int, n: int, cols):
    """dict name -> array for the requested lineitem columns of rows [row0, row0+n)"""
    out = {}
    ship = B.umod(B.h(5, row0, n), D_SHIP_HI - D_SHIP_LO + 1) + D_SHIP_LO
    for c in cols:
        if c == "l_quantity":
            out[c] = (B.umod(B.h(1, row0, n), 50) + 1) * 100
        elif c == "l_extendedprice":
            out[c] = B.umod(B.h(2, row0, n), 10494950 - 90100 + 1) + 90100
        elif c == "l_discount":
            out[c] = B.umod(B.h(3, row0, n), 11)
        elif c == "l_tax":
            out[c] = B.umod(B.h(4, row0, n), 9)
        elif c == "l_shipdate":
            out[c] = B.cast(ship, "i32")
        elif c == "l_returnflag_code":
            # receipt = ship + 1..30 days; 'R'/'A' (50/50) when received by the cutoff, else 'N'   codes: A=0 N=1 R=2
            h = B.h(6, row0, n)
            receipt = ship + 1 + B.umod(h, 30)
            ra = B.umod(B.h(7, row0, n), 2) * 2
            out[c] = B.cast(B.where(receipt <= D_CUTOFF, ra, ra * 0 + 1), "i32")
        elif c == "l_linestatus_code":
            out[c] = B.cast(B.where(ship <= D_CUTOFF, ship * 0, ship * 0 + 1), "i32")  # F=0 O=1
        elif c == "l_orderkey":
            # ~4 lines per order, non-decreasing; order ids drawn from the sparse o_orderkey space (8 of every 32 ids)
            idx = B.arange(row0, n) // 4
            out[c] = B.cast((idx // 8) * 32 + (idx % 8) + 1, "i32")
        else:
            raise KeyError(c)
    return out


LINEITEM_Q1 = ["l_quantity", "l_extendedprice", "l_discount", "l_tax", "l_shipdate", "l_returnflag_code", "l_linestatus_code"]
LINEITEM_Q6 = ["l_quantity", "l_extendedprice", "l_discount", "l_shipdate"]
_SLOT = {"l_orderkey": (Q.L_ORDERKEY, K.TYPE_INT, 0, 0), "l_quantity": (Q.L_QUANTITY, K.TYPE_DECIMAL64, 15, 2),
         "l_extendedprice": (Q.L_EXTENDEDPRICE, K.TYPE_DECIMAL64, 15, 2), "l_discount": (Q.L_DISCOUNT, K.TYPE_DECIMAL64, 15, 2),
         "l_tax": (Q.L_TAX, K.TYPE_DECIMAL64, 15, 2), "l_shipdate": (Q.L_SHIPDATE, K.TYPE_DATE, 0, 0),
         "l_returnflag_code": (Q.L_RETURNFLAG_CODE, K.TYPE_INT, 0, 0), "l_linestatus_code": (Q.L_LINESTATUS_CODE, K.TYPE_INT, 0, 0),
         "o_orderkey": (Q.O_ORDERKEY, K.TYPE_INT, 0, 0), "o_custkey": (Q.O_CUSTKEY, K.TYPE_INT, 0, 0),
         "o_orderdate": (Q.O_ORDERDATE, K.TYPE_DATE, 0, 0), "o_shippriority": (Q.O_SHIPPRIORITY, K.TYPE_INT, 0, 0),
         "c_custkey": (Q.C_CUSTKEY, K.TYPE_INT, 0, 0), "c_mktsegment_code": (Q.C_MKTSEGMENT_CODE, K.TYPE_TINYINT, 0, 0),
         "k": (Q.HC_K, K.TYPE_BIGINT, 0, 0), "v": (Q.HC_V, K.TYPE_BIGINT, 0, 0),
         "lo_custkey": (Q.LO_CUSTKEY, K.TYPE_INT, 0, 0), "lo_partkey": (Q.LO_PARTKEY, K.TYPE_INT, 0, 0), "lo_suppkey": (Q.LO_SUPPKEY, K.TYPE_INT, 0, 0),
         "lo_orderdate": (Q.LO_ORDERDATE, K.TYPE_INT, 0, 0), "lo_revenue": (Q.LO_REVENUE, K.TYPE_INT, 0, 0),
         "lo_supplycost": (Q.LO_SUPPLYCOST, K.TYPE_INT, 0, 0), "sc_custkey": (Q.SC_CUSTKEY, K.TYPE_INT, 0, 0),
         "sc_nation_code": (Q.SC_NATION_CODE, K.TYPE_INT, 0, 0), "sc_region_code": (Q.SC_REGION_CODE, K.TYPE_INT, 0, 0),
         "ss_suppkey": (Q.SS_SUPPKEY, K.TYPE_INT, 0, 0), "ss_region_code": (Q.SS_REGION_CODE, K.TYPE_INT, 0, 0),
         "sp_partkey": (Q.SP_PARTKEY, K.TYPE_INT, 0, 0), "sp_mfgr_code": (Q.SP_MFGR_CODE, K.TYPE_INT, 0, 0),
         "sd_datekey": (Q.SD_DATEKEY, K.TYPE_INT, 0, 0), "sd_year": (Q.SD_YEAR, K.TYPE_INT, 0, 0)}
LINEORDER_Q41 = ["lo_custkey", "lo_partkey", "lo_suppkey", "lo_orderdate", "lo_revenue", "lo_supplycost"]


def _orders_cols(B, row0, n, cols, n_customers):
    out = {}
    idx = B.arange(row0, n)
    for c in cols:
        if c == "o_orderkey":
            out[c] = B.cast((idx // 8) * 32 + (idx % 8) + 1, "i32")
        elif c == "o_custkey":
            # uniform over customers whose key is not a multiple of 3 (TPC-H rule)
            u = B.umod(B.h(21, row0, n), (n_customers // 3) * 2)
            out[c] = B.cast((u // 2) * 3 + (u % 2) + 1, "i32")
        elif c == "o_orderdate":
            out[c] = B.cast(B.umod(B.h(22, row0, n), D_ORDER_HI - D_ORDER_LO + 1) + D_ORDER_LO, "i32")
        elif c == "o_shippriority":
            out[c] = B.cast(idx * 0, "i32")
        else:
            raise KeyError(c)
    return out


def _customer_cols(B, row0, n, cols):
    out = {}
    idx = B.arange(row0, n)
    for c in cols:
        if c == "c_custkey":
            out[c] = B.cast(idx + 1, "i32")
        elif c == "c_mktsegment_code":
            out[c] = B.cast(B.umod(B.h(31, row0, n), 5), "i8")
        else:
            raise KeyError(c)
    return out


def _highcard_cols(B, row0, n, cols, n_keys):
    out = {}
    for c in cols:
        if c == "k":
            out[c] = B.umod(B.h(40, row0, n), n_keys) if n_keys < (1 << 31) else None
        elif c == "v":
            out[c] = B.umod(B.h(41, row0, n), 1 << 20)
        else:
            raise KeyError(c)
    return out


def _ssb_datekey(B, day):
    """yyyymmdd of SSB_DATE_LO + day (Fliegel-Van Flandern inverse of from_date), and the year"""
    j = day + SSB_DATE_LO + 32044
    g = j // 146097
    dg = j % 146097
    c = (dg // 36524 + 1) * 3 // 4
    dc = dg - c * 36524
    b = dc // 1461
    db = dc % 1461
    a = (db // 365 + 1) * 3 // 4
    da = db - a * 365
    y = g * 400 + c * 100 + b * 4 + a
    m = (da * 5 + 308) // 153 - 2
    d = da - (m + 4) * 153 // 5 + 122
    year = y - 4800 + (m + 2) // 12
    month = (m + 2) % 12 + 1
    return year * 10000 + month * 100 + d + 1, year


def _ssb_cols(table, B, row0, n, cols, kw):
    out = {}
    idx = B.arange(row0, n)
    n_cust, n_supp, n_part = kw.get("n_customers", 30000), kw.get("n_suppliers", 2000), kw.get("n_parts", 200000)
    for c in cols:
        if c == "lo_custkey":
            out[c] = B.cast(B.umod(B.h(50, row0, n), n_cust) + 1, "i32")
        elif c == "lo_partkey":
            out[c] = B.cast(B.umod(B.h(51, row0, n), n_part) + 1, "i32")
        elif c == "lo_suppkey":
            out[c] = B.cast(B.umod(B.h(52, row0, n), n_supp) + 1, "i32")
        elif c == "lo_orderdate":
            out[c] = B.cast(_ssb_datekey(B, B.umod(B.h(53, row0, n), SSB_DAYS))[0], "i32")
        elif c == "lo_revenue":
            out[c] = B.cast(B.umod(B.h(54, row0, n), 10_000_000) + 90_000, "i32")
        elif c == "lo_supplycost":
            out[c] = B.cast(B.umod(B.h(55, row0, n), 120_000) + 6_000, "i32")
        elif c in ("sc_custkey", "ss_suppkey", "sp_partkey"):
            out[c] = B.cast(idx + 1, "i32")
        elif c == "sc_nation_code":  # nation = region * 5 + 0..4
            out[c] = B.cast(B.umod(B.h(60, row0, n), 25), "i32")
        elif c == "sc_region_code":
            out[c] = B.cast(B.umod(B.h(60, row0, n), 25) // 5, "i32")
        elif c == "ss_region_code":
            out[c] = B.cast(B.umod(B.h(65, row0, n), 25) // 5, "i32")
        elif c == "sp_mfgr_code":
            out[c] = B.cast(B.umod(B.h(70, row0, n), 5), "i32")
        elif c == "sd_datekey":
            out[c] = B.cast(_ssb_datekey(B, idx)[0], "i32")
        elif c == "sd_year":
            out[c] = B.cast(_ssb_datekey(B, idx)[1], "i32")
        else:
            raise KeyError(c)
    return out


def _gen(table: str, B, row0: int, n: int, cols, **kw):
    if table in ("lineorder", "ssb_customer", "ssb_supplier", "ssb_part", "ssb_dates"):
        return _ssb_cols(table, B, row0, n, cols, kw)
    if table == "lineitem":
        return _lineitem_cols(B, row0, n, cols)
    if table == "orders":
        return _orders_cols(B, row0, n, cols, kw.get("n_customers", 150000))
    if table == "customer":
        return _customer_cols(B, row0, n, cols)
    if table == "highcard":
        return _highcard_cols(B, row0, n, cols, kw.get("n_keys", 1000))
    raise KeyError(table)


def host_chunk(table: str, row0: int, n: int, cols, **kw) -> K.Chunk:
    """numpy-backed Chunk of rows [row0, row0+n) (slots per queries.py)"""
    arrs = _gen(table, _Np, row0, n, cols, **kw)
    ch = K.Chunk(n)
    for name in cols:
        slot, t, p, s = _SLOT[name]
        ch.columns[slot] = K.Column.from_ints(t, arrs[name], None, p, s)
    return ch


def device_chunk(table: str, row0: int, n: int, cols, device, piece: int = 1 << 26, **kw):
    """Chunk of MEM_DEVICE columns backed by torch tensors (returned too, to keep them alive). Generated piecewise so the
    temporaries stay small next to the table itself."""
    import torch
    tensors = {}
    for name in cols:
        slot, t, p, s = _SLOT[name]
        dt = {1: torch.int8, 4: torch.int32, 8: torch.int64}[K.type_width(t)]
        tensors[name] = torch.empty(n, dtype=dt, device=device)
    B = _Torch(device)
    for a in range(0, n, piece):
        m = min(piece, n - a)
        part = _gen(table, B, row0 + a, m, cols, **kw)
        for name in cols:
            tensors[name][a:a + m] = part[name]
        del part
    ch = K.Chunk(n)
    for name in cols:
        slot, t, p, s = _SLOT[name]
        ch.columns[slot] = K.Column(t, tensors[name].data_ptr(), None, None, p, s, K.MEM_DEVICE)
    return ch, tensors


def add_binary_keys(ch: K.Chunk) -> K.Chunk:
    """Adds l_returnflag / l_linestatus as VARCHAR(1) BinaryColumns decoded from the code columns (host chunks only)."""
    rf = ch.columns[Q.L_RETURNFLAG_CODE].data
    ls = ch.columns[Q.L_LINESTATUS_CODE].data
    n = ch.n_rows
    offs = np.arange(n + 1, dtype=np.uint32)
    for slot, codes, names in ((Q.L_RETURNFLAG, rf, RETURNFLAGS), (Q.L_LINESTATUS, ls, LINESTATUS)):
        lut = np.frombuffer(b"".join(names), dtype=np.uint8)
        data = np.concatenate([lut[codes], np.zeros(16, np.uint8)])
        ch.columns[slot] = K.Column(K.TYPE_VARCHAR, data, None, offs.copy(), n_bytes=n)
    return ch
