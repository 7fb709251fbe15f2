"""ctypes mirror of include/srb/srb_c_api.h plus small builders for plans, columns and chunks.

This is the Python half of the boundary: the structs are byte-for-byte the C PODs, the constants are the
reference's enum values (LogicalType be/src/types/logical_type.h:27-77, PredicateType
be/src/storage/column_predicate.h:66-84, TStatusCode gensrc/thrift/StatusCode.thrift:53-83).
Nothing here computes anything; compute lives in libsrb.so (CUDA) only.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np

# ---- constants ------------------------------------------------------------------------------------------
OK, CANCELLED, NOT_IMPLEMENTED, RUNTIME_ERROR, MEM_LIMIT_EXCEEDED, INTERNAL_ERROR = 0, 1, 3, 4, 5, 6
END_OF_FILE, CORRUPTION, INVALID_ARGUMENT = 30, 32, 33

TYPE_TINYINT, TYPE_SMALLINT, TYPE_INT, TYPE_BIGINT, TYPE_LARGEINT, TYPE_DOUBLE = 1, 3, 5, 7, 9, 11
TYPE_VARCHAR, TYPE_BOOLEAN, TYPE_VARBINARY = 17, 24, 46
TYPE_DECIMAL32, TYPE_DECIMAL64, TYPE_DECIMAL128, TYPE_DATE = 47, 48, 49, 50

MEM_HOST, MEM_PINNED, MEM_DEVICE = 0, 1, 2

PRED_EQ, PRED_NE, PRED_GT, PRED_GE, PRED_LT, PRED_LE, PRED_IS_NULL, PRED_NOT_NULL = 1, 2, 3, 4, 5, 6, 9, 10
EXPR_SLOT_REF, EXPR_LITERAL, EXPR_CAST, EXPR_ADD, EXPR_SUB, EXPR_MUL = 1, 2, 3, 4, 5, 6
AGG_SUM, AGG_AVG, AGG_COUNT, AGG_MIN, AGG_MAX = 1, 2, 3, 4, 5
AGG_UPDATE_FINALIZE, AGG_UPDATE_SERIALIZE, AGG_MERGE_FINALIZE, AGG_MERGE_SERIALIZE = 0, 1, 2, 3
JOIN_INNER, JOIN_LEFT_SEMI, JOIN_LEFT_ANTI = 0, 2, 8
RF_KIND_BLOOM, RF_KIND_BITSET = 1, 3  # RuntimeFilterSerializeType (exprs/runtime_filter.h:46-51)
JOIN_LEFT_OUTER, JOIN_RIGHT_OUTER, JOIN_FULL_OUTER, JOIN_RIGHT_SEMI, JOIN_RIGHT_ANTI, JOIN_NULL_AWARE_LEFT_ANTI = 1, 3, 4, 7, 9, 10
STREAMING_AUTO, STREAMING_FORCE_STREAMING, STREAMING_FORCE_PREAGGREGATION, STREAMING_LIMITED_MEM = 0, 1, 2, 3  # TStreamingPreaggregationMode

_WIDTH = {
    TYPE_TINYINT: 1, TYPE_BOOLEAN: 1, TYPE_SMALLINT: 2, TYPE_INT: 4, TYPE_DATE: 4, TYPE_DECIMAL32: 4,
    TYPE_BIGINT: 8, TYPE_DECIMAL64: 8, TYPE_DOUBLE: 8, TYPE_LARGEINT: 16, TYPE_DECIMAL128: 16,
}
_NP = {1: np.int8, 2: np.int16, 4: np.int32, 8: np.int64}


def type_width(t: int) -> int:
    return _WIDTH.get(t, 0)


def is_binary(t: int) -> bool:
    return t in (TYPE_VARCHAR, TYPE_VARBINARY)


def is_decimal(t: int) -> bool:
    return t in (TYPE_DECIMAL32, TYPE_DECIMAL64, TYPE_DECIMAL128)


# ---- C structs ------------------------------------------------------------------------------------------
class srb_column(C.Structure):
    _fields_ = [("type", C.c_int32), ("precision", C.c_int32), ("scale", C.c_int32), ("mem_kind", C.c_int32),
                ("data", C.c_void_p), ("nulls", C.c_void_p), ("offsets", C.c_void_p), ("n_bytes", C.c_int64)]


class srb_chunk(C.Structure):
    _fields_ = [("n_rows", C.c_int64), ("n_cols", C.c_int32), ("slot_ids", C.POINTER(C.c_int32)),
                ("cols", C.POINTER(srb_column))]


class srb_predicate(C.Structure):
    _fields_ = [("slot_id", C.c_int32), ("op", C.c_int32), ("type", C.c_int32), ("or_group", C.c_int32),
                ("lit_lo", C.c_int64), ("lit_hi", C.c_int64), ("lit_bytes", C.c_void_p), ("lit_len", C.c_int64)]


class srb_expr_node(C.Structure):
    _fields_ = [("op", C.c_int32), ("type", C.c_int32), ("precision", C.c_int32), ("scale", C.c_int32),
                ("lhs", C.c_int32), ("rhs", C.c_int32), ("slot_id", C.c_int32), ("_pad", C.c_int32),
                ("lit_lo", C.c_int64), ("lit_hi", C.c_int64)]


class srb_agg_fn(C.Structure):
    _fields_ = [("fn", C.c_int32), ("arg", C.c_int32)]


class srb_scan_agg_plan(C.Structure):
    _fields_ = [("n_preds", C.c_int32), ("preds", C.POINTER(srb_predicate)),
                ("n_nodes", C.c_int32), ("nodes", C.POINTER(srb_expr_node)),
                ("n_keys", C.c_int32), ("key_nodes", C.POINTER(C.c_int32)),
                ("n_aggs", C.c_int32), ("aggs", C.POINTER(srb_agg_fn)),
                ("phase", C.c_int32), ("n_out_slots", C.c_int32), ("out_slots", C.POINTER(C.c_int32)),
                ("group_hint", C.c_int64)]


class srb_join_plan(C.Structure):
    _fields_ = [("join_type", C.c_int32), ("n_keys", C.c_int32),
                ("build_key_slots", C.POINTER(C.c_int32)), ("probe_key_slots", C.POINTER(C.c_int32)),
                ("n_build_out", C.c_int32), ("build_out_slots", C.POINTER(C.c_int32)),
                ("n_probe_out", C.c_int32), ("probe_out_slots", C.POINTER(C.c_int32)),
                ("emit_indexes", C.c_int32), ("n_conjuncts", C.c_int32), ("conjuncts", C.c_void_p)]


class srb_join_conjunct(C.Structure):
    _fields_ = [("left_slot", C.c_int32), ("op", C.c_int32), ("right_slot", C.c_int32), ("_pad", C.c_int32)]


class srb_topn_plan(C.Structure):
    _fields_ = [("n_keys", C.c_int32), ("_pad", C.c_int32), ("key_slots", C.POINTER(C.c_int32)), ("is_asc", C.POINTER(C.c_uint8)),
                ("nulls_first", C.POINTER(C.c_uint8)), ("offset", C.c_int64), ("limit", C.c_int64)]


class srb_config(C.Structure):
    _fields_ = [("device", C.c_int32), ("chunk_size", C.c_int32), ("staging_rows", C.c_int64),
                ("rank", C.c_int32), ("world_size", C.c_int32), ("mem_limit", C.c_int64)]


# ---- int128 helpers --------------------------------------------------------------------------------------
I128 = np.dtype([("lo", "<u8"), ("hi", "<i8")])


def split128(x: int):
    x &= (1 << 128) - 1
    lo, hi = x & ((1 << 64) - 1), x >> 64
    if lo >= 1 << 63:
        lo -= 1 << 64
    if hi >= 1 << 63:
        hi -= 1 << 64
    return lo, hi


def join128(lo: int, hi: int) -> int:
    return (int(lo) & ((1 << 64) - 1)) | (int(hi) << 64)


def to_i128_array(values: Sequence[int]) -> np.ndarray:
    a = np.zeros(len(values), dtype=I128)
    for i, v in enumerate(values):
        lo, hi = split128(int(v))
        a["lo"][i] = lo & ((1 << 64) - 1)
        a["hi"][i] = hi
    return a


def i128_to_pyints(a: np.ndarray) -> List[int]:
    return [join128(int(lo), int(hi)) for lo, hi in zip(a["lo"].tolist(), a["hi"].tolist())]


# ---- host-side column / chunk -------------------------------------------------------------------------
@dataclass
class Column:
    """One column in the reference's layout. `data`/`nulls`/`offsets` are numpy arrays (host) or integer device
    pointers (mem_kind == MEM_DEVICE)."""
    type: int
    data: object
    nulls: object = None
    offsets: object = None
    precision: int = 0
    scale: int = 0
    mem_kind: int = MEM_HOST
    n_bytes: int = 0

    @staticmethod
    def from_ints(type_: int, values, nulls=None, precision=0, scale=0) -> "Column":
        w = type_width(type_)
        if w == 16:
            data = values if isinstance(values, np.ndarray) and values.dtype == I128 else to_i128_array(values)
        elif type_ == TYPE_DOUBLE:
            data = np.ascontiguousarray(values, dtype=np.float64)
        else:
            data = np.ascontiguousarray(values, dtype=_NP[w])
        nl = None if nulls is None else np.ascontiguousarray(nulls, dtype=np.uint8)
        return Column(type_, data, nl, None, precision, scale)

    @staticmethod
    def from_strings(values: Sequence[bytes], nulls=None, type_: int = TYPE_VARCHAR, max_len: int = 0) -> "Column":
        """max_len: the column's declared length (VARCHAR(n): TypeDescriptor::len), carried in `precision`; 0 = not declared"""
        offs = np.zeros(len(values) + 1, dtype=np.uint32)
        for i, v in enumerate(values):
            offs[i + 1] = offs[i] + len(v)
        data = np.frombuffer(b"".join(values) + b"\0" * 16, dtype=np.uint8).copy()  # 16-byte tail padding, column/bytes.h:27
        nl = None if nulls is None else np.ascontiguousarray(nulls, dtype=np.uint8)
        return Column(type_, data, nl, offs, precision=max_len, n_bytes=int(offs[-1]))

    def ptr(self, what) -> Optional[int]:
        if what is None:
            return None
        if isinstance(what, np.ndarray):
            return what.ctypes.data
        return int(what)


@dataclass
class Chunk:
    n_rows: int
    columns: Dict[int, Column] = field(default_factory=dict)

    def to_c(self):
        """Returns (srb_chunk, keepalive)."""
        n = len(self.columns)
        slots = (C.c_int32 * n)(*self.columns.keys())
        cols = (srb_column * n)()
        for i, c in enumerate(self.columns.values()):
            cols[i].type, cols[i].precision, cols[i].scale, cols[i].mem_kind = c.type, c.precision, c.scale, c.mem_kind
            cols[i].data = c.ptr(c.data)
            cols[i].nulls = c.ptr(c.nulls)
            cols[i].offsets = c.ptr(c.offsets)
            cols[i].n_bytes = c.n_bytes
        ch = srb_chunk(self.n_rows, n, slots, cols)
        return ch, (slots, cols, self)


def decode_chunk(ch: srb_chunk) -> Chunk:
    """Copy a C chunk whose buffers are host-readable into numpy-backed columns."""
    out = Chunk(int(ch.n_rows))
    n = int(ch.n_rows)
    for i in range(ch.n_cols):
        c = ch.cols[i]
        w = type_width(c.type)
        nulls = None
        if c.nulls:
            nulls = np.ctypeslib.as_array(C.cast(c.nulls, C.POINTER(C.c_uint8)), shape=(n,)).copy() if n else np.zeros(0, np.uint8)
        if is_binary(c.type):
            offs = np.ctypeslib.as_array(C.cast(c.offsets, C.POINTER(C.c_uint32)), shape=(n + 1,)).copy()
            nb = int(offs[-1])
            data = np.ctypeslib.as_array(C.cast(c.data, C.POINTER(C.c_uint8)), shape=(nb,)).copy() if nb else np.zeros(0, np.uint8)
            col = Column(c.type, data, nulls, offs, c.precision, c.scale, n_bytes=nb)
        else:
            if n == 0:
                raw = np.zeros(0, np.uint8)
            else:
                raw = np.ctypeslib.as_array(C.cast(c.data, C.POINTER(C.c_uint8)), shape=(n * w,)).copy()
            if w == 16:
                data = raw.view(I128)
            elif c.type == TYPE_DOUBLE:
                data = raw.view(np.float64)
            else:
                data = raw.view(_NP[w])
            col = Column(c.type, data, nulls, None, c.precision, c.scale)
        out.columns[int(ch.slot_ids[i])] = col
    return out


def column_pyvalues(c: Column) -> list:
    """Python values of a host column: ints, floats, bytes; None for NULL rows."""
    if is_binary(c.type):
        o = c.offsets
        vals = [bytes(c.data[int(o[i]):int(o[i + 1])]) for i in range(len(o) - 1)]
    elif type_width(c.type) == 16:
        vals = i128_to_pyints(c.data)
        vals = [v - (1 << 128) if v >= (1 << 127) else v for v in vals]
    else:
        vals = c.data.tolist()
    if c.nulls is not None:
        vals = [None if nl else v for v, nl in zip(vals, c.nulls.tolist())]
    return vals


def chunk_rows(ch: Chunk, sort: bool = True) -> list:
    """Rows as tuples in slot order; sorted for order-normalised comparison."""
    cols = [column_pyvalues(c) for c in ch.columns.values()]
    rows = list(zip(*cols)) if cols else []
    if sort:
        rows.sort(key=lambda r: tuple((x is None, 0 if x is None else x) if not isinstance(x, bytes) else (False, x) for x in r))
    return rows
