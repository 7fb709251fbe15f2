"""Encoded segment pages decoded on the device in front of the scan (srb_op_push_pages, csrc/srb_pages.cuh) against the oracle's restatement of
the reference's page decoders (oracle/orc_page.cpp, pinned to be/test/util/frame_of_reference_coding_test.cpp and
be/test/storage/rowset/frame_of_reference_page_test.cpp): the decoded columns bit for bit, and whole pipelines fed from pages."""
import numpy as np
import pytest
import torch

from starrocks_b200 import capi as K
from starrocks_b200 import pages as PG
from starrocks_b200 import queries as Q
from starrocks_b200 import tpch_synth as T
from starrocks_b200.plan import ScanAggPlan, pred

pytestmark = pytest.mark.gpu

NP_OF = {K.TYPE_TINYINT: np.int8, K.TYPE_SMALLINT: np.int16, K.TYPE_INT: np.int32, K.TYPE_DATE: np.int32, K.TYPE_BIGINT: np.int64,
         K.TYPE_DECIMAL64: np.int64, K.TYPE_DECIMAL32: np.int32}


@pytest.fixture(scope="module")
def ctx():
    from starrocks_b200 import ops
    c = ops.Context(device=0)
    yield c
    c.close()


def oracle_page_column(oracle, type_, values, page_rows, precision=0, scale=0):
    """FOR pages built by the oracle's encoder (any width, int128 included)"""
    parts, pages, off = [], [], 0
    for s in range(0, max(len(values), 1), page_rows):
        p = oracle.for_encode(values[s:s + page_rows])
        parts.append(p)
        pages.append((off, len(p), len(values[s:s + page_rows])))
        off += len(p)
    raw = np.frombuffer(b"".join(parts) + b"\0" * 16, np.uint8).copy()
    return PG.PageColumn(type_, raw, off, pages, PG.ENC_FOR, precision, scale)


def torch_page_column(type_, values, page_rows, precision=0, scale=0, plain=False):
    enc, pages = (PG.plain_encode_column if plain else PG.for_encode_column)(torch.from_numpy(values), page_rows)
    raw = np.concatenate([enc.numpy(), np.zeros(16, np.uint8)])
    return PG.PageColumn(type_, raw, len(raw) - 16, pages, PG.ENC_PLAIN if plain else PG.ENC_FOR, precision, scale)


def decode_through_a_scan(ctx, columns, pushes=None):
    """a predicate-free scan operator fed with page sets -> its output columns (the decoded values)"""
    from starrocks_b200 import ops
    op = ops.ScanAggOperator(ctx, ScanAggPlan(preds=[], out_slots=list(columns.keys())))
    try:
        for cols in (pushes or [columns]):
            op.push_pages(cols)
        op.set_finishing()
        return op.pull_all()
    finally:
        op.close()


@pytest.mark.parametrize("type_,lo,hi", [(K.TYPE_TINYINT, -30, 30), (K.TYPE_SMALLINT, -8000, 8000), (K.TYPE_INT, -2**29, 2**29), (K.TYPE_DATE, 2448000, 2451500),
                                          (K.TYPE_BIGINT, -2**61, 2**61), (K.TYPE_DECIMAL64, 0, 11), (K.TYPE_DECIMAL64, 90100, 10494951)])
def test_for_pages_decode_to_the_encoded_values(ctx, oracle, type_, lo, hi):
    rng = np.random.default_rng(lo % 1000 + 3)
    dt = NP_OF[type_]
    for n, page_rows in ((1, 128), (127, 128), (128, 128), (129, 1 << 12), (70_001, 1 << 12), (300_000, 1 << 16)):
        a = rng.integers(lo, hi, n).astype(dt)
        for s in range(0, n, 640):  # some non-decreasing frames (delta format) and some constant ones (zero bits)
            a[s:s + 128] = np.sort(a[s:s + 128])
            a[s + 256:s + 384] = a[min(s + 256, n - 1)]
        col = oracle_page_column(oracle, type_, a, page_rows, 15 if type_ == K.TYPE_DECIMAL64 else 0, 2 if type_ == K.TYPE_DECIMAL64 else 0)
        got = decode_through_a_scan(ctx, {4: col})
        assert got.n_rows == n and (np.asarray(got.columns[4].data) == a).all(), (n, page_rows)
        # the same pages built by the torch builder used for the synthetic tables
        col2 = torch_page_column(type_, a, page_rows)
        assert col2.pages == col.pages and (col2.data[:col2.nbytes] == col.data[:col.nbytes]).all()


def test_int128_overflow_frames_and_plain_pages(ctx, oracle):
    rng = np.random.default_rng(8)
    n = 5000
    wide = K.to_i128_array([int(x) * 10**22 + int(y) for x, y in zip(rng.integers(-10**6, 10**6, n), rng.integers(0, 10**9, n))])
    col = oracle_page_column(oracle, K.TYPE_DECIMAL128, wide, 1 << 10, 38, 4)
    got = decode_through_a_scan(ctx, {1: col})
    d = np.asarray(got.columns[1].data)
    assert (d["lo"] == wide["lo"]).all() and (d["hi"] == wide["hi"]).all()
    # the overflow format (2) as the last frame of a page: values over the whole int32 range
    a = rng.integers(-2**31, 2**31 - 1, 100).astype(np.int32)
    got = decode_through_a_scan(ctx, {1: oracle_page_column(oracle, K.TYPE_INT, a, 128)})
    assert (np.asarray(got.columns[1].data) == a).all()
    # plain pages of every width, values at odd multiples of four bytes
    for type_, dt in ((K.TYPE_BIGINT, np.int64), (K.TYPE_INT, np.int32), (K.TYPE_SMALLINT, np.int16), (K.TYPE_DOUBLE, np.float64)):
        a = (rng.normal(0, 1e6, 10_001)).astype(dt)
        got = decode_through_a_scan(ctx, {2: torch_page_column(type_, a, 4096, plain=True)})
        assert (np.asarray(got.columns[2].data) == a).all()


def test_several_page_sets_and_host_rows_keep_their_order(ctx, oracle):
    from starrocks_b200 import ops
    rng = np.random.default_rng(9)
    n = 200_000
    a = rng.integers(0, 5000, n).astype(np.int64)
    b = rng.integers(2448000, 2451500, n).astype(np.int32)
    ca, cb = torch_page_column(K.TYPE_BIGINT, a, 1 << 12), torch_page_column(K.TYPE_DATE, b, 1 << 13)
    # page boundaries differ between the columns: split where both have one (every 8192 rows)
    cut_rows = 8192 * 10
    pa, pb = cut_rows // (1 << 12), cut_rows // (1 << 13)
    pushes = [{0: PG.slice_pages(ca, 0, pa), 1: PG.slice_pages(cb, 0, pb)}, {0: PG.slice_pages(ca, pa, len(ca.pages)), 1: PG.slice_pages(cb, pb, len(cb.pages))}]
    got = decode_through_a_scan(ctx, {0: ca, 1: cb}, pushes)
    assert (np.asarray(got.columns[0].data) == a).all() and (np.asarray(got.columns[1].data) == b).all()
    # a filter over page-fed rows, with plain host rows pushed in between
    plan = ScanAggPlan(preds=[pred(1, K.PRED_GE, K.TYPE_DATE, 2450000)], out_slots=[0, 1])
    op = ops.ScanAggOperator(ctx, plan)
    extra = K.Chunk(1000, {0: K.Column.from_ints(K.TYPE_BIGINT, a[:1000]), 1: K.Column.from_ints(K.TYPE_DATE, b[:1000])})
    op.push_pages(pushes[0])
    op.push_chunk(extra)
    op.push_pages(pushes[1])
    op.set_finishing()
    got = op.pull_all()
    op.close()
    aa = np.concatenate([a[:cut_rows], a[:1000], a[cut_rows:]])
    bb = np.concatenate([b[:cut_rows], b[:1000], b[cut_rows:]])
    keep = bb >= 2450000
    assert (np.asarray(got.columns[0].data) == aa[keep]).all() and (np.asarray(got.columns[1].data) == bb[keep]).all()


@pytest.mark.parametrize("query", ["q1", "q6"])
@pytest.mark.parametrize("where", ["host", "device"])
def test_q1_and_q6_fed_from_pages_match_the_oracle(ctx, oracle, query, where):
    from starrocks_b200 import ops
    n = 1_300_000
    cols = T.LINEITEM_Q1 if query == "q1" else T.LINEITEM_Q6
    plan_fn = Q.q1_plan if query == "q1" else Q.q6_plan
    ch = T.host_chunk("lineitem", 11, n, cols)
    want = K.chunk_rows(oracle.scan_agg(plan_fn(), ch, n_threads=4))
    pcs, ptrs = {}, []
    for slot, c in ch.columns.items():
        pc = torch_page_column(c.type, np.asarray(c.data), 1 << 16, c.precision, c.scale)
        if where == "device":
            d = ctx.to_device(pc.data)
            ptrs.append(d)
            pc = PG.PageColumn(pc.type, d, pc.nbytes, pc.pages, pc.encoding, pc.precision, pc.scale, K.MEM_DEVICE)
        pcs[slot] = pc
    try:
        op = ops.ScanAggOperator(ctx, plan_fn())
        # three page sets: pages [0, 7), [7, 8), [8, end)
        for a, b in ((0, 7), (7, 8), (8, 20)):
            op.push_pages({s: PG.slice_pages(c, a, min(b, len(c.pages))) for s, c in pcs.items()})
        op.set_finishing()
        got = K.chunk_rows(op.pull_all())
        op.close()
    finally:
        for p in ptrs:
            ctx.device_free(p)
    assert got == want
    encoded = sum(c.nbytes for c in pcs.values())
    assert encoded * 4 < sum(np.asarray(c.data).nbytes for c in ch.columns.values())  # the pages are several times smaller than the columns


def test_corrupt_pages_are_reported(ctx, oracle):
    from starrocks_b200 import ops
    a = np.arange(10_000, dtype=np.int32)[::-1].copy()
    good = torch_page_column(K.TYPE_INT, a, 1 << 12)

    def run(col):
        op = ops.ScanAggOperator(ctx, ScanAggPlan(preds=[], out_slots=[0]))
        try:
            op.push_pages({0: col})
            op.set_finishing()
            return op.pull_all()
        finally:
            op.close()

    assert (np.asarray(run(good).columns[0].data) == a).all()
    bad = PG.PageColumn(good.type, good.data.copy(), good.nbytes, good.pages)
    off, nb, nr = bad.pages[1]
    bad.data[off + nb - 4] ^= 0x10  # the footer's value count no longer matches
    with pytest.raises(ops.SrbError) as ei:
        run(bad)
    assert ei.value.code == K.CORRUPTION
    bad = PG.PageColumn(good.type, good.data.copy(), good.nbytes, good.pages)
    bad.data[off + nb - 6] = 31  # a frame claims more bits than the page holds
    with pytest.raises(ops.SrbError) as ei:
        run(bad)
    assert ei.value.code == K.CORRUPTION
    with pytest.raises(ops.SrbError) as ei:  # the page set's row count must be what the pages hold
        op = ops.ScanAggOperator(ctx, ScanAggPlan(preds=[], out_slots=[0]))
        try:
            n, slots, cols, keep = PG.to_c({0: good})
            ops._check(op.L.srb_op_push_pages(op.h, n, slots, cols, 123))
        finally:
            op.close()
    assert ei.value.code == K.INVALID_ARGUMENT


def _page_table(chunk, page_rows=1 << 12, pages_per_set=5):
    cols = {slot: torch_page_column(c.type, np.asarray(c.data), page_rows, c.precision, c.scale) for slot, c in chunk.columns.items()}
    return PG.PageTable(cols, pages_per_set)


def test_join_pipelines_fed_from_pages(ctx, oracle):
    """Q3 and SSB Q4.1 with their fact tables arriving as page tables (several page sets each) instead of chunks: the scan operators decode
    them on the device (with the runtime filters of the joins applied to the decoded rows); same result as the oracle's operator chain"""
    from starrocks_b200 import pipelines as P
    from tests import oracle_chains as OC
    n_c, n_o, n_l = 6000, 60000, 240000
    cust = T.host_chunk("customer", 0, n_c, ["c_custkey", "c_mktsegment_code"])
    orders = T.host_chunk("orders", 0, n_o, ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"], n_customers=n_c)
    line = T.host_chunk("lineitem", 0, n_l, ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"])
    want = K.chunk_rows(OC.q3_oracle(oracle, cust, orders, line))
    q = P.TpchQ3(ctx)
    try:
        got = K.chunk_rows(q.run(cust, _page_table(orders), _page_table(line)).pull_all())
    finally:
        q.close()
    assert got == want and len(got) > 100
    n_lo, n_cu, n_s, n_p = 300_000, 30_000, 2_000, 200_000
    kw = dict(n_customers=n_cu, n_suppliers=n_s, n_parts=n_p)
    dims = {name: T.host_chunk(name, 0, n, P.SSB_DIM_COLS[name]) for name, n in
            (("ssb_customer", n_cu), ("ssb_supplier", n_s), ("ssb_part", n_p), ("ssb_dates", T.SSB_DAYS))}
    lo = T.host_chunk("lineorder", 0, n_lo, T.LINEORDER_Q41, **kw)
    want = K.chunk_rows(OC.ssb_q41_oracle(oracle, dims, lo))
    for use_rf in (True, False):
        q = P.SsbQ41(ctx, runtime_filter=use_rf).build(dims)
        agg = q.probe(_page_table(lo))
        try:
            assert K.chunk_rows(agg.pull_all()) == want
        finally:
            agg.close()
            q.close()
