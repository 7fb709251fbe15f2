"""GPU parity: every result of libsrb.so (through the C ABI) is compared with the CPU oracle on the same inputs.
Bit-exact for integer / DECIMAL / key / row-set outputs (order-normalised); 1e-9 relative for DOUBLE.
The known-answer vectors of the reference's own unit tests (be/test/exprs/agg/aggregate_test.cpp,
be/test/exec/join_hash_map_test.cpp) are replayed through the CUDA path as well."""
import math
import os

import numpy as np
import pytest

from starrocks_b200 import capi as K
from tests import oracle_chains as OC
from starrocks_b200 import queries as Q
from starrocks_b200 import tpch_synth as T
from starrocks_b200.plan import ExprBuilder, JoinPlan, ScanAggPlan, pred, pred_in

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from starrocks_b200 import ops
    c = ops.Context(device=0, chunk_size=4096)
    yield c
    c.close()


def rows_equal(a, b, rel=1e-9):
    if len(a) != len(b):
        return False
    for ra, rb in zip(a, b):
        if len(ra) != len(rb):
            return False
        for x, y in zip(ra, rb):
            if isinstance(x, float) or isinstance(y, float):
                if x is None or y is None:
                    if x is not y:
                        return False
                elif not (x == y or abs(x - y) <= rel * max(abs(x), abs(y))):
                    return False
            elif x != y:
                return False
    return True


def run_gpu(ctx, plan, chunk, piece_rows=0, device=False):
    from starrocks_b200 import ops
    if device:
        dch, ptrs = ctx.device_chunk(chunk)
        try:
            return ops.run_scan_agg(ctx, plan, dch)
        finally:
            for p in ptrs:
                ctx.device_free(p)
    return ops.run_scan_agg(ctx, plan, chunk, piece_rows)


def _decode_blobs(rows):
    """AVG intermediate blobs (avg.h:168-181) -> (sum, count) so that DOUBLE sums compare within the tolerance"""
    import struct
    out = []
    for r in rows:
        rr = []
        for x in r:
            if isinstance(x, bytes) and len(x) == 16:
                x = ("avg", struct.unpack("<d", x[:8])[0], struct.unpack("<q", x[8:])[0])
            rr.append(x)
        out.append(tuple(rr))
    return out


def _flat(rows):
    return [tuple(y for x in r for y in (x if isinstance(x, tuple) else (x,))) for r in rows]


def check(ctx, oracle, plan, chunk, sort=True, **kw):
    want = oracle.scan_agg(plan, chunk)
    got = run_gpu(ctx, plan, chunk, **kw)
    assert list(got.columns.keys()) == list(want.columns.keys())
    for s in want.columns:
        assert got.columns[s].type == want.columns[s].type, s
        assert got.columns[s].scale == want.columns[s].scale, s
    a, b = _flat(_decode_blobs(K.chunk_rows(got, sort))), _flat(_decode_blobs(K.chunk_rows(want, sort)))
    assert rows_equal(a, b), (a[:5], b[:5])
    return got


# ---- the reference's aggregate known-answer vectors, through the CUDA path ------------------------------------------------
COL1 = list(range(1024)) + [100, 200]
COL2 = list(range(2000, 3000))


def _agg_plan(fn, type_, p=0, s=0, phase=K.AGG_UPDATE_FINALIZE):
    e = ExprBuilder()
    return ScanAggPlan(exprs=e, aggs=[(fn, e.slot(0, type_, p, s))], phase=phase)


def _one(ch):
    return K.column_pyvalues(ch.columns[0])[0]


def _gpu_agg(ctx, fn, type_, vals, p=0, s=0, nulls=None, phase=K.AGG_UPDATE_FINALIZE, **kw):
    ch = K.Chunk(len(vals), {0: K.Column.from_ints(type_, vals, nulls, p, s)})
    return run_gpu(ctx, _agg_plan(fn, type_, p, s, phase), ch, **kw)


def _gpu_two_phase(ctx, fn, type_, v1, v2, p=0, s=0):
    inter = [_gpu_agg(ctx, fn, type_, v, p, s, phase=K.AGG_UPDATE_SERIALIZE) for v in (v1, v2)]
    c0, c1 = inter[0].columns[0], inter[1].columns[0]
    if K.is_binary(c0.type):
        col = K.Column.from_strings(K.column_pyvalues(c0) + K.column_pyvalues(c1), type_=K.TYPE_VARBINARY)
        col.precision, col.scale = c0.precision, c0.scale
    else:
        col = K.Column.from_ints(c0.type, K.column_pyvalues(c0) + K.column_pyvalues(c1), None, c0.precision, c0.scale)
    e = ExprBuilder()
    plan = ScanAggPlan(exprs=e, aggs=[(fn, e.slot(0, col.type, col.precision, col.scale))], phase=K.AGG_MERGE_FINALIZE)
    return _one(run_gpu(ctx, plan, K.Chunk(2, {0: col})))


@pytest.mark.parametrize("type_", [K.TYPE_SMALLINT, K.TYPE_INT, K.TYPE_BIGINT, K.TYPE_LARGEINT, K.TYPE_DOUBLE])
def test_reference_agg_vectors(ctx, type_):
    # aggregate_test.cpp:331-367,399-414,957-1012
    assert _one(_gpu_agg(ctx, K.AGG_COUNT, type_, COL1)) == 1026
    assert _gpu_two_phase(ctx, K.AGG_COUNT, type_, COL1, COL2) == 2026
    assert _one(_gpu_agg(ctx, K.AGG_SUM, type_, COL1)) == 524076
    assert _one(_gpu_agg(ctx, K.AGG_SUM, type_, COL2)) == 2499500
    assert _gpu_two_phase(ctx, K.AGG_SUM, type_, COL1, COL2) == 3023576
    assert _one(_gpu_agg(ctx, K.AGG_AVG, type_, COL1)) == 524076 / 1026.0
    assert _gpu_two_phase(ctx, K.AGG_AVG, type_, COL1, COL2) == 3023576 / 2026.0
    assert _one(_gpu_agg(ctx, K.AGG_MAX, type_, COL1)) == 1023
    assert _gpu_two_phase(ctx, K.AGG_MAX, type_, COL1, COL2) == 2999
    assert _one(_gpu_agg(ctx, K.AGG_MIN, type_, COL2)) == 2000
    assert _gpu_two_phase(ctx, K.AGG_MIN, type_, COL1, COL2) == 0


@pytest.mark.parametrize("type_,p,s", [(K.TYPE_DECIMAL32, 9, 9), (K.TYPE_DECIMAL64, 9, 3), (K.TYPE_DECIMAL128, 38, 15)])
def test_reference_decimal_vectors(ctx, oracle, type_, p, s):
    # aggregate_test.cpp:369-397,431-468
    r = _gpu_agg(ctx, K.AGG_SUM, type_, COL1, p, s)
    assert r.columns[0].type == K.TYPE_DECIMAL128 and r.columns[0].scale == s and _one(r) == 524076
    assert _gpu_two_phase(ctx, K.AGG_SUM, type_, COL1, COL2, p, s) == 3023576
    ddi = oracle.decimal_div_integer
    assert _one(_gpu_agg(ctx, K.AGG_AVG, type_, COL1, p, s)) == ddi(524076, 1026, s)
    assert _one(_gpu_agg(ctx, K.AGG_AVG, type_, COL2, p, s)) == ddi(2499500, 1000, s)
    assert _gpu_two_phase(ctx, K.AGG_AVG, type_, COL1, COL2, p, s) == ddi(3023576, 2026, s)


def test_reference_date_avg_and_nullable(ctx):
    d1 = [Q.from_date(2000 + j, 1, 1) for j in range(20)] + [Q.from_date(2000, 1, 1), Q.from_date(2001, 1, 1)]
    assert _one(_gpu_agg(ctx, K.AGG_AVG, K.TYPE_DATE, d1)) == 2454716  # aggregate_test.cpp:428
    assert _one(_gpu_agg(ctx, K.AGG_MAX, K.TYPE_DATE, d1)) == Q.from_date(2019, 1, 1)
    vals = list(range(100))
    nulls = [1 if i % 3 == 0 else 0 for i in range(100)]
    assert _one(_gpu_agg(ctx, K.AGG_SUM, K.TYPE_INT, vals, nulls=nulls)) == sum(v for v, n in zip(vals, nulls) if not n)
    assert _one(_gpu_agg(ctx, K.AGG_COUNT, K.TYPE_INT, vals, nulls=nulls)) == 100 - sum(nulls)
    assert _one(_gpu_agg(ctx, K.AGG_SUM, K.TYPE_INT, vals, nulls=[1] * 100)) is None  # aggregate_test.cpp:1250-1297
    assert _one(_gpu_agg(ctx, K.AGG_COUNT, K.TYPE_INT, vals, nulls=[1] * 100)) == 0


def test_empty_input_no_group_by(ctx):
    # aggregator.cpp:938-945: one row, SUM NULL, COUNT 0 — with a chunk of zero rows and with no chunk at all
    from starrocks_b200 import ops
    assert _one(_gpu_agg(ctx, K.AGG_SUM, K.TYPE_INT, [])) is None
    assert _one(_gpu_agg(ctx, K.AGG_COUNT, K.TYPE_INT, [])) == 0
    op = ops.ScanAggOperator(ctx, _agg_plan(K.AGG_MIN, K.TYPE_BIGINT))
    op.set_finishing()
    out = op.pull_all()
    op.close()
    assert out.n_rows == 1 and _one(out) is None


# ---- predicates, expressions, keys: randomised plans against the oracle --------------------------------------------------
def _rand_col(rng, type_, n, null_frac=0.0, small=False, p=0, s=0):
    w = K.type_width(type_)
    if type_ == K.TYPE_DOUBLE:
        vals = rng.normal(0, 1e6, n)
    elif w == 16:
        vals = [int(rng.integers(-2**62, 2**62)) * int(rng.integers(1, 2**40)) for _ in range(n)]
        if small:
            vals = [int(x) for x in rng.integers(-50, 50, n)]
    else:
        hi = 50 if small else 2 ** (8 * w - 2)
        vals = rng.integers(-hi, hi, n)
    nulls = (rng.random(n) < null_frac).astype(np.uint8) if null_frac > 0 else None
    return K.Column.from_ints(type_, vals, nulls, p, s)


@pytest.mark.parametrize("type_", [K.TYPE_TINYINT, K.TYPE_SMALLINT, K.TYPE_INT, K.TYPE_BIGINT, K.TYPE_DATE])
def test_reference_predicate_and_filter_vectors(ctx, type_):
    """the reference's own known answers (be/test/storage/column_predicate_test.cpp test_gt/ge/lt/le/is_null/not_null over the nullable
    column {1, 2, 3, NULL, 4}; be/test/column/fixed_length_column_test.cpp:58-74, 130-163 filter vectors) through the CUDA filter path,
    host and device chunks"""
    from tests.test_oracle_golden import PREDICATE_VECTORS, predicate_vector_chunk, selection_of, check_filter_vectors
    for device in (False, True):
        for op, want in PREDICATE_VECTORS:
            got = run_gpu(ctx, Q.filter_plan([pred(0, op, type_, 3)], [1]), predicate_vector_chunk(type_), device=device)
            assert selection_of(got, 5) == want, (op, device)
        if type_ == K.TYPE_INT:
            check_filter_vectors(lambda plan, ch: run_gpu(ctx, plan, ch, device=device))


@pytest.mark.parametrize("type_", [K.TYPE_TINYINT, K.TYPE_SMALLINT, K.TYPE_INT, K.TYPE_BIGINT, K.TYPE_DATE, K.TYPE_DECIMAL32,
                                   K.TYPE_DECIMAL64, K.TYPE_DECIMAL128, K.TYPE_DOUBLE])
def test_predicates_all_ops(ctx, oracle, type_):
    rng = np.random.default_rng(type_)
    n = 20000
    col = _rand_col(rng, type_, n, null_frac=0.1, small=True)
    ch = K.Chunk(n, {0: col, 1: K.Column.from_ints(K.TYPE_BIGINT, np.arange(n))})
    e = ExprBuilder()
    arg = e.slot(1, K.TYPE_BIGINT)
    for op in (K.PRED_EQ, K.PRED_NE, K.PRED_GT, K.PRED_GE, K.PRED_LT, K.PRED_LE, K.PRED_IS_NULL, K.PRED_NOT_NULL):
        lit = 3.5 if type_ == K.TYPE_DOUBLE else 3
        plan = ScanAggPlan(preds=[pred(0, op, type_, lit)], exprs=e, aggs=[(K.AGG_SUM, arg), (K.AGG_COUNT, None)])
        check(ctx, oracle, plan, ch)


def test_predicate_conjunction_and_or_groups(ctx, oracle):
    rng = np.random.default_rng(5)
    n = 50000
    ch = K.Chunk(n, {0: _rand_col(rng, K.TYPE_INT, n, 0.05, small=True), 1: _rand_col(rng, K.TYPE_BIGINT, n, 0.05, small=True),
                     2: _rand_col(rng, K.TYPE_DECIMAL128, n, 0.0, small=True), 3: K.Column.from_ints(K.TYPE_BIGINT, np.arange(n))})
    e = ExprBuilder()
    arg = e.slot(3, K.TYPE_BIGINT)
    preds = [pred(0, K.PRED_GE, K.TYPE_INT, -20), pred(1, K.PRED_LT, K.TYPE_BIGINT, 5, or_group=7), pred(2, K.PRED_EQ, K.TYPE_DECIMAL128, 3, or_group=7),
             pred(0, K.PRED_NE, K.TYPE_INT, 0), pred(1, K.PRED_IS_NULL, K.TYPE_BIGINT, 0, or_group=9), pred(2, K.PRED_GT, K.TYPE_DECIMAL128, -10, or_group=9)]
    plan = ScanAggPlan(preds=preds, exprs=e, aggs=[(K.AGG_SUM, arg), (K.AGG_COUNT, None), (K.AGG_MIN, arg), (K.AGG_MAX, arg)])
    check(ctx, oracle, plan, ch)
    check(ctx, oracle, plan, ch, device=True)


def test_string_predicates(ctx, oracle):
    # BinaryColumnPredicateCmpBase (storage/column_predicate_cmp.cpp:559-700): every comparison on a VARCHAR column, NULL rows, empty
    # strings, prefixes, IN lists of strings; the reference's own vectors first (column_predicate_test.cpp:564-660)
    ids = lambda n: K.Column.from_ints(K.TYPE_INT, np.arange(n, dtype=np.int32))
    keep = lambda plan, c, **kw: [r[0] for r in K.chunk_rows(run_gpu(ctx, plan, c, **kw), sort=False)]
    ch = K.Chunk(3, {0: K.Column.from_strings([b"aa", b"ab", b"ac"]), 1: ids(3)})
    assert keep(ScanAggPlan(preds=[pred(0, K.PRED_EQ, K.TYPE_VARCHAR, b"aa")], out_slots=[1]), ch) == [0]
    ch = K.Chunk(5, {0: K.Column.from_strings([b"xxx", b"yyy", b"abc", b"", b"abc"], nulls=np.array([0, 0, 0, 1, 0], np.uint8)), 1: ids(5)})
    assert keep(ScanAggPlan(preds=[pred(0, K.PRED_EQ, K.TYPE_VARCHAR, b"abc")], out_slots=[1]), ch) == [2, 4]
    rng = np.random.default_rng(16)
    n = 150_001
    words = [b"", b"a", b"ab", b"abc", b"abd", b"AUTOMOBILE", b"BUILDING", b"BUILDINGS", b"BUILD", b"MACHINERY", b"zz", b"0123456789abcdef", b"0123456789abcdefX"]
    vals = [words[i] for i in rng.integers(0, len(words), n)]
    ch = K.Chunk(n, {0: K.Column.from_strings(vals, nulls=(rng.random(n) < 0.05).astype(np.uint8)), 1: ids(n), 2: _rand_col(rng, K.TYPE_BIGINT, n, 0.0, small=True)})
    for op in (K.PRED_EQ, K.PRED_NE, K.PRED_LT, K.PRED_LE, K.PRED_GT, K.PRED_GE):
        for lit in (b"BUILDING", b"", b"abc", b"0123456789abcdef"):
            plan = ScanAggPlan(preds=[pred(0, op, K.TYPE_VARCHAR, lit)], out_slots=[1, 2])
            check(ctx, oracle, plan, ch, sort=False)
    plan = ScanAggPlan(preds=pred_in(0, K.TYPE_VARCHAR, [b"BUILDING", b"zz", b"a"], or_group=1) + [pred(2, K.PRED_GT, K.TYPE_BIGINT, 0), pred(0, K.PRED_NOT_NULL, K.TYPE_VARCHAR)],
                       out_slots=[1, 2])
    check(ctx, oracle, plan, ch, sort=False)
    check(ctx, oracle, plan, ch, sort=False, device=True)
    check(ctx, oracle, plan, ch, sort=False, piece_rows=4096)
    e = ExprBuilder()
    aggp = ScanAggPlan(preds=[pred(0, K.PRED_GE, K.TYPE_VARCHAR, b"BUILD")], exprs=e, keys=[e.slot(0, K.TYPE_VARCHAR)], aggs=[(K.AGG_SUM, e.slot(2, K.TYPE_BIGINT)), (K.AGG_COUNT, None)])
    check(ctx, oracle, aggp, ch)
    from starrocks_b200 import ops
    with pytest.raises(ops.SrbError) as ei:  # literals ride inside the lowered program: at most 16 bytes
        run_gpu(ctx, ScanAggPlan(preds=[pred(0, K.PRED_EQ, K.TYPE_VARCHAR, b"0123456789abcdefX")], out_slots=[1]), ch)
    assert ei.value.code == K.NOT_IMPLEMENTED


def test_in_list_and_not_in_predicates(ctx, oracle):
    # ColumnInPredicate / ColumnNotInPredicate (storage/column_predicate.h:249-277): IN = OR of EQ, NOT IN = AND of NE; NULL rows never pass
    rng = np.random.default_rng(14)
    n = 200_003
    ch = K.Chunk(n, {0: _rand_col(rng, K.TYPE_INT, n, 0.1, small=True), 1: _rand_col(rng, K.TYPE_BIGINT, n, 0.0, small=True),
                     2: _rand_col(rng, K.TYPE_DECIMAL128, n, 0.05, small=True, p=27, s=2)})
    for preds in (pred_in(0, K.TYPE_INT, [3, -7, 11, 49], or_group=1),
                  pred_in(0, K.TYPE_INT, [3, -7, 11], or_group=1) + pred_in(1, K.TYPE_BIGINT, [0, 1, 2, 3, 4], or_group=2),
                  pred_in(1, K.TYPE_BIGINT, [5, 6, 7, -1], 0, negate=True) + [pred(0, K.PRED_GT, K.TYPE_INT, -10)],
                  pred_in(2, K.TYPE_DECIMAL128, [1, 2, 3, -40], or_group=3)):
        check(ctx, oracle, ScanAggPlan(preds=preds, out_slots=[0, 1, 2]), ch, sort=False)
        check(ctx, oracle, ScanAggPlan(preds=preds, out_slots=[0, 1, 2]), ch, sort=False, device=True)
        e = ExprBuilder()
        plan = ScanAggPlan(preds=preds, exprs=e, keys=[e.slot(0, K.TYPE_INT)], aggs=[(K.AGG_SUM, e.slot(1, K.TYPE_BIGINT)), (K.AGG_COUNT, None)])
        check(ctx, oracle, plan, ch, device=True)


def test_having_is_a_filter_over_the_aggregate_result_on_the_device(ctx, oracle):
    # aggregate_blocking_source_operator.cpp:62-68: HAVING conjuncts run over the emitted groups
    from starrocks_b200 import ops, pipelines as P
    rng = np.random.default_rng(15)
    n = 300_000
    ch = K.Chunk(n, {0: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 5000, n).astype(np.int32)),
                     1: K.Column.from_ints(K.TYPE_DECIMAL64, rng.integers(-1000, 100000, n), (rng.random(n) < 0.01).astype(np.uint8), 15, 2)})
    e = ExprBuilder()
    plan = ScanAggPlan(exprs=e, keys=[e.slot(0, K.TYPE_INT)], aggs=[(K.AGG_SUM, e.slot(1, K.TYPE_DECIMAL64, 15, 2)), (K.AGG_COUNT, None)], group_hint=5000)
    groups = oracle.scan_agg(plan, ch)
    # HAVING sum(v) > 3,100,000 (DECIMAL128 scaled) AND count(*) <> 60
    hv = [pred(1, K.PRED_GT, K.TYPE_DECIMAL128, 3_100_000), pred(2, K.PRED_NE, K.TYPE_BIGINT, 60)]
    want = K.chunk_rows(oracle.scan_agg(ScanAggPlan(preds=hv, out_slots=[0, 1, 2]), groups))
    assert 0 < len(want) < groups.n_rows
    dch, ptrs = ctx.device_chunk(ch)
    agg = ops.ScanAggOperator(ctx, plan)
    try:
        agg.push_chunk(dch)
        agg.set_finishing()
        f = P.having(ctx, agg, hv, [0, 1, 2])
        try:
            assert K.chunk_rows(f.pull_all()) == want
        finally:
            f.close()
    finally:
        agg.close()
        for p in ptrs:
            ctx.device_free(p)


def test_filter_only_compaction_preserves_order(ctx, oracle):
    # Chunk::filter (column/chunk.cpp:302): surviving rows keep their order, every column compacted by the same mask
    rng = np.random.default_rng(11)
    n = 100003
    ch = K.Chunk(n, {0: _rand_col(rng, K.TYPE_INT, n, 0.1, small=True), 1: _rand_col(rng, K.TYPE_DECIMAL128, n, 0.2),
                     2: _rand_col(rng, K.TYPE_SMALLINT, n), 3: _rand_col(rng, K.TYPE_DOUBLE, n), 4: _rand_col(rng, K.TYPE_TINYINT, n, 0.3)})
    for preds in ([pred(0, K.PRED_GT, K.TYPE_INT, 0)], [pred(0, K.PRED_GT, K.TYPE_INT, 48)], [pred(0, K.PRED_GE, K.TYPE_INT, -1000)], []):
        plan = ScanAggPlan(preds=preds, out_slots=[1, 0, 2, 3, 4])
        check(ctx, oracle, plan, ch, sort=False)
        check(ctx, oracle, plan, ch, sort=False, piece_rows=4096)
        check(ctx, oracle, plan, ch, sort=False, device=True)


DEC_EDGE = [0, 1, -1, 99, -100, 2**31 - 1, -2**31, 2**63 - 1, -2**63, 2**64, 10**18, -10**18, 10**37, -10**37, 2**126, -2**126, 2**127 - 1, -2**127]


@pytest.mark.parametrize("rtype", [K.TYPE_DECIMAL32, K.TYPE_DECIMAL64, K.TYPE_DECIMAL128])
def test_decimal_arithmetic_overflow_to_null(ctx, oracle, rtype):
    # decimal_binary_function.h:94-109: an overflowing row becomes NULL and is skipped by the nullable aggregate
    w = K.type_width(rtype)
    lim = 2 ** (8 * w - 1)
    edge = [x for x in DEC_EDGE if -lim <= x < lim]
    rng = np.random.default_rng(w)
    a = edge * len(edge) + [int(x) for x in rng.integers(-1000, 1000, 500)]
    b = [y for y in edge for _ in edge] + [int(x) for x in rng.integers(-1000, 1000, 500)]
    n = len(a)
    ch = K.Chunk(n, {0: K.Column.from_ints(rtype, a, None, 9, 2), 1: K.Column.from_ints(rtype, b, None, 9, 1),
                     2: K.Column.from_ints(K.TYPE_INT, np.arange(n) % 7)})
    for op in ("add", "sub", "mul"):
        e = ExprBuilder()
        x, y, k = e.slot(0, rtype, 9, 2), e.slot(1, rtype, 9, 1), e.slot(2, K.TYPE_INT)
        r = getattr(e, op)(x, y, rtype)
        plan = ScanAggPlan(exprs=e, keys=[k], aggs=[(K.AGG_SUM, r), (K.AGG_COUNT, r), (K.AGG_MIN, r), (K.AGG_MAX, r), (K.AGG_AVG, r)])
        check(ctx, oracle, plan, ch)
        plan2 = ScanAggPlan(exprs=e, aggs=[(K.AGG_SUM, r), (K.AGG_COUNT, r), (K.AGG_MIN, r), (K.AGG_MAX, r)])
        check(ctx, oracle, plan2, ch)


def test_cast_and_fast_mul(ctx, oracle):
    rng = np.random.default_rng(3)
    n = 30000
    a = [int(x) for x in rng.integers(-2**62, 2**62, n)]
    b = [int(x) for x in rng.integers(-2**62, 2**62, n)]
    a[:4] = [2**63 - 1, -2**63, -2**63, 2**63 - 1]
    b[:4] = [2**63 - 1, -2**63, 2**63 - 1, -1]
    ch = K.Chunk(n, {0: K.Column.from_ints(K.TYPE_DECIMAL64, a, None, 18, 2), 1: K.Column.from_ints(K.TYPE_DECIMAL64, b, None, 18, 2),
                     2: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 5, n))})
    e = ExprBuilder()
    x, y, k = e.slot(0, K.TYPE_DECIMAL64, 18, 2), e.slot(1, K.TYPE_DECIMAL64, 18, 2), e.slot(2, K.TYPE_INT)
    prod = e.mul(e.cast(x, K.TYPE_DECIMAL128, 18, 2), e.cast(y, K.TYPE_DECIMAL128, 18, 2), K.TYPE_DECIMAL128, 38)  # fast path
    up = e.cast(x, K.TYPE_DECIMAL128, 38, 10)  # scale-up by 10^8 (checked)
    up64 = e.cast(y, K.TYPE_DECIMAL64, 18, 6)  # overflows int64 for most rows -> NULL
    plan = ScanAggPlan(exprs=e, keys=[k], aggs=[(K.AGG_SUM, prod), (K.AGG_SUM, up), (K.AGG_COUNT, up64), (K.AGG_MAX, prod), (K.AGG_MIN, up)])
    check(ctx, oracle, plan, ch)


def test_integer_arithmetic_wraps(ctx, oracle):
    rng = np.random.default_rng(4)
    n = 10000
    ch = K.Chunk(n, {0: _rand_col(rng, K.TYPE_INT, n), 1: _rand_col(rng, K.TYPE_INT, n), 2: _rand_col(rng, K.TYPE_BIGINT, n)})
    e = ExprBuilder()
    x, y, z = e.slot(0, K.TYPE_INT), e.slot(1, K.TYPE_INT), e.slot(2, K.TYPE_BIGINT)
    plan = ScanAggPlan(exprs=e, aggs=[(K.AGG_SUM, e.mul(x, y)), (K.AGG_SUM, e.add(x, y)), (K.AGG_SUM, e.mul(z, z)), (K.AGG_MIN, e.sub(z, e.cast(x, K.TYPE_BIGINT)))])
    check(ctx, oracle, plan, ch)


def test_low_cardinality_group_by_through_the_interpreter_shared_states(ctx, oracle, monkeypatch):
    """a few groups through the interpreter: rows are added into the CTA's shared-memory mirror of the table's states and flushed
    once per CTA — carries out of the low 64 bits (DECIMAL128 sums of large positive and negative values), AVG, COUNT of a nullable
    column and a BIGINT sum that wraps; the same rows through per-row table atomics (SRB_NO_SHARED_AGG) give the same result"""
    rng = np.random.default_rng(77)
    n = 300_000
    big = rng.integers(-(1 << 62), 1 << 62, n).astype(np.int64)
    big[::2] = np.abs(big[::2]) | (1 << 62)  # every other row adds ~2^62: the low word carries every few rows
    nulls = (rng.random(n) < 0.1).astype(np.uint8)
    ch = K.Chunk(n, {0: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 7, n).astype(np.int32)),
                     1: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 5, n).astype(np.int32)),
                     2: K.Column.from_ints(K.TYPE_DECIMAL64, big, None, 18, 2),
                     3: K.Column.from_ints(K.TYPE_BIGINT, big, nulls),
                     4: K.Column.from_ints(K.TYPE_INT, rng.integers(-1000, 1000, n).astype(np.int32))})
    e = ExprBuilder()
    k0, k1 = e.slot(0, K.TYPE_INT), e.slot(1, K.TYPE_INT)
    d, b, i = e.slot(2, K.TYPE_DECIMAL64, 18, 2), e.slot(3, K.TYPE_BIGINT), e.slot(4, K.TYPE_INT)
    plan = ScanAggPlan(exprs=e, keys=[k0, k1], aggs=[(K.AGG_SUM, d), (K.AGG_AVG, d), (K.AGG_SUM, b), (K.AGG_COUNT, b), (K.AGG_COUNT, None),
                                                      (K.AGG_SUM, e.sub(e.cast(i, K.TYPE_BIGINT), b, K.TYPE_BIGINT))], group_hint=64)
    got = K.chunk_rows(check(ctx, oracle, plan, ch))
    assert len(got) == 35
    monkeypatch.setenv("SRB_NO_SHARED_AGG", "1")
    assert K.chunk_rows(run_gpu(ctx, plan, ch)) == got
    monkeypatch.delenv("SRB_NO_SHARED_AGG")


@pytest.mark.parametrize("groups", [3, 4, 11])
def test_low_cardinality_group_by_with_thread_private_accumulators(ctx, oracle, monkeypatch, groups):
    """a handful of groups through the interpreter (k_scan_agg_lowcard): every aggregate function — SUM / AVG of DECIMAL64 with carries,
    SUM of a wrapping BIGINT, COUNT of a nullable column, MIN / MAX of DECIMAL128, DATE and DOUBLE, SUM / AVG of DOUBLE — over nullable
    arguments, with more groups than the hint promised (rows beyond the CTA's directory update the table directly), a predicate, and
    VARCHAR(1) keys; same result as the oracle and as the row-atomic path (SRB_NO_LOWCARD)"""
    rng = np.random.default_rng(79 + groups)
    n = 700_001
    big = rng.integers(-(1 << 62), 1 << 62, n).astype(np.int64)
    big[::2] = np.abs(big[::2]) | (1 << 62)
    wide = K.to_i128_array([int(x) * 10**19 for x in rng.integers(-10**6, 10**6, n)])
    codes = rng.integers(0, groups, n)
    ch = K.Chunk(n, {0: K.Column.from_ints(K.TYPE_INT, codes.astype(np.int32)),
                     1: K.Column.from_ints(K.TYPE_DECIMAL64, big, (rng.random(n) < 0.05).astype(np.uint8), 18, 2),
                     2: K.Column.from_ints(K.TYPE_BIGINT, big, (rng.random(n) < 0.1).astype(np.uint8)),
                     3: K.Column.from_ints(K.TYPE_DECIMAL128, wide, (rng.random(n) < 0.2).astype(np.uint8), 38, 3),
                     4: K.Column.from_ints(K.TYPE_DATE, rng.integers(2440000, 2460000, n).astype(np.int32)),
                     5: K.Column.from_ints(K.TYPE_DOUBLE, rng.normal(0, 1e3, n), (rng.random(n) < 0.01).astype(np.uint8)),
                     6: K.Column.from_strings([bytes([65 + int(c)]) for c in codes])})
    for key_slot, key_type in ((0, K.TYPE_INT), (6, K.TYPE_VARCHAR)):
        e = ExprBuilder()
        k0 = e.slot(key_slot, key_type)
        d, b, w, dt, f = e.slot(1, K.TYPE_DECIMAL64, 18, 2), e.slot(2, K.TYPE_BIGINT), e.slot(3, K.TYPE_DECIMAL128, 38, 3), e.slot(4, K.TYPE_DATE), e.slot(5, K.TYPE_DOUBLE)
        aggs = [(K.AGG_SUM, d), (K.AGG_AVG, d), (K.AGG_SUM, b), (K.AGG_COUNT, b), (K.AGG_COUNT, None), (K.AGG_MIN, w), (K.AGG_MAX, w), (K.AGG_MIN, dt),
                (K.AGG_MAX, f), (K.AGG_SUM, f), (K.AGG_AVG, f)]
        plan = ScanAggPlan(preds=[pred(4, K.PRED_GE, K.TYPE_DATE, 2441000)], exprs=e, keys=[k0], aggs=aggs, group_hint=3)
        monkeypatch.setenv("SRB_LOWCARD", "1")  # (opt-in: measured slower than the shared mirror, kept for A/B)
        got = check(ctx, oracle, plan, ch, device=True)
        assert got.n_rows == groups
        monkeypatch.delenv("SRB_LOWCARD")
        other = run_gpu(ctx, plan, ch, device=True)  # the default: the shared-memory mirror (MIN / MAX of DATE and DOUBLE, DOUBLE sums included)
        assert rows_equal(_flat(_decode_blobs(K.chunk_rows(other))), _flat(_decode_blobs(K.chunk_rows(got))))


@pytest.mark.parametrize("keys", [[K.TYPE_INT], [K.TYPE_BIGINT], [K.TYPE_DATE], [K.TYPE_TINYINT, K.TYPE_SMALLINT], [K.TYPE_INT, K.TYPE_INT],
                                  [K.TYPE_INT, K.TYPE_DATE, K.TYPE_INT], [K.TYPE_DECIMAL128], [K.TYPE_BIGINT, K.TYPE_INT]])
@pytest.mark.parametrize("null_frac", [0.0, 0.2])
def test_group_by_key_shapes(ctx, oracle, keys, null_frac):
    # AggHashMap variants (agg_hash_variant.h:40-105): single number key, fixed-size packed keys, nullable keys
    rng = np.random.default_rng(len(keys) * 10 + int(null_frac * 10))
    n = 40000
    cols = {i: _rand_col(rng, t, n, null_frac, small=True) for i, t in enumerate(keys)}
    cols[10] = _rand_col(rng, K.TYPE_DECIMAL64, n, 0.1, p=15, s=2)
    cols[11] = _rand_col(rng, K.TYPE_DOUBLE, n, 0.1)
    ch = K.Chunk(n, cols)
    e = ExprBuilder()
    ks = [e.slot(i, t) for i, t in enumerate(keys)]
    m, d = e.slot(10, K.TYPE_DECIMAL64, 15, 2), e.slot(11, K.TYPE_DOUBLE)
    aggs = [(K.AGG_SUM, m), (K.AGG_AVG, m), (K.AGG_COUNT, m), (K.AGG_COUNT, None), (K.AGG_MIN, m), (K.AGG_MAX, m), (K.AGG_SUM, d), (K.AGG_AVG, d),
            (K.AGG_MIN, d), (K.AGG_MAX, d)]
    plan = ScanAggPlan(exprs=e, keys=ks, aggs=aggs)
    # (a nullable DECIMAL128 key packs to 17 bytes: a wide key, three words)
    check(ctx, oracle, plan, ch)
    check(ctx, oracle, plan, ch, piece_rows=4096)


def test_group_by_binary_keys(ctx, oracle):
    # AggHashMapWithSerializedKey (agg_hash_map.h:652-898): two VARCHAR(1) keys serialise to 10 bytes
    rng = np.random.default_rng(8)
    n = 30000
    a = [bytes([65 + int(x)]) for x in rng.integers(0, 3, n)]
    b = [b"" if x == 0 else bytes([70 + int(x)]) * int(x) for x in rng.integers(0, 3, n)]
    nulls = (rng.random(n) < 0.1).astype(np.uint8)
    ch = K.Chunk(n, {0: K.Column.from_strings(a), 1: K.Column.from_strings(b, nulls), 2: _rand_col(rng, K.TYPE_BIGINT, n, small=True)})
    e = ExprBuilder()
    k1, k2, v = e.slot(0, K.TYPE_VARCHAR), e.slot(1, K.TYPE_VARCHAR), e.slot(2, K.TYPE_BIGINT)
    plan = ScanAggPlan(exprs=e, keys=[k1, k2], aggs=[(K.AGG_SUM, v), (K.AGG_COUNT, None)])
    check(ctx, oracle, plan, ch)
    check(ctx, oracle, plan, ch, piece_rows=1000)
    check(ctx, oracle, plan, ch, device=True)


WIDE_KEY_CASES = {
    "three BIGINT (24 bytes)": [K.TYPE_BIGINT, K.TYPE_BIGINT, K.TYPE_BIGINT],
    "four BIGINT (32 bytes)": [K.TYPE_BIGINT] * 4,
    "nullable BIGINT, DECIMAL128, INT (30 bytes)": [(K.TYPE_BIGINT, True), K.TYPE_DECIMAL128, (K.TYPE_INT, True)],
    "VARCHAR(40) and INT (up to 48 bytes)": [("str", 40, True), K.TYPE_INT],
    "two VARCHAR(20) (up to 48 bytes)": [("str", 20, False), ("str", 20, False)],
    "eight words: DECIMAL128 x 4": [K.TYPE_DECIMAL128] * 4,
}


def _wide_key_chunk(rng, spec, n, n_groups):
    """key columns drawn from a pool of n_groups distinct key tuples + a BIGINT and a DECIMAL64 measure"""
    pool = []
    for _ in range(n_groups):
        t = []
        for s_ in spec:
            if isinstance(s_, tuple) and s_[0] == "str":
                t.append(bytes(rng.integers(97, 123, int(rng.integers(0, s_[1] + 1))).astype(np.uint8)))
            else:
                ty = s_[0] if isinstance(s_, tuple) else s_
                t.append(int(rng.integers(-(1 << 62), 1 << 62)) * (10**15 if ty == K.TYPE_DECIMAL128 else 1) if ty != K.TYPE_INT else int(rng.integers(-1000, 1000)))
        pool.append(t)
    pick = rng.integers(0, n_groups, n)
    cols = {}
    for i, s_ in enumerate(spec):
        vals = [pool[g][i] for g in pick]
        if isinstance(s_, tuple) and s_[0] == "str":
            cols[i] = K.Column.from_strings(vals, (rng.random(n) < 0.03).astype(np.uint8) if s_[2] else None, max_len=s_[1])
        else:
            ty, nullable = s_ if isinstance(s_, tuple) else (s_, False)
            cols[i] = K.Column.from_ints(ty, vals, (rng.random(n) < 0.03).astype(np.uint8) if nullable else None, 38 if ty == K.TYPE_DECIMAL128 else 0, 0)
    cols[20] = K.Column.from_ints(K.TYPE_BIGINT, rng.integers(-1000, 1000, n))
    cols[21] = K.Column.from_ints(K.TYPE_DECIMAL64, rng.integers(0, 10**6, n), (rng.random(n) < 0.05).astype(np.uint8), 15, 2)
    return K.Chunk(n, cols)


def _wide_key_plan(spec, phase=K.AGG_UPDATE_FINALIZE, hint=0):
    e = ExprBuilder()
    keys = []
    for i, s_ in enumerate(spec):
        if isinstance(s_, tuple) and s_[0] == "str":
            keys.append(e.slot(i, K.TYPE_VARCHAR))
        else:
            ty = s_[0] if isinstance(s_, tuple) else s_
            keys.append(e.slot(i, ty, 38 if ty == K.TYPE_DECIMAL128 else 0, 0))
    v, d = e.slot(20, K.TYPE_BIGINT), e.slot(21, K.TYPE_DECIMAL64, 15, 2)
    return ScanAggPlan(preds=[pred(20, K.PRED_GE, K.TYPE_BIGINT, -900)], exprs=e, keys=keys, phase=phase, group_hint=hint,
                       aggs=[(K.AGG_SUM, v), (K.AGG_COUNT, None), (K.AGG_AVG, d), (K.AGG_MIN, v), (K.AGG_MAX, d)])


@pytest.mark.parametrize("name", list(WIDE_KEY_CASES))
def test_group_keys_beyond_16_bytes(ctx, oracle, name):
    # AggHashMapWithSerializedKey (agg_hash_map.h:652-898) with 17-64 serialized bytes: the slot carries the further key words behind pair B;
    # few groups (every row on the same slots), many groups (the table grows from 1024 slots: rehash + replay), host pieces and device chunks
    spec = WIDE_KEY_CASES[name]
    rng = np.random.default_rng(31)
    few = _wide_key_chunk(rng, spec, 40_000, 50)
    check(ctx, oracle, _wide_key_plan(spec, hint=50), few)
    check(ctx, oracle, _wide_key_plan(spec), few, device=True)
    many = _wide_key_chunk(rng, spec, 200_000, 60_000)
    got = check(ctx, oracle, _wide_key_plan(spec), many)
    assert got.n_rows > 30_000
    check(ctx, oracle, _wide_key_plan(spec), many, piece_rows=30_000)
    check(ctx, oracle, _wide_key_plan(spec, hint=100_000), many, device=True)


def test_wide_group_keys_serialize_phase_and_too_long(ctx, oracle):
    from starrocks_b200 import ops
    spec = WIDE_KEY_CASES["VARCHAR(40) and INT (up to 48 bytes)"]
    rng = np.random.default_rng(32)
    ch = _wide_key_chunk(rng, spec, 60_000, 3000)
    want = K.chunk_rows(oracle.scan_agg(_wide_key_plan(spec), ch))
    check(ctx, oracle, _wide_key_plan(spec, phase=K.AGG_UPDATE_SERIALIZE), ch)  # intermediate rows (AVG blobs) keyed by the wide keys
    # a string longer than the declared length does not fit the key area: reported, not truncated
    bad = K.Chunk(3, {0: K.Column.from_strings([b"a", b"b" * 41, b"c"], max_len=40), 1: K.Column.from_ints(K.TYPE_INT, np.arange(3, dtype=np.int32)),
                      20: K.Column.from_ints(K.TYPE_BIGINT, np.arange(3)), 21: K.Column.from_ints(K.TYPE_DECIMAL64, np.arange(3), np.zeros(3, np.uint8), 15, 2)})
    with pytest.raises(ops.SrbError) as ex:
        run_gpu(ctx, _wide_key_plan(spec), bad)
    assert ex.value.code == K.NOT_IMPLEMENTED
    assert len(want) > 2000


def test_table_growth_and_deferred_rows(ctx, oracle):
    # no cardinality hint: the table starts at 1024 slots and must grow (rehash) while rows are deferred and replayed
    rng = np.random.default_rng(9)
    n = 600000
    keys = rng.integers(0, 200000, n)
    ch = K.Chunk(n, {0: K.Column.from_ints(K.TYPE_BIGINT, keys), 1: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(-1000, 1000, n))})
    e = ExprBuilder()
    k, v = e.slot(0, K.TYPE_BIGINT), e.slot(1, K.TYPE_BIGINT)
    plan = ScanAggPlan(exprs=e, keys=[k], aggs=[(K.AGG_SUM, v), (K.AGG_COUNT, None), (K.AGG_MIN, v)])
    got = check(ctx, oracle, plan, ch)
    assert got.n_rows == len(np.unique(keys))
    check(ctx, oracle, plan, ch, piece_rows=50000)


def _push_device_chunks(ctx, plan, chunks):
    from starrocks_b200 import ops
    op = ops.ScanAggOperator(ctx, plan)
    ptrs = []
    try:
        for ch in chunks:
            dch, pp = ctx.device_chunk(ch)
            ptrs += pp
            op.push_chunk(dch)
        op.set_finishing()
        return op.pull_all()
    finally:
        op.close()
        for q in ptrs:
            ctx.device_free(q)


def test_deferral_bitmaps_are_clean_when_device_chunks_grow(ctx, oracle):
    # regression (round-1 advice): the two deferral bitmaps come dirty from the buffer pool with a rounded-up size; a second, larger
    # device chunk that still fits must not see stale bits when its rows are deferred and replayed (rows would be counted twice)
    rng = np.random.default_rng(77)
    e = ExprBuilder()
    k, v = e.slot(0, K.TYPE_BIGINT), e.slot(1, K.TYPE_BIGINT)
    plan = ScanAggPlan(exprs=e, keys=[k], aggs=[(K.AGG_SUM, v), (K.AGG_COUNT, None)])
    for attempt in range(3):
        # dirty 16 KB buffers into the pool: filter-only results of 4096 all-ones INT rows
        for _ in range(3):
            ones = K.Chunk(4096, {0: K.Column.from_ints(K.TYPE_INT, np.full(4096, -1, np.int32))})
            _push_device_chunks(ctx, ScanAggPlan(preds=[], out_slots=[0]), [ones])
        n1, n2 = 100_000, 120_000
        c1 = K.Chunk(n1, {0: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(0, 100, n1)), 1: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(0, 1000, n1))})
        c2 = K.Chunk(n2, {0: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(0, 50_000, n2)), 1: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(0, 1000, n2))})
        got = _push_device_chunks(ctx, plan, [c1, c2])
        both = K.Chunk(n1 + n2, {s: K.Column.from_ints(K.TYPE_BIGINT, np.concatenate([c1.columns[s].data, c2.columns[s].data])) for s in (0, 1)})
        want = oracle.scan_agg(plan, both)
        assert K.chunk_rows(got, True) == K.chunk_rows(want, True)


def test_filter_only_device_chunk_without_the_null_map_of_an_earlier_one(ctx, oracle):
    # regression (round-1 advice): a slot that was nullable in the first device chunk arrives with nulls == NULL later:
    # its rows are all non-NULL and the result's null flags must say so
    rng = np.random.default_rng(78)
    n = 50_000
    plan = ScanAggPlan(preds=[pred(0, K.PRED_GT, K.TYPE_INT, 10)], out_slots=[1, 0])
    a = K.Chunk(n, {0: _rand_col(rng, K.TYPE_INT, n, 0.0, small=True), 1: _rand_col(rng, K.TYPE_BIGINT, n, 0.3)})
    b = K.Chunk(n, {0: _rand_col(rng, K.TYPE_INT, n, 0.0, small=True), 1: _rand_col(rng, K.TYPE_BIGINT, n, 0.0)})
    assert a.columns[1].nulls is not None and b.columns[1].nulls is None
    # dirty buffers of the result's size classes into the pool first
    _push_device_chunks(ctx, ScanAggPlan(preds=[], out_slots=[1, 0]), [K.Chunk(n, {0: a.columns[0], 1: K.Column.from_ints(K.TYPE_BIGINT, a.columns[1].data, np.ones(n, np.uint8))})])
    got = _push_device_chunks(ctx, plan, [a, b])
    both = K.Chunk(2 * n, {0: K.Column.from_ints(K.TYPE_INT, np.concatenate([a.columns[0].data, b.columns[0].data])),
                           1: K.Column.from_ints(K.TYPE_BIGINT, np.concatenate([a.columns[1].data, b.columns[1].data]),
                                                 np.concatenate([a.columns[1].nulls, np.zeros(n, np.uint8)]))})
    want = oracle.scan_agg(plan, both)
    assert K.chunk_rows(got, False) == K.chunk_rows(want, False)


def test_all_phases_with_groups(ctx, oracle):
    rng = np.random.default_rng(10)
    n = 20000
    ch = K.Chunk(n, {0: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 50, n)), 1: _rand_col(rng, K.TYPE_DECIMAL64, n, 0.1, p=15, s=2),
                     2: _rand_col(rng, K.TYPE_DOUBLE, n, 0.1), 3: _rand_col(rng, K.TYPE_INT, n, 0.1)})
    e = ExprBuilder()
    k, m, d, i = e.slot(0, K.TYPE_INT), e.slot(1, K.TYPE_DECIMAL64, 15, 2), e.slot(2, K.TYPE_DOUBLE), e.slot(3, K.TYPE_INT)
    aggs = [(K.AGG_SUM, m), (K.AGG_AVG, m), (K.AGG_COUNT, m), (K.AGG_MIN, m), (K.AGG_AVG, d), (K.AGG_SUM, d), (K.AGG_AVG, i), (K.AGG_MAX, i), (K.AGG_COUNT, None)]
    p1 = ScanAggPlan(exprs=e, keys=[k], aggs=aggs, phase=K.AGG_UPDATE_SERIALIZE)
    halves = [K.Chunk(n // 2, {s: K.Column(c.type, c.data[:n // 2], None if c.nulls is None else c.nulls[:n // 2], None, c.precision, c.scale)
                               for s, c in ch.columns.items()}),
              K.Chunk(n - n // 2, {s: K.Column(c.type, c.data[n // 2:], None if c.nulls is None else c.nulls[n // 2:], None, c.precision, c.scale)
                                   for s, c in ch.columns.items()})]
    inter = [check(ctx, oracle, p1, h) for h in halves]
    # concatenate the two intermediate chunks and merge them (phase 2)
    cols = {}
    for s in inter[0].columns:
        c0, c1 = inter[0].columns[s], inter[1].columns[s]
        vals = K.column_pyvalues(c0) + K.column_pyvalues(c1)
        nulls = [1 if v is None else 0 for v in vals]
        if K.is_binary(c0.type):
            col = K.Column.from_strings([b"" if v is None else v for v in vals], nulls, type_=c0.type)
        elif c0.type == K.TYPE_DOUBLE:
            col = K.Column.from_ints(c0.type, [0.0 if v is None else v for v in vals], nulls)
        else:
            col = K.Column.from_ints(c0.type, [0 if v is None else v for v in vals], nulls)
        col.precision, col.scale = c0.precision, c0.scale
        cols[s] = col
    merged_in = K.Chunk(inter[0].n_rows + inter[1].n_rows, cols)
    e2 = ExprBuilder()
    k2 = e2.slot(0, K.TYPE_INT)
    args = [e2.slot(1 + a, cols[1 + a].type, cols[1 + a].precision, cols[1 + a].scale) for a in range(len(aggs))]
    p2 = ScanAggPlan(exprs=e2, keys=[k2], aggs=[(fn, args[a]) for a, (fn, _) in enumerate(aggs)], phase=K.AGG_MERGE_FINALIZE)
    two_phase = check(ctx, oracle, p2, merged_in)
    single = run_gpu(ctx, ScanAggPlan(exprs=e, keys=[k], aggs=aggs), ch)
    assert rows_equal(K.chunk_rows(two_phase), K.chunk_rows(single))


# ---- the benchmark pipelines on synthetic TPC-H-shaped data: static (TMA) kernels vs generic kernel vs oracle ---------------
def _both_kernel_families(ctx, oracle, plan_fn, chunk, **kw):
    want = K.chunk_rows(oracle.scan_agg(plan_fn(), chunk, n_threads=4))
    outs = []
    # static kernel (narrow tile body where the data allows), static kernel with the general body only, generic interpreter
    for env in (None, "SRB_NO_NARROW", "SRB_FORCE_GENERIC"):
        if env:
            os.environ[env] = "1"
        try:
            for dev in (True, False):
                outs.append(K.chunk_rows(run_gpu(ctx, plan_fn(), chunk, device=dev, **kw)))
        finally:
            if env:
                os.environ.pop(env, None)
    for o in outs:
        assert rows_equal(o, want), (o[:2], want[:2])


@pytest.mark.parametrize("n", [1024, 5000, 300001, 2_000_000])
def test_q6_pipeline(ctx, oracle, n):
    _both_kernel_families(ctx, oracle, Q.q6_plan, T.host_chunk("lineitem", 12345, n, T.LINEITEM_Q6))


@pytest.mark.parametrize("n", [1024, 70000, 1_500_000])
def test_q1_pipeline_dict_keys(ctx, oracle, n):
    ch = T.host_chunk("lineitem", 777, n, T.LINEITEM_Q1)
    _both_kernel_families(ctx, oracle, Q.q1_plan, ch)
    got = run_gpu(ctx, Q.q1_plan(), ch, device=True)
    assert got.n_rows == 4 or n < 5000


def test_q1_pipeline_binary_keys(ctx, oracle):
    ch = T.add_binary_keys(T.host_chunk("lineitem", 0, 200000, T.LINEITEM_Q1))
    plan = Q.q1_plan(dict_keys=False)
    got = check(ctx, oracle, plan, ch)
    assert got.n_rows == 4
    check(ctx, oracle, plan, ch, piece_rows=4096)


def test_q1_more_groups_than_register_banks(ctx, oracle):
    # the static Q1 kernel keeps 4 groups in registers; extra groups must take the global-table path and still be exact
    ch = T.host_chunk("lineitem", 0, 400000, T.LINEITEM_Q1)
    rf = ch.columns[Q.L_RETURNFLAG_CODE].data
    rf[:] = np.arange(len(rf)) % 11
    _both_kernel_families(ctx, oracle, Q.q1_plan, ch)


def test_q1_overflowing_charge_is_null(ctx, oracle):
    ch = T.host_chunk("lineitem", 0, 100000, T.LINEITEM_Q1)
    ext = ch.columns[Q.L_EXTENDEDPRICE].data
    ext[::7] = 2**62  # ext*(1-disc) fits int128, times (1+tax) overflows -> NULL rows inside sum(charge)
    ext[5::11] = -2**62
    _both_kernel_families(ctx, oracle, Q.q1_plan, ch)


@pytest.mark.parametrize("case", ["wide_rows", "negative", "discount_above_one", "carries", "carries_filtered"])
def test_q1_narrow_body_escapes_and_carries(ctx, oracle, case):
    # The static Q1 kernel runs a narrow tile body (32x32->64 products, plain 64-bit accumulation) while every value of a warp is a
    # small non-negative integer, and redoes the tile with the general body otherwise. Values around the borders of that assumption:
    n = 1_500_000
    ch = T.host_chunk("lineitem", 31, n, T.LINEITEM_Q1)
    qty, ext = ch.columns[Q.L_QUANTITY].data, ch.columns[Q.L_EXTENDEDPRICE].data
    disc, tax = ch.columns[Q.L_DISCOUNT].data, ch.columns[Q.L_TAX].data
    if case == "wide_rows":          # a few rows per many warps leave [0, 2^32)
        ext[::1999] = 2**32
        ext[7::4001] = 2**40 + 12345
        qty[11::997] = 2**33
    elif case == "negative":
        qty[::313] = -qty[::313]
        ext[5::1201] = -ext[5::1201]
    elif case == "discount_above_one":  # 1 - l_discount < 0
        disc[::777] = 250
    else:
        # products just below 2^64: every accumulator word wraps many times (carries recovered after the fact)
        ext[:] = 42_000_000
        disc[:] = 0
        tax[:] = 4_100_000_000
        if case == "carries_filtered":  # half the rows fail the shipdate predicate: their adds land in the trash word
            ch.columns[Q.L_SHIPDATE].data[::2] = Q.from_date(1999, 1, 1)
    _both_kernel_families(ctx, oracle, Q.q1_plan, ch)


@pytest.mark.parametrize("n,nk", [(1024, 100), (500000, 1000), (3_000_000, 1_000_000)])
def test_highcard_pipeline(ctx, oracle, n, nk):
    ch = T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=nk)
    _both_kernel_families(ctx, oracle, lambda: Q.highcard_plan(group_hint=0), ch)
    _both_kernel_families(ctx, oracle, lambda: Q.highcard_plan(group_hint=nk), ch)


def test_highcard_l2_blocked_partition_path(ctx, oracle, monkeypatch):
    # a table far larger than L2 (2^24 slots x 32 B = 512 MB) and a device batch >= 4 Mi rows take the partition-then-aggregate
    # path (srb_hcpart.cuh; the reference's answer to the same wall is its two-level maps, agg_hash_map.h:92-98)
    from starrocks_b200 import ops
    n, nk = 6_000_000, 5_000_000
    ch = T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=nk)
    want = K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), ch))
    dch, ptrs = ctx.device_chunk(ch)
    try:
        launches = {}
        for mode in ("partitioned", "direct"):
            if mode == "direct":
                monkeypatch.setenv("SRB_NO_PARTITION", "1")
            op = ops.ScanAggOperator(ctx, Q.highcard_plan(group_hint=nk))
            op.push_chunk(dch)
            op.set_finishing()
            got = K.chunk_rows(op.pull_all())
            launches[mode] = op.counters()["kernel_launches"]
            op.close()
            assert rows_equal(got, want), mode
        assert launches["partitioned"] == launches["direct"] + 3  # histogram, scan, scatter
    finally:
        monkeypatch.delenv("SRB_NO_PARTITION", raising=False)
        for p in ptrs:
            ctx.device_free(p)


def test_highcard_merge_phase_static_shape_and_l2_blocked_path(ctx, oracle, monkeypatch):
    """phase 2 of the two-phase high-cardinality GROUP BY (the tuples a rank receives from the shuffle): intermediate (k, sum, count)
    rows from two phase-1 operators are merged by the statically specialised merge kernel — through the partition-then-aggregate path
    when the table is far larger than L2, directly otherwise, and by the generic interpreter — all equal to one-phase aggregation"""
    from starrocks_b200 import ops
    n, nk = 9_000_000, 5_000_000
    ch = T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=nk)
    want = K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), ch))
    inters = []
    for a, b in ((0, n // 2), (n // 2, n)):
        op = ops.ScanAggOperator(ctx, Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE, group_hint=nk))
        op.push_chunk(ops.slice_chunk(ch, a, b))
        op.set_finishing()
        inters.append(ctx.fetch_chunk(op.pull_device()))
        op.close()
    from starrocks_b200 import distributed as D
    inter = D.concat_chunks(inters)
    assert inter.n_rows > (1 << 22)
    # the serialized SUM column is nullable (NullableAggregateFunctionUnary) but carries no NULL: the map is dropped per batch
    sum_col = [c for c in inter.columns.values() if c.nulls is not None]
    assert sum_col and not any(c.nulls.any() for c in sum_col)
    dch, ptrs = ctx.device_chunk(inter)
    try:
        seen = {}
        for mode, env in (("partitioned", None), ("direct", "SRB_NO_PARTITION"), ("generic", "SRB_FORCE_GENERIC")):
            if env:
                monkeypatch.setenv(env, "1")
            op = ops.ScanAggOperator(ctx, Q.highcard_merge_plan(group_hint=nk))
            op.push_chunk(dch)
            op.set_finishing()
            got = K.chunk_rows(op.pull_all())
            prof = op.profile()
            seen[mode] = (op.counters()["kernel_launches"], prof["shape"], prof["static_launches"])
            op.close()
            if env:
                monkeypatch.delenv(env)
            assert rows_equal(got, want), mode
        assert seen["partitioned"][1] == 4 and seen["direct"][1] == 4
        assert seen["partitioned"][2] >= 1 and seen["direct"][2] >= 1 and seen["generic"][2] == 0
        assert seen["partitioned"][0] == seen["direct"][0] + 3  # histogram, scan, scatter
    finally:
        for p in ptrs:
            ctx.device_free(p)
    # partial sums that really are NULL (a phase-1 group whose every value was NULL): the batch keeps its null map, is not
    # permuted, and the interpreter merges it — equal to the oracle's merge of the same tuples
    for c in sum_col:
        c.nulls[::7] = 1
    want_n = K.chunk_rows(oracle.scan_agg(Q.highcard_merge_plan(), inter))
    op = ops.ScanAggOperator(ctx, Q.highcard_merge_plan(group_hint=nk))
    op.push_chunk(inter)
    op.set_finishing()
    got_n = K.chunk_rows(op.pull_all())
    assert op.profile()["static_launches"] == 0
    op.close()
    assert rows_equal(got_n, want_n) and not rows_equal(got_n, want)


def test_host_chunks_of_4096_rows_like_the_pipeline_driver(ctx, oracle):
    # pipeline_driver.cpp:345-351: chunks of <= chunk_size rows; results must not depend on the chunking
    ch = T.host_chunk("lineitem", 99, 123457, T.LINEITEM_Q1)
    want = K.chunk_rows(oracle.scan_agg(Q.q1_plan(), ch))
    for piece in (4096, 1000, 50000):
        assert rows_equal(K.chunk_rows(run_gpu(ctx, Q.q1_plan(), ch, piece_rows=piece)), want)


def test_pull_chunk_slices_of_chunk_size(ctx, oracle):
    from starrocks_b200 import ops
    ch = T.host_chunk("highcard", 0, 100000, ["k", "v"], n_keys=10000)
    op = ops.ScanAggOperator(ctx, Q.highcard_plan())
    assert op.need_input() and not op.has_output()
    op.push_chunk(ch)
    op.set_finishing()
    op.set_finishing()  # idempotent (operator.h:66)
    rows, n_chunks = [], 0
    while op.has_output():
        c = op.pull_chunk()
        assert c is not None and 0 < c.n_rows <= 4096
        rows += K.chunk_rows(c, sort=False)
        n_chunks += 1
    assert op.is_finished() and op.pull_chunk() is None
    cnt = op.counters()
    assert cnt["rows_in"] == 100000 and cnt["rows_selected"] == 100000 and cnt["kernel_launches"] > 0
    op.close()
    assert n_chunks == math.ceil(len(rows) / 4096)
    assert sorted(rows) == K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), ch))


def test_error_convention(ctx):
    from starrocks_b200 import ops
    op = ops.ScanAggOperator(ctx, Q.q6_plan())
    with pytest.raises(ops.SrbError) as ei:
        op.pull_chunk()  # blocking aggregate: no output before set_finishing
    assert ei.value.code == K.INTERNAL_ERROR
    op.set_finishing()
    with pytest.raises(ops.SrbError):
        op.push_chunk(T.host_chunk("lineitem", 0, 10, T.LINEITEM_Q6))
    op.close()
    op = ops.ScanAggOperator(ctx, Q.q6_plan())
    with pytest.raises(ops.SrbError) as ei:
        op.push_chunk(K.Chunk(3, {0: K.Column.from_ints(K.TYPE_INT, [1, 2, 3])}))  # slots of the plan missing
    assert ei.value.code == K.INVALID_ARGUMENT
    op.set_cancelled()
    assert op.is_finished() and not op.has_output() and not op.need_input()
    op.close()
    e = ExprBuilder()
    long_keys = [e.slot(i, K.TYPE_DECIMAL128, 38, 0) for i in range(5)]  # an 80-byte packed key: beyond the 64 bytes of a wide key
    op = ops.ScanAggOperator(ctx, ScanAggPlan(exprs=e, keys=long_keys, aggs=[(K.AGG_COUNT, None)]))
    with pytest.raises(ops.SrbError) as ei:
        op.push_chunk(K.Chunk(1, {i: K.Column.from_ints(K.TYPE_DECIMAL128, [1], None, 38, 0) for i in range(5)}))
    assert ei.value.code == K.NOT_IMPLEMENTED
    op.close()


# ---- joins ------------------------------------------------------------------------------------------------------------------
@pytest.fixture(params=["tiled", "general"])
def join_mode(request, monkeypatch):
    """Joins with at most one match per probe row (unique build keys, LEFT SEMI) take the tiled two-pass probe; SRB_JOIN_GENERAL
    forces the chain-walking count/append probe that serves every other join. Both must give the oracle's rows."""
    if request.param == "general":
        monkeypatch.setenv("SRB_JOIN_GENERAL", "1")
    else:
        monkeypatch.delenv("SRB_JOIN_GENERAL", raising=False)
    return request.param


def run_join(ctx, plan, build, probe, build_pieces=1, device=False):
    from starrocks_b200 import ops
    op = ops.HashJoinOperator(ctx, plan)
    ptrs = []
    try:
        assert not op.probe.need_input()  # hash_join_probe_operator.cpp:55-77: gated on the build
        step = max(1, build.n_rows // build_pieces)
        for a in range(0, build.n_rows, step):
            op.push_chunk(ops.slice_chunk(build, a, min(build.n_rows, a + step)))
        op.set_finishing()
        assert op.probe.need_input()
        if device:
            probe, ptrs = ctx.device_chunk(probe)
        op.probe.push_chunk(probe)
        op.probe.set_finishing()
        out = op.probe.pull_all()
        assert op.probe.is_finished()
        return out
    finally:
        for p in ptrs:
            ctx.device_free(p)
        op.close()


def run_join_chunks(ctx, plan, build, probes, device=False, build_pieces=2):
    """build in pieces, one probe chunk after the other (every chunk's output pulled before the next push, as the driver does), then
    probe set_finishing and the final pull: the build-side remainder of RIGHT / FULL OUTER and RIGHT SEMI / ANTI joins"""
    from starrocks_b200 import ops
    op = ops.HashJoinOperator(ctx, plan)
    ptrs, rows = [], []
    try:
        step = max(1, build.n_rows // build_pieces)
        for a in range(0, max(build.n_rows, 1), step):
            op.push_chunk(ops.slice_chunk(build, a, min(build.n_rows, a + step)))
        op.set_finishing()
        for pr in probes:
            if device:
                pr, pp = ctx.device_chunk(pr)
                ptrs += pp
            op.probe.push_chunk(pr)
            rows += K.chunk_rows(op.probe.pull_all(), sort=False)
        op.probe.set_finishing()
        rows += K.chunk_rows(op.probe.pull_all(), sort=False)
        assert op.probe.is_finished()
        return rows
    finally:
        for p in ptrs:
            ctx.device_free(p)
        op.close()


OUTER_TYPES = [K.JOIN_LEFT_OUTER, K.JOIN_RIGHT_OUTER, K.JOIN_FULL_OUTER, K.JOIN_RIGHT_SEMI, K.JOIN_RIGHT_ANTI]


@pytest.mark.parametrize("jt", OUTER_TYPES)
@pytest.mark.parametrize("ktypes", [[K.TYPE_INT], [K.TYPE_BIGINT, K.TYPE_SMALLINT], [K.TYPE_DECIMAL128]])
def test_outer_and_right_side_joins(ctx, oracle, jt, ktypes):
    # join_hash_map.tpp:1187-1986: LEFT / RIGHT / FULL OUTER, RIGHT SEMI / ANTI; duplicate keys, NULL keys on both sides, several probe chunks
    from starrocks_b200 import ops
    rng = np.random.default_rng(23 + jt)
    nb, npr = 40_000, 150_000

    def keycols(n, base_slot, span):
        cols = {}
        for i, t in enumerate(ktypes):
            v = rng.integers(0, span if i == 0 else 3, n)
            if K.type_width(t) == 16:
                v = [int(x) * 10**20 for x in v]
            cols[base_slot + i] = K.Column.from_ints(t, v, (rng.random(n) < 0.03).astype(np.uint8), 38 if K.type_width(t) == 16 else 0, 0)
        return cols

    bcols = keycols(nb, 0, 30_000)
    bcols[8] = K.Column.from_ints(K.TYPE_BIGINT, np.arange(nb))
    bcols[9] = K.Column.from_ints(K.TYPE_DECIMAL64, rng.integers(0, 999, nb), (rng.random(nb) < 0.1).astype(np.uint8), 15, 2)
    build = K.Chunk(nb, bcols)
    pcols = keycols(npr, 20, 60_000)
    pcols[28] = K.Column.from_ints(K.TYPE_INT, np.arange(npr, dtype=np.int32))
    probe = K.Chunk(npr, pcols)
    nk = len(ktypes)
    plan = JoinPlan(list(range(nk)), list(range(20, 20 + nk)), build_out=[8, 9], probe_out=[28, 20], join_type=jt)
    want = sorted(K.chunk_rows(oracle.join(plan, build, probe), sort=False), key=repr)
    assert len(want) > 1000
    pieces = [ops.slice_chunk(probe, 0, 50_000), ops.slice_chunk(probe, 50_000, 50_001), ops.slice_chunk(probe, 50_001, npr)]
    for device in (False, True):
        got = run_join_chunks(ctx, plan, build, pieces, device=device)
        assert sorted(got, key=repr) == want, (jt, device)


@pytest.mark.parametrize("unique_build", [False, True])
def test_inner_join_with_other_conjuncts(ctx, oracle, unique_build):
    # HashJoiner::_process_other_conjunct (hash_joiner.cpp:488-507): `ON b.k = p.k AND p.d > b.d AND p.x <> b.y` — the conjuncts compare columns of
    # the two sides on every joined row; NULLs fail them; duplicate and unique build keys (the latter would otherwise take the tiled probe)
    from starrocks_b200 import ops
    rng = np.random.default_rng(41)
    nb, npr = 30_000, 120_000
    bk = np.arange(nb, dtype=np.int32) if unique_build else rng.integers(0, 8000, nb).astype(np.int32)
    build = K.Chunk(nb, {0: K.Column.from_ints(K.TYPE_INT, bk, (rng.random(nb) < 0.02).astype(np.uint8)),
                         1: K.Column.from_ints(K.TYPE_DATE, rng.integers(9000, 9400, nb).astype(np.int32), (rng.random(nb) < 0.05).astype(np.uint8)),
                         2: K.Column.from_ints(K.TYPE_DECIMAL64, rng.integers(0, 50, nb), None, 15, 2),
                         3: K.Column.from_ints(K.TYPE_BIGINT, np.arange(nb))})
    probe = K.Chunk(npr, {10: K.Column.from_ints(K.TYPE_INT, rng.integers(0, nb if unique_build else 9000, npr).astype(np.int32)),
                          11: K.Column.from_ints(K.TYPE_DATE, rng.integers(9000, 9400, npr).astype(np.int32), (rng.random(npr) < 0.05).astype(np.uint8)),
                          12: K.Column.from_ints(K.TYPE_DECIMAL64, rng.integers(0, 50, npr), None, 15, 2),
                          13: K.Column.from_ints(K.TYPE_INT, np.arange(npr, dtype=np.int32))})
    for conj in ([(11, K.PRED_GT, 1)], [(11, K.PRED_GE, 1), (12, K.PRED_NE, 2)], [(2, K.PRED_LT, 12), (10, K.PRED_EQ, 0)]):
        plan = JoinPlan([0], [10], build_out=[1, 2, 3], probe_out=[11, 12, 13], conjuncts=conj)
        want = sorted(K.chunk_rows(oracle.join(plan, build, probe), sort=False), key=repr)
        plain = oracle.join(JoinPlan([0], [10], build_out=[1, 2, 3], probe_out=[11, 12, 13]), build, probe).n_rows
        assert 100 < len(want) < plain
        pieces = [ops.slice_chunk(probe, 0, 70_000), ops.slice_chunk(probe, 70_000, npr)]
        for device in (False, True):
            got = run_join_chunks(ctx, plan, build, pieces, device=device)
            assert sorted(got, key=repr) == want, (conj, device)
    # a conjunct over a column the join does not carry, or on an outer join, is refused
    with pytest.raises(ops.SrbError):
        run_join_chunks(ctx, JoinPlan([0], [10], build_out=[3], probe_out=[13], conjuncts=[(11, K.PRED_GT, 1)]), build, [probe])
    with pytest.raises(ops.SrbError):
        ops.HashJoinOperator(ctx, JoinPlan([0], [10], build_out=[1], probe_out=[11], join_type=K.JOIN_LEFT_OUTER, conjuncts=[(11, K.PRED_GT, 1)]))


@pytest.mark.parametrize("build_nulls", [False, True])
@pytest.mark.parametrize("ktypes", [[K.TYPE_INT], [K.TYPE_BIGINT, K.TYPE_SMALLINT], [K.TYPE_DECIMAL128]])
def test_null_aware_left_anti_join(ctx, oracle, build_nulls, ktypes):
    # NOT IN (join_hash_map.tpp:1356-1383, hash_joiner.h:386-396): NULL probe keys are dropped; one NULL key on the build side and nothing
    # qualifies; plain LEFT ANTI over the same rows keeps the NULL probe keys
    from starrocks_b200 import ops
    rng = np.random.default_rng(5)
    nb, npr = 30_000, 120_000

    def keycols(n, base_slot, span, p_null):
        cols = {}
        for i, t in enumerate(ktypes):
            v = rng.integers(0, span if i == 0 else 3, n)
            if K.type_width(t) == 16:
                v = [int(x) * 10**20 for x in v]
            cols[base_slot + i] = K.Column.from_ints(t, v, (rng.random(n) < p_null).astype(np.uint8), 38 if K.type_width(t) == 16 else 0, 0)
        return cols

    bcols = keycols(nb, 0, 30_000, 0.001 if build_nulls else 0.0)
    bcols[8] = K.Column.from_ints(K.TYPE_BIGINT, np.arange(nb))
    build = K.Chunk(nb, bcols)
    pcols = keycols(npr, 20, 60_000, 0.05)
    pcols[28] = K.Column.from_ints(K.TYPE_INT, np.arange(npr, dtype=np.int32))
    probe = K.Chunk(npr, pcols)
    nk = len(ktypes)
    pieces = [ops.slice_chunk(probe, 0, 70_000), ops.slice_chunk(probe, 70_000, npr)]
    counts = {}
    for jt in (K.JOIN_NULL_AWARE_LEFT_ANTI, K.JOIN_LEFT_ANTI):
        plan = JoinPlan(list(range(nk)), list(range(20, 20 + nk)), build_out=[], probe_out=[28, 20], join_type=jt)
        want = sorted(K.chunk_rows(oracle.join(plan, build, probe), sort=False), key=repr)
        for device in (False, True):
            got = run_join_chunks(ctx, plan, build, pieces, device=device)
            assert sorted(got, key=repr) == want, (jt, device)
        counts[jt] = len(want)
    assert counts[K.JOIN_LEFT_ANTI] > 1000
    assert counts[K.JOIN_NULL_AWARE_LEFT_ANTI] == 0 if build_nulls else 0 < counts[K.JOIN_NULL_AWARE_LEFT_ANTI] < counts[K.JOIN_LEFT_ANTI]


@pytest.mark.parametrize("jt,rows", [(K.JOIN_LEFT_OUTER, 5), (K.JOIN_FULL_OUTER, 5), (K.JOIN_LEFT_ANTI, 5), (K.JOIN_INNER, 0), (K.JOIN_LEFT_SEMI, 0),
                                     (K.JOIN_RIGHT_SEMI, 0), (K.JOIN_RIGHT_OUTER, 0), (K.JOIN_RIGHT_ANTI, 0), (K.JOIN_NULL_AWARE_LEFT_ANTI, 5)])
def test_joins_against_an_empty_build_side(ctx, oracle, jt, rows):
    # be/test/exec/join_hash_map_test.cpp:2376-2388 EmptyHashMapTest: 5 probe rows against an empty hash table
    build = K.Chunk(0, {0: K.Column.from_ints(K.TYPE_INT, np.zeros(0, np.int32)), 1: K.Column.from_ints(K.TYPE_BIGINT, np.zeros(0, np.int64))})
    probe = K.Chunk(5, {2: K.Column.from_ints(K.TYPE_INT, np.arange(5, dtype=np.int32)), 3: K.Column.from_ints(K.TYPE_BIGINT, np.arange(5) * 7)})
    plan = JoinPlan([0], [2], build_out=[1], probe_out=[2, 3], join_type=jt)
    want = K.chunk_rows(oracle.join(plan, build, probe), sort=False)
    assert len(want) == rows
    got = run_join_chunks(ctx, plan, build, [probe])
    assert sorted(got, key=repr) == sorted(want, key=repr)
    # and an empty probe side against a real build side: only the right-side remainder comes out
    build = K.Chunk(3, {0: K.Column.from_ints(K.TYPE_INT, np.array([4, 5, 5], np.int32)), 1: K.Column.from_ints(K.TYPE_BIGINT, np.array([1, 2, 3]))})
    none = K.Chunk(0, {2: K.Column.from_ints(K.TYPE_INT, np.zeros(0, np.int32)), 3: K.Column.from_ints(K.TYPE_BIGINT, np.zeros(0, np.int64))})
    want = K.chunk_rows(oracle.join(plan, build, none), sort=False)
    got = run_join_chunks(ctx, plan, build, [none])
    assert sorted(got, key=repr) == sorted(want, key=repr)


def test_join_reference_vectors(ctx, oracle, join_mode):
    # be/test/exec/join_hash_map_test.cpp semantics: 1-based build indexes, NULL keys never match, one-to-many chains
    build = K.Chunk(6, {0: K.Column.from_ints(K.TYPE_INT, [1, 2, 2, 3, 5, 7], nulls=[0, 0, 0, 0, 1, 0]),
                        1: K.Column.from_ints(K.TYPE_BIGINT, [10, 20, 21, 30, 50, 70])})
    probe = K.Chunk(5, {2: K.Column.from_ints(K.TYPE_INT, [2, 3, 4, 5, 7], nulls=[0, 0, 0, 0, 1]),
                        3: K.Column.from_ints(K.TYPE_BIGINT, [200, 300, 400, 500, 700])})
    plan = JoinPlan([0], [2], build_out=[1], probe_out=[3], emit_indexes=True)
    rows = K.chunk_rows(run_join(ctx, plan, build, probe))
    assert rows == [(200, 20, 0, 2), (200, 21, 0, 3), (300, 30, 1, 4)]
    assert rows == K.chunk_rows(oracle.join(plan, build, probe))
    semi = JoinPlan([0], [2], probe_out=[3], join_type=K.JOIN_LEFT_SEMI)
    assert K.chunk_rows(run_join(ctx, semi, build, probe)) == [(200,), (300,)]
    # LEFT ANTI: the probe rows without a match, a NULL probe key included (join_hash_map.tpp:1349-1420)
    anti = JoinPlan([0], [2], probe_out=[3], join_type=K.JOIN_LEFT_ANTI)
    assert K.chunk_rows(run_join(ctx, anti, build, probe)) == [(400,), (500,), (700,)] == K.chunk_rows(oracle.join(anti, build, probe))


@pytest.mark.parametrize("ktypes", [[K.TYPE_INT], [K.TYPE_BIGINT, K.TYPE_INT], [K.TYPE_DECIMAL128]])
def test_left_semi_and_anti_partition_the_probe_side(ctx, oracle, ktypes):
    """LEFT SEMI and LEFT ANTI against the oracle on duplicate-heavy build sides with NULL keys on both sides; together they return
    every probe row exactly once, in probe order; an empty build side makes ANTI return everything"""
    rng = np.random.default_rng(17 + len(ktypes))
    nb, npb = 20000, 150000
    bcols = {i: _rand_col(rng, t, nb, 0.05, small=True) for i, t in enumerate(ktypes)}
    pcols = {10 + i: _rand_col(rng, t, npb, 0.05, small=True) for i, t in enumerate(ktypes)}
    if ktypes == [K.TYPE_INT]:
        bcols[0] = K.Column.from_ints(K.TYPE_INT, rng.integers(0, 50000, nb), nulls=(rng.random(nb) < 0.02).astype(np.uint8))
        pcols[10] = K.Column.from_ints(K.TYPE_INT, rng.integers(-100, 60000, npb), nulls=(rng.random(npb) < 0.02).astype(np.uint8))
    pcols[15] = K.Column.from_ints(K.TYPE_BIGINT, np.arange(npb))
    build, probe = K.Chunk(nb, bcols), K.Chunk(npb, pcols)
    keys = (list(range(len(ktypes))), [10 + i for i in range(len(ktypes))])
    got = {}
    for name, jt in (("semi", K.JOIN_LEFT_SEMI), ("anti", K.JOIN_LEFT_ANTI)):
        plan = JoinPlan(*keys, probe_out=[15], join_type=jt)
        want = K.chunk_rows(oracle.join(plan, build, probe), sort=False)
        for device in (False, True):
            assert K.chunk_rows(run_join(ctx, plan, build, probe, build_pieces=3, device=device), sort=False) == want
        got[name] = [r[0] for r in want]
    assert len(got["semi"]) > 1000 and len(got["anti"]) > 1000
    assert sorted(got["semi"] + got["anti"]) == list(range(npb))
    empty = K.Chunk(0, {i: K.Column.from_ints(t, []) for i, t in enumerate(ktypes)})
    plan = JoinPlan(*keys, probe_out=[15], join_type=K.JOIN_LEFT_ANTI)
    assert [r[0] for r in K.chunk_rows(run_join(ctx, plan, empty, probe), sort=False)] == list(range(npb))


@pytest.mark.parametrize("ktypes", [[K.TYPE_INT], [K.TYPE_BIGINT], [K.TYPE_INT, K.TYPE_INT], [K.TYPE_SMALLINT], [K.TYPE_BIGINT, K.TYPE_INT],
                                    [K.TYPE_DECIMAL128]])
def test_join_key_widths_against_oracle(ctx, oracle, ktypes, join_mode):
    rng = np.random.default_rng(len(ktypes) + ktypes[0])
    nb, npb = 30000, 100000
    bcols = {i: _rand_col(rng, t, nb, 0.05, small=True) for i, t in enumerate(ktypes)}
    pcols = {10 + i: _rand_col(rng, t, npb, 0.05, small=True) for i, t in enumerate(ktypes)}
    if len(ktypes) == 1 and ktypes[0] != K.TYPE_SMALLINT:  # wider domain: mostly unique build keys
        bcols[0] = K.Column.from_ints(ktypes[0], rng.permutation(nb * 3)[:nb] - nb)
        pcols[10] = K.Column.from_ints(ktypes[0], rng.integers(-nb, 2 * nb, npb))
    bcols[5] = _rand_col(rng, K.TYPE_DECIMAL64, nb, 0.1, p=15, s=2)
    bcols[6] = K.Column.from_ints(K.TYPE_INT, np.arange(nb))
    pcols[15] = K.Column.from_ints(K.TYPE_BIGINT, np.arange(npb))
    build, probe = K.Chunk(nb, bcols), K.Chunk(npb, pcols)
    plan = JoinPlan(list(range(len(ktypes))), [10 + i for i in range(len(ktypes))], build_out=[5, 6], probe_out=[15, 10])
    if ktypes == [K.TYPE_SMALLINT]:  # 100 distinct keys x 30000 x 100000 rows would explode: shrink
        build = K.Chunk(300, {s: K.Column(c.type, c.data[:300], None if c.nulls is None else c.nulls[:300], None, c.precision, c.scale) for s, c in bcols.items()})
        probe = K.Chunk(2000, {s: K.Column(c.type, c.data[:2000], None if c.nulls is None else c.nulls[:2000], None, c.precision, c.scale) for s, c in pcols.items()})
    want = K.chunk_rows(oracle.join(plan, build, probe))
    assert K.chunk_rows(run_join(ctx, plan, build, probe)) == want
    assert K.chunk_rows(run_join(ctx, plan, build, probe, build_pieces=7, device=True)) == want
    assert len(want) > 0


def test_join_empty_sides(ctx, oracle, join_mode):
    build = K.Chunk(3, {0: K.Column.from_ints(K.TYPE_INT, [1, 2, 3]), 1: K.Column.from_ints(K.TYPE_INT, [7, 8, 9])})
    probe = K.Chunk(2, {2: K.Column.from_ints(K.TYPE_INT, [10, 11])})
    plan = JoinPlan([0], [2], build_out=[1], probe_out=[2])
    assert run_join(ctx, plan, build, probe).n_rows == 0
    empty_build = K.Chunk(0, {0: K.Column.from_ints(K.TYPE_INT, []), 1: K.Column.from_ints(K.TYPE_INT, [])})
    assert run_join(ctx, plan, empty_build, probe).n_rows == 0


def test_q3_pipeline_chained_on_device(ctx, oracle, join_mode):
    """customer(filter) -> build; orders(filter) -> probe -> build; lineitem(filter) -> probe -> group by, all operator outputs
    handed over as device chunks; compared with the same chain through the oracle."""
    from starrocks_b200 import ops
    n_c, n_o, n_l = 15000, 150000, 600000
    cust = T.host_chunk("customer", 0, n_c, ["c_custkey", "c_mktsegment_code"])
    orders = T.host_chunk("orders", 0, n_o, ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"], n_customers=n_c)
    line = T.host_chunk("lineitem", 0, n_l, ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"])

    # oracle chain
    oc = oracle.scan_agg(Q.q3_customer_filter(), cust)
    oo = oracle.scan_agg(Q.q3_orders_filter(), orders)
    j1 = oracle.join(Q.q3_join_customer_orders(), oc, oo)
    ol = oracle.scan_agg(Q.q3_lineitem_filter(), line)
    j2 = oracle.join(Q.q3_join_orders_lineitem(), j1, ol)
    want = K.chunk_rows(oracle.scan_agg(Q.q3_agg_plan(), j2))
    assert len(want) > 100

    def filt(plan, ch):
        op = ops.ScanAggOperator(ctx, plan)
        op.push_chunk(ch)
        op.set_finishing()
        return op, op.pull_device()

    fc, dc = filt(Q.q3_customer_filter(), cust)
    jo1 = ops.HashJoinOperator(ctx, Q.q3_join_customer_orders())
    jo1.push_chunk(dc)
    jo1.set_finishing()
    fo, do = filt(Q.q3_orders_filter(), orders)
    jo1.probe.push_chunk(do)
    d1 = jo1.probe.pull_device()
    jo2 = ops.HashJoinOperator(ctx, Q.q3_join_orders_lineitem())
    jo2.push_chunk(d1)
    jo2.set_finishing()
    fl, dl = filt(Q.q3_lineitem_filter(), line)
    jo2.probe.push_chunk(dl)
    d2 = jo2.probe.pull_device()
    agg = ops.ScanAggOperator(ctx, Q.q3_agg_plan())
    agg.push_chunk(d2)
    agg.set_finishing()
    got = K.chunk_rows(agg.pull_all())
    for o in (fc, fo, fl, jo1, jo2, agg):
        o.close()
    assert got == want


def _ssb_tables(n_lo, n_c, n_s, n_p, row0=0):
    kw = dict(n_customers=n_c, n_suppliers=n_s, n_parts=n_p)
    from starrocks_b200 import pipelines as P
    dims = {"ssb_customer": T.host_chunk("ssb_customer", 0, n_c, P.SSB_DIM_COLS["ssb_customer"]),
            "ssb_supplier": T.host_chunk("ssb_supplier", 0, n_s, P.SSB_DIM_COLS["ssb_supplier"]),
            "ssb_part": T.host_chunk("ssb_part", 0, n_p, P.SSB_DIM_COLS["ssb_part"]),
            "ssb_dates": T.host_chunk("ssb_dates", 0, T.SSB_DAYS, P.SSB_DIM_COLS["ssb_dates"])}
    return dims, T.host_chunk("lineorder", row0, n_lo, T.LINEORDER_Q41, **kw)


@pytest.mark.parametrize("use_rf", [False, True])
def test_ssb_q41_star_join_chained_on_device(ctx, oracle, join_mode, use_rf):
    """SSB Q4.1: three filtered dimensions + dates as build sides, lineorder probing them in turn with late
    materialisation, GROUP BY d_year, c_nation — every operator output handed on as a device chunk; compared with the oracle chain."""
    from starrocks_b200 import pipelines as P
    dims, lo = _ssb_tables(600_000, 30_000, 2_000, 200_000)
    want = K.chunk_rows(OC.ssb_q41_oracle(oracle, dims, lo))
    assert len(want) == 7 * 5  # 7 years x the 5 nations of AMERICA
    q = P.SsbQ41(ctx, runtime_filter=use_rf).build(dims)
    stage_rows = []
    agg = q.probe(lo, stage_rows=stage_rows)
    got = K.chunk_rows(agg.pull_all())
    agg.close()
    # a second fact-table pass through the same build sides (the next lineorder morsel range) gives the same answer
    agg2 = q.probe(lo)
    got2 = K.chunk_rows(agg2.pull_all())
    agg2.close()
    q.close()
    assert got == want and got2 == want
    if use_rf:  # the scan's three runtime filters already drop nearly every row a probe would drop (false positives remain)
        assert lo.n_rows // 20 > stage_rows[0] >= stage_rows[1] >= stage_rows[2] >= stage_rows[3] == stage_rows[4] > 0
        assert stage_rows[0] < stage_rows[3] * 1.2
        stage_rows = stage_rows[1:]
    else:
        assert stage_rows[0] > stage_rows[1] > stage_rows[2] == stage_rows[3] > 0  # 1/5 supplier, 1/5 customer, 2/5 part, every date matches
    oracle_chain = [oracle.join(Q.ssb_join_supplier(), oracle.scan_agg(Q.ssb_supplier_filter(), dims["ssb_supplier"]), lo).n_rows]
    assert stage_rows[0] == oracle_chain[0] or use_rf  # without filters the first probe's output is the oracle's join output


def test_device_pull_of_aggregate_result(ctx, oracle):
    from starrocks_b200 import ops
    ch = T.host_chunk("highcard", 0, 200000, ["k", "v"], n_keys=5000)
    op = ops.ScanAggOperator(ctx, Q.highcard_plan())
    op.push_chunk(ch)
    op.set_finishing()
    d = op.pull_device()
    got = ctx.fetch_chunk(d)
    op.close()
    assert K.chunk_rows(got) == K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), ch))


# ---- runtime filters: build-side bloom filter + min/max, applied by the probe-side scan ---------------------------
def _build_join(ctx, type_, keys, nulls=None):
    from starrocks_b200 import ops
    build = K.Chunk(len(keys), {0: K.Column.from_ints(type_, keys, nulls=nulls)})
    j = ops.HashJoinOperator(ctx, JoinPlan([0], [1], build_out=[], probe_out=[1]))
    j.push_chunk(build)
    j.set_finishing()
    return j, build


@pytest.mark.parametrize("type_,np_t", [(K.TYPE_INT, np.int32), (K.TYPE_BIGINT, np.int64), (K.TYPE_DATE, np.int32), (K.TYPE_SMALLINT, np.int16)])
def test_runtime_filter_directory_is_the_reference_filter(ctx, oracle, type_, np_t):
    """the device-built SimdBlockFilter holds exactly the bits of the oracle's (= the reference's) filter of the same size, and
    min / max / has_null / inserted agree — for several key types, negative keys, NULL keys and an oversized expected_rows"""
    from tests.oracle_lib import OracleRuntimeFilter
    rng = np.random.default_rng(11)
    n = 40_000
    lo, hi = (-(1 << 14), 1 << 14) if np_t == np.int16 else (-(1 << 30), 1 << 30)
    keys = rng.integers(lo, hi, n).astype(np_t)
    nulls = (rng.random(n) < 0.01).astype(np.uint8)
    for use_nulls, expected, mix in ((False, 0, False), (True, 0, False), (True, 500_000, False), (True, 0, True)):
        j, build = _build_join(ctx, type_, keys, nulls if use_nulls else None)
        rf = j.build_runtime_filter(0, expected, mix=mix)
        want = OracleRuntimeFilter(oracle, expected or n, mix=mix).insert_column(build.columns[0])
        try:
            assert rf.info() == want.info()
            assert (rf.directory() == want.directory()).all()
        finally:
            want.close()
            rf.close()
            j.close()


def test_runtime_filter_of_an_empty_build_side_rejects_everything(ctx, oracle):
    from starrocks_b200 import ops
    j, _ = _build_join(ctx, K.TYPE_INT, np.array([], dtype=np.int32))
    rf = j.build_runtime_filter()
    assert rf.info()["inserted"] == 0 and rf.info()["min"] > rf.info()["max"]
    probe = K.Chunk(1000, {1: K.Column.from_ints(K.TYPE_INT, np.arange(1000, dtype=np.int32))})
    op = ops.ScanAggOperator(ctx, Q.filter_plan([], [1]))
    op.add_runtime_filter(rf, 1)
    op.push_chunk(probe)
    op.set_finishing()
    assert op.pull_all().n_rows == 0
    op.close()
    rf.close()
    j.close()


@pytest.mark.parametrize("pred_kind", ["range", "generic", "none"])
@pytest.mark.parametrize("nullable,mix", [(False, False), (False, True), (True, False), (True, True)])
def test_scan_filter_with_runtime_filter_matches_the_oracle(ctx, oracle, pred_kind, nullable, mix):
    """scan predicates AND the runtime filter on the key column, order-preserving compaction: the kept rows are exactly those the
    oracle's predicate + RuntimeBloomFilter::evaluate keep (false positives included), on the vectorised and the generic mask path,
    with NULL keys following has_null"""
    from starrocks_b200 import ops
    from tests.oracle_lib import OracleRuntimeFilter
    rng = np.random.default_rng(5)
    n_build, n = 20_000, 300_000
    bkeys = rng.integers(0, 1 << 20, n_build).astype(np.int32)
    for build_has_null in ((False, True) if nullable else (False,)):
        bnulls = None
        if build_has_null:
            bnulls = np.zeros(n_build, np.uint8)
            bnulls[7] = 1
        j, build = _build_join(ctx, K.TYPE_INT, bkeys, bnulls)
        rf = j.build_runtime_filter(mix=mix)
        want_rf = OracleRuntimeFilter(oracle, n_build, mix=mix).insert_column(build.columns[0])
        pk = rng.integers(-1000, (1 << 20) + 1000, n).astype(np.int32)
        pnull = (rng.random(n) < 0.05).astype(np.uint8) if nullable else None
        probe = K.Chunk(n, {1: K.Column.from_ints(K.TYPE_INT, pk, nulls=pnull),
                            2: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(0, 1000, n).astype(np.int64)),
                            3: K.Column.from_ints(K.TYPE_TINYINT, rng.integers(0, 5, n).astype(np.int8))})
        preds = {"range": [pred(2, K.PRED_LT, K.TYPE_BIGINT, 700), pred(2, K.PRED_GE, K.TYPE_BIGINT, 100)],
                 "generic": [pred(3, K.PRED_NE, K.TYPE_TINYINT, 2)], "none": []}[pred_kind]
        plan = Q.filter_plan(preds, [1, 2])
        # oracle: predicate selection, then the filter's evaluate on the key column
        v2, v3 = probe.columns[2].data, probe.columns[3].data
        sel = {"range": (v2 < 700) & (v2 >= 100), "generic": v3 != 2, "none": np.ones(n, bool)}[pred_kind].astype(np.uint8)
        # the filter is evaluated on its own all-ones selection (RunningContext::selection) and merged with the predicates' afterwards
        keep = (want_rf.evaluate(probe.columns[1]) & sel).astype(bool)
        assert 0 < keep.sum() < sel.sum()
        want = [r for r, k in zip(K.chunk_rows(K.Chunk(n, {1: probe.columns[1], 2: probe.columns[2]}), sort=False), keep) if k]
        for device in (False, True):
            op = ops.ScanAggOperator(ctx, plan)
            op.add_runtime_filter(rf, 1)
            ptrs = []
            src = probe
            if device:
                src, ptrs = ctx.device_chunk(probe)
            op.push_chunk(src)
            op.set_finishing()
            got = K.chunk_rows(op.pull_all(), sort=False)
            op.close()
            for p in ptrs:
                ctx.device_free(p)
            assert got == want  # same rows, same order
        want_rf.close()
        rf.close()
        j.close()


@pytest.mark.parametrize("type_,np_t,lo,hi", [(K.TYPE_TINYINT, np.int8, -50, 60), (K.TYPE_SMALLINT, np.int16, -9000, 12000),
                                              (K.TYPE_BIGINT, np.int64, -(1 << 40), 1 << 41)])
def test_runtime_filter_on_full_tiles_of_every_key_width(ctx, oracle, type_, np_t, lo, hi):
    """a scan without predicates: every tile reaches the filter fully selected (the full-tile body of k_filter_rf), for 1-, 2- and
    8-byte keys, probe keys on both sides of [min, max]; two and a half tiles so the last one takes the list path"""
    from starrocks_b200 import ops
    from tests.oracle_lib import OracleRuntimeFilter
    rng = np.random.default_rng(11)
    n_build, n = 40, 8192 * 2 + 4000
    bkeys = rng.integers(lo, hi, n_build).astype(np_t)
    j, build = _build_join(ctx, type_, bkeys, None)
    rf = j.build_runtime_filter()
    want_rf = OracleRuntimeFilter(oracle, n_build).insert_column(build.columns[0])
    info = np.iinfo(np_t)
    span = hi - lo
    pk = rng.integers(max(info.min, lo - span), min(info.max, hi + span), n).astype(np_t)
    pk[::3] = bkeys[rng.integers(0, n_build, len(pk[::3]))]  # a third of the probe keys are build keys
    probe = K.Chunk(n, {1: K.Column.from_ints(type_, pk), 2: K.Column.from_ints(K.TYPE_INT, np.arange(n, dtype=np.int32))})
    keep = want_rf.evaluate(probe.columns[1]).astype(bool)
    assert n // 3 <= keep.sum() < n
    want = [r for r, k in zip(K.chunk_rows(probe, sort=False), keep) if k]
    op = ops.ScanAggOperator(ctx, Q.filter_plan([], [1, 2]))
    op.add_runtime_filter(rf, 1)
    op.push_chunk(probe)
    op.set_finishing()
    got = K.chunk_rows(op.pull_all(), sort=False)
    op.close()
    assert got == want
    want_rf.close()
    rf.close()
    j.close()


@pytest.mark.parametrize("type_,np_t,lo,hi,n_build", [(K.TYPE_INT, np.int32, -3000, 40_000, 9000), (K.TYPE_BIGINT, np.int64, 1 << 40, (1 << 40) + 100_000, 30_000),
                                                      (K.TYPE_SMALLINT, np.int16, -20_000, 20_000, 6000), (K.TYPE_TINYINT, np.int8, -50, 60, 40),
                                                      (K.TYPE_DATE, np.int32, 8000, 10_600, 2000)])
@pytest.mark.parametrize("nullable", [False, True])
def test_bitset_runtime_filter_matches_the_oracle(ctx, oracle, type_, np_t, lo, hi, n_build, nullable):
    """a builder that holds the whole build side may take a RuntimeBitsetFilter (create_join_runtime_filter's rule): the device bitset holds
    the oracle's bits, and a scan with it keeps exactly the rows RuntimeBitsetFilter::evaluate keeps (exact membership inside [min, max],
    NULL keys following has_null), on full and on partly selected tiles"""
    from starrocks_b200 import ops
    from tests.oracle_lib import OracleBitsetFilter
    rng = np.random.default_rng(17)
    bkeys = rng.integers(lo, hi, n_build).astype(np_t)
    bnulls = None
    if nullable:
        bnulls = np.zeros(n_build, np.uint8)
        bnulls[3] = 1
    j, build = _build_join(ctx, type_, bkeys, bnulls)
    rf = j.build_runtime_filter(bitset=True)
    info = rf.info()
    live = bkeys[bnulls == 0] if nullable else bkeys
    assert rf.kind() == K.RF_KIND_BITSET and info["min"] == int(live.min()) and info["max"] == int(live.max())
    assert OracleBitsetFilter.chosen(oracle, info["min"], info["max"], n_build, 126 << 20)  # the cache of the rule: the GPU's L2
    want_rf = OracleBitsetFilter(oracle, info["min"], info["max"]).insert_column(build.columns[0])
    bits = rf.bitset().view(np.uint8)
    assert bits[:len(want_rf.data())].tolist() == want_rf.data().tolist() and not bits[len(want_rf.data()):].any()
    n = 8192 * 3 + 1234
    tinfo = np.iinfo(np_t)
    span = hi - lo
    pk = rng.integers(max(tinfo.min, lo - span // 2), min(tinfo.max, hi + span // 2), n).astype(np_t)
    pnull = (rng.random(n) < 0.04).astype(np.uint8) if nullable else None
    probe = K.Chunk(n, {1: K.Column.from_ints(type_, pk, nulls=pnull), 2: K.Column.from_ints(K.TYPE_INT, np.arange(n, dtype=np.int32))})
    for preds, sel in (([], np.ones(n, bool)), ([pred(2, K.PRED_GE, K.TYPE_INT, 5000)], np.arange(n) >= 5000)):
        keep = (want_rf.evaluate(probe.columns[1]).astype(bool)) & sel
        assert 0 < keep.sum() < sel.sum()
        want = [r for r, k in zip(K.chunk_rows(probe, sort=False), keep) if k]
        for device in (False, True):
            op = ops.ScanAggOperator(ctx, Q.filter_plan(preds, [1, 2]))
            op.add_runtime_filter(rf, 1)
            src, ptrs = (ctx.device_chunk(probe) if device else (probe, []))
            op.push_chunk(src)
            op.set_finishing()
            got = K.chunk_rows(op.pull_all(), sort=False)
            op.close()
            for p in ptrs:
                ctx.device_free(p)
            assert got == want
    want_rf.close()
    rf.close()
    j.close()


def test_bitset_runtime_filter_is_only_taken_when_the_rule_says_so(ctx, oracle):
    # be/test/exprs/runtime_filter_test.cpp:1409-1455 through the device; sparse keys and partial filters stay bloom filters
    from starrocks_b200 import ops
    j, build = _build_join(ctx, K.TYPE_INT, np.arange(0, 200, 2, dtype=np.int32))
    rf = j.build_runtime_filter(bitset=True)
    assert rf.kind() == K.RF_KIND_BITSET
    probe = K.Chunk(9, {1: K.Column.from_ints(K.TYPE_INT, np.array([-1, 5, 10, 15, 20, 25, 200, 0, 0], np.int32), np.array([0] * 7 + [1, 1], np.uint8)),
                        2: K.Column.from_ints(K.TYPE_INT, np.arange(9, dtype=np.int32))})
    op = ops.ScanAggOperator(ctx, Q.filter_plan([], [2]))
    op.add_runtime_filter(rf, 1)
    op.push_chunk(probe)
    op.set_finishing()
    assert [r[0] for r in K.chunk_rows(op.pull_all(), sort=False)] == [2, 4]
    op.close()
    rf.close()
    partial = j.build_runtime_filter(expected_rows=1000, bitset=True)  # a partial filter that will be merged: never a bitset
    assert partial.kind() == K.RF_KIND_BLOOM
    partial.close()
    j.close()
    rng = np.random.default_rng(2)
    j, _ = _build_join(ctx, K.TYPE_INT, rng.integers(0, 1 << 30, 5000).astype(np.int32))  # 128 MB of bitset for 5000 rows
    rf = j.build_runtime_filter(bitset=True)
    assert rf.kind() == K.RF_KIND_BLOOM
    rf.close()
    j.close()


def test_runtime_filter_misuse_is_reported(ctx):
    from starrocks_b200 import ops
    j, _ = _build_join(ctx, K.TYPE_INT, np.arange(10, dtype=np.int32))
    rf = j.build_runtime_filter()
    agg = ops.ScanAggOperator(ctx, Q.highcard_plan())
    with pytest.raises(ops.SrbError):
        agg.add_runtime_filter(rf, Q.HC_K)  # only scan-filter plans take runtime filters
    agg.close()
    f = ops.ScanAggOperator(ctx, Q.filter_plan([], [1]))
    f.add_runtime_filter(rf, 1)
    with pytest.raises(ops.SrbError):  # BIGINT column against a filter over INT keys
        f.push_chunk(K.Chunk(3, {1: K.Column.from_ints(K.TYPE_BIGINT, [1, 2, 3])}))
    f.close()
    j2 = ops.HashJoinOperator(ctx, JoinPlan([0], [1]))
    with pytest.raises(ops.SrbError):
        j2.build_runtime_filter()  # build side not finished
    j2.close()
    rf.close()
    j.close()


@pytest.mark.parametrize("use_rf", [True, False])
def test_q3_pipeline_with_and_without_runtime_filter(ctx, oracle, use_rf):
    from starrocks_b200 import pipelines as P
    n_c, n_o, n_l = 6000, 60000, 240000
    cust = T.host_chunk("customer", 0, n_c, ["c_custkey", "c_mktsegment_code"])
    orders = T.host_chunk("orders", 0, n_o, ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"], n_customers=n_c)
    line = T.host_chunk("lineitem", 0, n_l, ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"])
    want = K.chunk_rows(OC.q3_oracle(oracle, cust, orders, line))
    q = P.TpchQ3(ctx, runtime_filter=use_rf)
    rows = {}
    try:
        got = K.chunk_rows(q.run(cust, orders, line, rows=rows).pull_all())
    finally:
        q.close()
    assert got == want
    plain = int((line.columns[Q.L_SHIPDATE].data > Q.Q3_DATE).sum())
    assert (rows["lineitem_kept"] < plain // 2) if use_rf else (rows["lineitem_kept"] == plain)  # the filter drops most non-matching lines


@pytest.mark.parametrize("n", [3000, 400_000])
def test_q3_aggregation_static_shape(ctx, oracle, n):
    """Q3's GROUP BY over the join output (3 INT/DATE keys, sum(ext * (1 - disc))): the statically specialised kernel, the same with
    the general body, and the generic interpreter all give the oracle's rows — including negative keys and amounts, and a
    discount above one; and the plan really takes the static shape"""
    from starrocks_b200 import ops
    rng = np.random.default_rng(n)
    okey = np.sort(rng.integers(-50, n // 2, n)).astype(np.int32)
    ch = K.Chunk(n, {Q.L_ORDERKEY: K.Column.from_ints(K.TYPE_INT, okey),
                     Q.O_ORDERDATE: K.Column.from_ints(K.TYPE_DATE, (okey % 2000 + 2448000).astype(np.int32)),
                     Q.O_SHIPPRIORITY: K.Column.from_ints(K.TYPE_INT, (okey % 3 - 1).astype(np.int32)),
                     Q.L_EXTENDEDPRICE: K.Column.from_ints(K.TYPE_DECIMAL64, rng.integers(-10**7, 10**9, n).astype(np.int64), precision=15, scale=2),
                     Q.L_DISCOUNT: K.Column.from_ints(K.TYPE_DECIMAL64, rng.integers(0, 120, n).astype(np.int64), precision=15, scale=2)})
    _both_kernel_families(ctx, oracle, Q.q3_agg_plan, ch)
    dch, ptrs = ctx.device_chunk(ch)
    op = ops.ScanAggOperator(ctx, Q.q3_agg_plan(group_hint=n))
    op.push_chunk(dch)
    op.set_finishing()
    op.pull_device()
    prof = op.profile()
    op.close()
    for p in ptrs:
        ctx.device_free(p)
    assert prof["shape"] == 3 and (prof["static_launches"] >= 1 or n < 4096)


# ---- streaming pre-aggregation sink: tests/test_streaming_cpu.py pins the oracle side -----------------------------
def _stream_gpu(ctx, plan, chunk, mode, piece_rows=0):
    from starrocks_b200 import ops
    op = ops.ScanAggOperator(ctx, plan)
    op.set_streaming_mode(mode)
    if piece_rows:
        for a in range(0, chunk.n_rows, piece_rows):
            op.push_chunk(ops.slice_chunk(chunk, a, min(chunk.n_rows, a + piece_rows)))
    else:
        op.push_chunk(chunk)
    op.set_finishing()
    out = op.pull_all()
    st = op.streaming_stats()
    op.close()
    return out, st


def test_streaming_mode_misuse_is_reported(ctx):
    from starrocks_b200 import ops
    op = ops.ScanAggOperator(ctx, Q.highcard_plan())  # update + finalize: not the first phase of a two-phase aggregation
    with pytest.raises(ops.SrbError) as ei:
        op.set_streaming_mode(K.STREAMING_AUTO)
    assert ei.value.code == K.INVALID_ARGUMENT
    op.close()
    op = ops.ScanAggOperator(ctx, Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE))
    with pytest.raises(ops.SrbError):
        op.set_streaming_mode(7)
    op.close()


def check_streaming_mode(ctx, oracle, mode, monkeypatch):
    """heavy repetition, then all-distinct keys, then moderate repetition, pushed in 4096-row chunks with a 4096-row AUTO unit: the
    AUTO machine leaves INIT_PREAGG, passes rows through and comes back; whatever it does, phase 2 over the sink's output equals
    the one-phase result (oracle), and FORCE_PREAGGREGATION equals the blocking first phase"""
    from tests.test_streaming_cpu import hc_chunk
    monkeypatch.setenv("SRB_STREAM_UNIT", "4096")
    monkeypatch.setenv("SRB_STREAM_L2_BYTES", "65536")  # the table outgrows "the L2" after a few thousand groups
    rng = np.random.default_rng(22)
    keys = np.concatenate([rng.integers(0, 100, 30_000), 10**6 + np.arange(300_000), rng.integers(0, 20_000, 50_000)])
    ch = hc_chunk(keys, rng.integers(-10**12, 10**12, len(keys)), (rng.random(len(keys)) < 0.05).astype(np.uint8))
    inter, st = _stream_gpu(ctx, Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE), ch, mode, piece_rows=65536)
    assert st["rows_streamed"] + st["groups"] == inter.n_rows
    if mode == K.STREAMING_AUTO:
        assert 0 < st["rows_streamed"] < len(keys)
    elif mode == K.STREAMING_FORCE_STREAMING:
        assert st["rows_streamed"] == len(keys) and st["groups"] == 0
    elif mode == K.STREAMING_FORCE_PREAGGREGATION:
        assert st["rows_streamed"] == 0
        assert K.chunk_rows(inter) == K.chunk_rows(oracle.scan_agg(Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE), ch))
    got = K.chunk_rows(run_gpu(ctx, Q.highcard_merge_plan(), inter))
    assert got == K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), ch))


def test_streaming_sink_force_preaggregation_is_the_blocking_first_phase(ctx, oracle, monkeypatch):
    # the modes that pass rows through run in tests/test_gpu_streaming.py
    check_streaming_mode(ctx, oracle, K.STREAMING_FORCE_PREAGGREGATION, monkeypatch)
