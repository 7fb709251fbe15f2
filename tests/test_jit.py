"""Run-time compiled kernel shapes (csrc/srb_jit.h): a plan that no built-in shape matches gets the static TMA-staged kernel compiled for it.

CPU part (no GPU needed: NVRTC cross-compiles): the generator + compiler accept the plan families below and reject what the static kernel
does not cover. GPU part: the compiled kernel's results through the C ABI against the oracle, bit for bit, with the operator's profile
confirming that the compiled kernel (not the interpreter) did the work."""
import os

import numpy as np
import pytest

from starrocks_b200 import capi as K
from starrocks_b200 import queries as Q
from starrocks_b200 import tpch_synth as T
from starrocks_b200.plan import ExprBuilder, ScanAggPlan, pred

D64, D128 = K.TYPE_DECIMAL64, K.TYPE_DECIMAL128
SHAPE_JIT = 100


def lineitem_plan(n_keys=2, hint=4, extra=True, phase=K.AGG_UPDATE_FINALIZE, charge=False, preds=True):
    """a Q1-like plan that is NOT one of the built-in shapes: other aggregates, other key lists, other cardinality estimates"""
    e = ExprBuilder()
    qty = e.slot(Q.L_QUANTITY, D64, 15, 2)
    ext = e.slot(Q.L_EXTENDEDPRICE, D64, 15, 2)
    disc = e.slot(Q.L_DISCOUNT, D64, 15, 2)
    tax = e.slot(Q.L_TAX, D64, 15, 2)
    keys = [e.slot(Q.L_RETURNFLAG_CODE, K.TYPE_INT), e.slot(Q.L_LINESTATUS_CODE, K.TYPE_INT)][:n_keys]
    one = e.lit(100, D64, 16, 2)
    omd = e.sub(one, disc, D64, 16)
    dp = e.mul(e.cast(ext, D128, 15, 2), e.cast(omd, D128, 16, 2), D128, 31)
    aggs = [(K.AGG_SUM, qty), (K.AGG_SUM, dp), (K.AGG_AVG, disc), (K.AGG_COUNT, None)]
    if charge:
        opt = e.add(one, tax, D64, 16)
        aggs.append((K.AGG_SUM, e.mul(dp, e.cast(opt, D128, 16, 2), D128, 38)))
    if extra:
        aggs += [(K.AGG_SUM, tax), (K.AGG_COUNT, tax)]
    ps = [pred(Q.L_SHIPDATE, K.PRED_LE, K.TYPE_DATE, Q.from_date(1998, 9, 2))] if preds else []
    return ScanAggPlan(preds=ps, exprs=e, keys=keys, aggs=aggs, phase=phase, group_hint=hint)


def orderkey_plan(hint):
    """GROUP BY l_orderkey (INT): SUM(l_extendedprice), COUNT(*), AVG(l_discount) — many groups"""
    e = ExprBuilder()
    k = e.slot(Q.L_ORDERKEY, K.TYPE_INT)
    ext = e.slot(Q.L_EXTENDEDPRICE, D64, 15, 2)
    disc = e.slot(Q.L_DISCOUNT, D64, 15, 2)
    return ScanAggPlan(preds=[], exprs=e, keys=[k], aggs=[(K.AGG_SUM, ext), (K.AGG_COUNT, None), (K.AGG_AVG, disc)], group_hint=hint)


def minmax_plan(n_keys):
    e = ExprBuilder()
    keys = [e.slot(Q.L_ORDERKEY, K.TYPE_INT)] if n_keys else []
    ext = e.slot(Q.L_EXTENDEDPRICE, D64, 15, 2)
    disc = e.slot(Q.L_DISCOUNT, D64, 15, 2)
    qty = e.slot(Q.L_QUANTITY, D64, 15, 2)
    neg = e.sub(disc, ext, D64, 16)  # negative values through the order-preserving encoding
    return ScanAggPlan(preds=[], exprs=e, keys=keys, group_hint=100_000 if n_keys else 0,
                       aggs=[(K.AGG_MIN, ext), (K.AGG_MAX, neg), (K.AGG_MIN, neg), (K.AGG_SUM, qty), (K.AGG_COUNT, None), (K.AGG_MAX, qty)])


PLANS = {
    "two keys, six aggregates": (lambda: lineitem_plan(), T.LINEITEM_Q1),
    "no group by": (lambda: lineitem_plan(n_keys=0), T.LINEITEM_Q1),
    "one key": (lambda: lineitem_plan(n_keys=1, extra=False), T.LINEITEM_Q1),
    "directory of 16": (lambda: lineitem_plan(hint=12), T.LINEITEM_Q1),
    "serialize phase": (lambda: lineitem_plan(phase=K.AGG_UPDATE_SERIALIZE), T.LINEITEM_Q1),
    "checked product": (lambda: lineitem_plan(charge=True, extra=False), T.LINEITEM_Q1),
    "no predicate": (lambda: lineitem_plan(preds=False), T.LINEITEM_Q1),
    "min and max": (lambda: Q.q1_plan(with_min_max=True), T.LINEITEM_Q1),
    "min and max, no group by": (lambda: minmax_plan(0), T.LINEITEM_Q1),
    "min and max, many groups": (lambda: minmax_plan(1), ["l_orderkey", "l_extendedprice", "l_discount", "l_quantity"]),
    "many groups": (lambda: orderkey_plan(100_000), ["l_orderkey", "l_extendedprice", "l_discount"]),
    "unknown cardinality": (lambda: orderkey_plan(0), ["l_orderkey", "l_extendedprice", "l_discount"]),
}


# ---- CPU: generator + NVRTC ------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def no_disk_cache(tmp_path_factory):
    old = os.environ.get("SRB_JIT_CACHE_DIR")
    os.environ["SRB_JIT_CACHE_DIR"] = str(tmp_path_factory.mktemp("srb_jit"))
    yield os.environ["SRB_JIT_CACHE_DIR"]
    if old is None:
        os.environ.pop("SRB_JIT_CACHE_DIR")
    else:
        os.environ["SRB_JIT_CACHE_DIR"] = old


@pytest.mark.parametrize("name", ["two keys, six aggregates", "no group by", "many groups"])
def test_generated_shapes_compile_without_a_gpu(no_disk_cache, name):
    from starrocks_b200 import ops
    plan_fn, cols = PLANS[name]
    first = T.host_chunk("lineitem", 0, 16, cols)
    d = ops.jit_check_plan(plan_fn(), first)
    assert d.startswith("compiled shape"), d
    st = ops.jit_stats()
    assert st["failed"] == 0 and st["pending"] == 0, st


def test_every_shape_family_of_the_gpu_tests_compiles_without_a_gpu(no_disk_cache):
    # what the GPU parity cases below run is generated and compiled here as well: null-map variants, MIN / MAX, VARCHAR keys (offsets +
    # byte window), DOUBLE aggregates, table-path and directory geometries
    from starrocks_b200 import ops
    first = T.host_chunk("lineitem", 0, 16, T.LINEITEM_Q1)
    nul = K.Chunk(16, dict(first.columns))
    for slot in (Q.L_QUANTITY, Q.L_DISCOUNT, Q.L_RETURNFLAG_CODE):
        c = first.columns[slot]
        nul.columns[slot] = K.Column(c.type, c.data, np.zeros(16, np.uint8), None, c.precision, c.scale)
    cases = [(Q.q1_plan(with_min_max=True), first), (Q.q1_plan(with_min_max=True), nul), (lineitem_plan(hint=12), nul), (lineitem_plan(n_keys=0), nul),
             (string_key_plan(4), string_key_chunk(16, lambda r, i: b"ab")), (string_key_plan(50_000, True), string_key_chunk(16, lambda r, i: b"k", 0.3, second_int_key=True)),
             (double_plan(0, 4), double_chunk(16)), (double_plan(4, 100_000), double_chunk(16)), (double_plan(None, 0), double_chunk(16))]
    for plan, chunk in cases:
        assert ops.jit_check_plan(plan, chunk).startswith("compiled shape")
    st = ops.jit_stats()
    assert st["failed"] == 0 and st["pending"] == 0, st


def test_builtin_shapes_are_recognised_and_uncovered_plans_are_refused(no_disk_cache):
    from starrocks_b200 import ops
    first = T.host_chunk("lineitem", 0, 16, T.LINEITEM_Q1)
    assert ops.jit_check_plan(Q.q1_plan(), first).startswith("built-in shape")
    assert ops.jit_check_plan(Q.q1_plan(with_min_max=True), first).startswith("compiled shape")  # MIN / MAX of up to 8 bytes: covered
    e = ExprBuilder()
    ext = e.slot(Q.L_EXTENDEDPRICE, D64, 15, 2)
    wide = ScanAggPlan(preds=[], exprs=e, keys=[], aggs=[(K.AGG_MAX, e.cast(ext, D128, 15, 2))])
    with pytest.raises(ops.SrbError) as ex:  # MAX over a 16-byte value: the interpreter's
        ops.jit_check_plan(wide, first)
    assert ex.value.code == K.NOT_IMPLEMENTED
    # columns that arrive with null maps get a variant of their own
    nul = K.Chunk(16, dict(first.columns))
    c = first.columns[Q.L_DISCOUNT]
    nul.columns[Q.L_DISCOUNT] = K.Column(c.type, c.data, np.zeros(16, np.uint8), None, c.precision, c.scale)
    assert ops.jit_check_plan(lineitem_plan(), nul).startswith("compiled shape")


def test_compiled_shapes_are_cached_on_disk(no_disk_cache):
    from starrocks_b200 import ops
    first = T.host_chunk("lineitem", 0, 16, T.LINEITEM_Q1)
    ops.jit_check_plan(lineitem_plan(hint=7), first)
    files = [f for f in os.listdir(no_disk_cache) if f.endswith(".cubin")]
    assert files and all(os.path.getsize(os.path.join(no_disk_cache, f)) > 10_000 for f in files)


# ---- GPU: parity of the compiled kernels ---------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def ctx():
    from starrocks_b200 import ops
    c = ops.Context(device=0, chunk_size=4096)
    yield c
    c.close()


@pytest.fixture()
def jit_sync():
    keep = {k: os.environ.get(k) for k in ("SRB_JIT", "SRB_JIT_MIN_ROWS")}
    os.environ["SRB_JIT"] = "sync"
    os.environ["SRB_JIT_MIN_ROWS"] = "0"
    yield
    for k, v in keep.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v


def run_device(ctx, plan, chunk):
    """-> (result chunk, profile of the operator)"""
    from starrocks_b200 import ops
    dch, ptrs = ctx.device_chunk(chunk)
    op = ops.ScanAggOperator(ctx, plan)
    try:
        op.push_chunk(dch)
        op.set_finishing()
        return op.pull_all(), op.profile()
    finally:
        op.close()
        for p in ptrs:
            ctx.device_free(p)


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(PLANS))
def test_compiled_shape_matches_the_oracle(ctx, oracle, jit_sync, name):
    plan_fn, cols = PLANS[name]
    n = 700_001  # whole tiles through the compiled kernel, the ragged tail through the interpreter
    chunk = T.host_chunk("lineitem", 0, n, cols)
    want = K.chunk_rows(oracle.scan_agg(plan_fn(), chunk))
    got, prof = run_device(ctx, plan_fn(), chunk)
    assert prof["shape"] == SHAPE_JIT and prof["static_launches"] >= 1, prof
    assert K.chunk_rows(got) == want
    assert len(want) >= 1


@pytest.mark.gpu
def test_compiled_twin_of_a_builtin_shape_gives_the_same_rows(ctx, oracle, jit_sync):
    # SRB_NO_BUILTIN_SHAPES: Q1 itself goes through the generator -> the same kernel source as the built-in ShapeQ1Dict
    chunk = T.host_chunk("lineitem", 0, 500_000, T.LINEITEM_Q1)
    want = K.chunk_rows(oracle.scan_agg(Q.q1_plan(), chunk))
    os.environ["SRB_NO_BUILTIN_SHAPES"] = "1"
    try:
        got, prof = run_device(ctx, Q.q1_plan(), chunk)
    finally:
        os.environ.pop("SRB_NO_BUILTIN_SHAPES")
    assert prof["shape"] == SHAPE_JIT, prof
    assert K.chunk_rows(got) == want
    got, prof = run_device(ctx, Q.q1_plan(), chunk)
    assert prof["shape"] == 1 and K.chunk_rows(got) == want


@pytest.mark.gpu
def test_asynchronous_compilation_serves_the_first_operator_from_the_interpreter(ctx, oracle):
    from starrocks_b200 import ops
    keep = {k: os.environ.get(k) for k in ("SRB_JIT", "SRB_JIT_MIN_ROWS")}
    os.environ["SRB_JIT"] = "async"
    os.environ["SRB_JIT_MIN_ROWS"] = "0"
    try:
        plan_fn = lambda: lineitem_plan(hint=3, extra=False)  # a shape no other test compiles
        chunk = T.host_chunk("lineitem", 0, 300_000, T.LINEITEM_Q1)
        want = K.chunk_rows(oracle.scan_agg(plan_fn(), chunk))
        got, prof = run_device(ctx, plan_fn(), chunk)  # compilation starts; this operator is (almost certainly) served by the interpreter
        assert K.chunk_rows(got) == want
        ops.jit_wait()
        st = ops.jit_stats()
        assert st["pending"] == 0 and st["failed"] == 0, st
        got, prof = run_device(ctx, plan_fn(), chunk)
        assert prof["shape"] == SHAPE_JIT and K.chunk_rows(got) == want
    finally:
        for k, v in keep.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


@pytest.mark.gpu
def test_compiled_shape_with_null_maps_and_many_chunks(ctx, oracle, jit_sync):
    # null maps without a set byte are stripped per batch (the compiled kernel takes it); a batch whose maps hold NULLs goes to the interpreter;
    # both update the same table
    from starrocks_b200 import ops
    n = 400_000
    chunk = T.host_chunk("lineitem", 0, n, T.LINEITEM_Q1)
    clean = K.Chunk(n, dict(chunk.columns))
    c = chunk.columns[Q.L_TAX]
    clean.columns[Q.L_TAX] = K.Column(c.type, c.data, np.zeros(n, np.uint8), None, c.precision, c.scale)
    dirty = K.Chunk(n, dict(chunk.columns))
    nm = (np.arange(n) % 97 == 0).astype(np.uint8)
    dirty.columns[Q.L_TAX] = K.Column(c.type, c.data, nm, None, c.precision, c.scale)
    plan_fn = lambda: lineitem_plan()
    from starrocks_b200 import distributed as D
    both = D.concat_chunks([clean, dirty])
    want = K.chunk_rows(oracle.scan_agg(plan_fn(), both))
    d1, p1 = ctx.device_chunk(clean)
    d2, p2 = ctx.device_chunk(dirty)
    op = ops.ScanAggOperator(ctx, plan_fn())
    try:
        op.push_chunk(d1)
        op.push_chunk(d2)
        op.set_finishing()
        got = op.pull_all()
        prof = op.profile()
    finally:
        op.close()
        for p in p1 + p2:
            ctx.device_free(p)
    assert prof["shape"] == SHAPE_JIT and K.chunk_rows(got) == want


def with_nulls(chunk, slots, frac, seed=3):
    rng = np.random.default_rng(seed)
    out = K.Chunk(chunk.n_rows, dict(chunk.columns))
    for slot in slots:
        c = chunk.columns[slot]
        out.columns[slot] = K.Column(c.type, c.data, (rng.random(chunk.n_rows) < frac).astype(np.uint8), None, c.precision, c.scale)
    return out


@pytest.mark.gpu
@pytest.mark.parametrize("name,slots", [
    ("two keys, six aggregates", [Q.L_QUANTITY, Q.L_DISCOUNT, Q.L_TAX]),       # nullable measures: SUM / AVG / COUNT(x) skip NULLs
    ("two keys, six aggregates", [Q.L_RETURNFLAG_CODE, Q.L_TAX]),               # a nullable key: NULL is a group of its own
    ("two keys, six aggregates", [Q.L_SHIPDATE, Q.L_EXTENDEDPRICE]),            # a nullable predicate column: NULL never passes a compare
    ("no group by", [Q.L_QUANTITY, Q.L_EXTENDEDPRICE, Q.L_DISCOUNT, Q.L_TAX]),
    ("min and max", [Q.L_EXTENDEDPRICE, Q.L_LINESTATUS_CODE]),
    ("min and max, many groups", [Q.L_EXTENDEDPRICE, Q.L_DISCOUNT, Q.L_ORDERKEY]),
    ("checked product", [Q.L_TAX, Q.L_DISCOUNT]),
])
@pytest.mark.parametrize("frac", [0.02, 0.6])
def test_compiled_shape_with_nulls_matches_the_oracle(ctx, oracle, jit_sync, name, slots, frac):
    plan_fn, cols = PLANS[name]
    chunk = with_nulls(T.host_chunk("lineitem", 0, 600_001, cols), slots, frac)
    want = K.chunk_rows(oracle.scan_agg(plan_fn(), chunk))
    got, prof = run_device(ctx, plan_fn(), chunk)
    assert prof["shape"] == SHAPE_JIT and prof["static_launches"] >= 1, prof
    assert K.chunk_rows(got) == want


# ---- BinaryColumn group keys on the compiled kernel: offsets + a window of the bytes staged per tile ------------------------------------
def string_key_chunk(n, maker, nulls=0.0, seed=11, second_int_key=False):
    rng = np.random.default_rng(seed)
    vals = [maker(rng, i) for i in range(n)]
    nl = (rng.random(n) < nulls).astype(np.uint8) if nulls else None
    cols = {0: K.Column.from_strings(vals, nulls=nl), 1: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(-1000, 1000, n)),
            2: K.Column.from_ints(D64, rng.integers(0, 10**6, n), None, 15, 2)}
    if second_int_key:
        cols[3] = K.Column.from_ints(K.TYPE_SMALLINT, rng.integers(0, 3, n).astype(np.int16))
    return K.Chunk(n, cols)


def string_key_plan(hint, second_int_key=False):
    e = ExprBuilder()
    keys = [e.slot(0, K.TYPE_VARCHAR)]
    v = e.slot(1, K.TYPE_BIGINT)
    d = e.slot(2, D64, 15, 2)
    if second_int_key:
        keys.append(e.slot(3, K.TYPE_SMALLINT))
    return ScanAggPlan(preds=[pred(1, K.PRED_GE, K.TYPE_BIGINT, -900)], exprs=e, keys=keys, group_hint=hint,
                       aggs=[(K.AGG_SUM, v), (K.AGG_COUNT, None), (K.AGG_AVG, d), (K.AGG_MAX, v)])


STRING_CASES = {
    "one byte, four groups": (lambda r, i: bytes([65 + int(r.integers(0, 4))]), 4, 0.0, False),
    "0 to 6 bytes, nullable": (lambda r, i: b"abcdef"[:int(r.integers(0, 7))], 8, 0.05, False),
    "with a SMALLINT key": (lambda r, i: [b"x", b"yy", b"zzz"][int(r.integers(0, 3))], 12, 0.0, True),
    "11 bytes: the tile's strings overflow the window": (lambda r, i: b"0123456789" + bytes([65 + int(r.integers(0, 3))]), 4, 0.0, False),
    "many groups": (lambda r, i: b"k%05d" % int(r.integers(0, 20000)), 50000, 0.0, False),
}


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(STRING_CASES))
def test_compiled_shape_with_string_keys_matches_the_oracle(ctx, oracle, jit_sync, name):
    maker, hint, nulls, second = STRING_CASES[name]
    chunk = string_key_chunk(150_003, maker, nulls, second_int_key=second)
    plan_fn = lambda: string_key_plan(hint, second)
    want = K.chunk_rows(oracle.scan_agg(plan_fn(), chunk))
    got, prof = run_device(ctx, plan_fn(), chunk)
    assert prof["shape"] == SHAPE_JIT and prof["static_launches"] >= 1, prof
    assert K.chunk_rows(got) == want and len(want) >= 3


@pytest.mark.gpu
def test_q1_with_varchar_keys_on_the_compiled_kernel(ctx, oracle, jit_sync):
    chunk = T.add_binary_keys(T.host_chunk("lineitem", 0, 400_003, T.LINEITEM_Q1))
    for s_ in (Q.L_RETURNFLAG_CODE, Q.L_LINESTATUS_CODE):
        chunk.columns.pop(s_)
    plan_fn = lambda: Q.q1_plan(dict_keys=False)
    want = K.chunk_rows(oracle.scan_agg(plan_fn(), chunk))
    got, prof = run_device(ctx, plan_fn(), chunk)
    assert prof["shape"] == SHAPE_JIT and K.chunk_rows(got) == want and len(want) == 4


@pytest.mark.gpu
def test_string_key_beyond_16_bytes_is_refused_on_the_compiled_kernel_too(ctx, jit_sync):
    from starrocks_b200 import ops
    chunk = string_key_chunk(100_000, lambda r, i: b"short" if i != 77_777 else b"a key of twenty bytes")
    with pytest.raises(ops.SrbError) as ex:
        run_device(ctx, string_key_plan(4), chunk)
    assert ex.value.code == K.NOT_IMPLEMENTED


# ---- a compilation that fails leaves the operator on the interpreter ---------------------------------------------------------------------
def test_a_failed_compilation_is_reported_by_check_plan(no_disk_cache, monkeypatch):
    from starrocks_b200 import ops
    monkeypatch.setenv("SRB_JIT_BREAK", "1")
    first = T.host_chunk("lineitem", 0, 16, T.LINEITEM_Q1)
    with pytest.raises(ops.SrbError) as ex:
        ops.jit_check_plan(lineitem_plan(hint=5), first)
    assert ex.value.code == K.INTERNAL_ERROR and "SRB_JIT_BREAK" in str(ex.value)
    assert ops.jit_stats()["failed"] >= 1


@pytest.mark.gpu
def test_a_failed_compilation_falls_back_to_the_interpreter(ctx, oracle, jit_sync, monkeypatch):
    monkeypatch.setenv("SRB_JIT_BREAK", "1")
    plan_fn = lambda: lineitem_plan(hint=6, charge=True)  # a shape of its own: nothing cached
    chunk = T.host_chunk("lineitem", 0, 300_000, T.LINEITEM_Q1)
    want = K.chunk_rows(oracle.scan_agg(plan_fn(), chunk))
    got, prof = run_device(ctx, plan_fn(), chunk)
    assert prof["shape"] != SHAPE_JIT and prof["static_launches"] == 0
    assert K.chunk_rows(got) == want


# ---- DOUBLE aggregates on the compiled kernel: sums kept as doubles per thread, MIN / MAX through the order-preserving encoding ----------
def double_chunk(n, seed=9):
    rng = np.random.default_rng(seed)
    return K.Chunk(n, {0: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 3, n).astype(np.int32)),
                       1: K.Column.from_ints(K.TYPE_DOUBLE, rng.normal(0, 1000, n)),
                       2: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(-100, 100, n)),
                       3: K.Column.from_ints(K.TYPE_DOUBLE, rng.random(n) * 50, (rng.random(n) < 0.2).astype(np.uint8)),
                       4: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 50_000, n).astype(np.int32))})


def double_plan(key_slot, hint):
    e = ExprBuilder()
    keys = [e.slot(key_slot, K.TYPE_INT)] if key_slot is not None else []
    d, b, dn = e.slot(1, K.TYPE_DOUBLE), e.slot(2, K.TYPE_BIGINT), e.slot(3, K.TYPE_DOUBLE)
    return ScanAggPlan(preds=[pred(2, K.PRED_GE, K.TYPE_BIGINT, -90)], exprs=e, keys=keys, group_hint=hint,
                       aggs=[(K.AGG_SUM, d), (K.AGG_AVG, d), (K.AGG_AVG, b), (K.AGG_SUM, b), (K.AGG_MIN, d), (K.AGG_MAX, dn), (K.AGG_SUM, dn),
                             (K.AGG_COUNT, dn), (K.AGG_COUNT, None)])


@pytest.mark.gpu
@pytest.mark.parametrize("key_slot,hint", [(0, 4), (4, 100_000), (None, 0)])
def test_compiled_shape_with_double_aggregates(ctx, oracle, jit_sync, key_slot, hint):
    from tests.test_gpu_parity import rows_equal
    chunk = double_chunk(500_003)
    want = K.chunk_rows(oracle.scan_agg(double_plan(key_slot, hint), chunk))
    got, prof = run_device(ctx, double_plan(key_slot, hint), chunk)
    assert prof["shape"] == SHAPE_JIT and prof["static_launches"] >= 1, prof
    assert rows_equal(K.chunk_rows(got), want, rel=1e-9)  # sums of doubles are added in another order (as everywhere)
