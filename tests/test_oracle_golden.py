"""Pins the CPU oracle to the reference's own known-answer tests. CPU only.

Each test names the reference test it restates (be/test/...:line). Expected values are the reference's
literals; inputs are rebuilt exactly as the reference's generators build them.
"""
import numpy as np
import pytest

from starrocks_b200 import capi as K
from starrocks_b200.plan import ExprBuilder, ScanAggPlan, JoinPlan, pred

I32MAX, I32MIN = 2**31 - 1, -2**31
I64MAX, I64MIN = 2**63 - 1, -2**63
I128MAX, I128MIN = 2**127 - 1, -2**127


def wrap(x, bits):
    x &= (1 << bits) - 1
    return x - (1 << bits) if x >= 1 << (bits - 1) else x


def cdiv(a, b):  # C truncating division
    q = abs(a) // abs(b)
    return q if (a < 0) == (b < 0) else -q


# ---- be/test/runtime/decimalv3_test.cpp:402-491 (prepare_add_data / prepare_sub_data / prepare_mul_data) ----
def _limits(width):
    return {4: (I32MAX, I32MIN), 8: (I64MAX, I64MIN), 16: (I128MAX, I128MIN)}[width]


def _add_cases(width):
    mx, mn = _limits(width)
    hmx, hmn = cdiv(mx, 2), cdiv(mn, 2)
    # decimalv3_test.cpp prepare_add_data (mirrors prepare_sub_data with the second operand negated)
    return [(0, 0, False), (hmx, -hmx, False), (hmn, -hmn, False), (hmx - 1, hmx - 1, False), (mx - hmx, hmx, False),
            (hmn + 1, hmn + 1, False), (mn - hmn, hmn, False), (mx, 1, True), (mx, hmx, True), (mx, mx, True),
            (mn, -1, True), (mn, hmn, True), (mn, mn, True)]


def _sub_cases(width):
    mx, mn = _limits(width)
    hmx, hmn = cdiv(mx, 2), cdiv(mn, 2)
    return [(0, 0, False), (hmx, hmx, False), (hmn, hmn, False), (hmx - 1, -(hmx - 1), False), (mx - hmx, -hmx, False),
            (hmn + 1, -(hmn + 1), False), (mn - hmn, -hmn, False), (mx, -1, True), (mx, -hmx, True), (mx, -mx, True),
            (mn, 1, True), (mn, -hmn, True), (mn, wrap(-mn, width * 8), False)]


def _mul_cases(width):
    mx, mn = _limits(width)
    v0 = 1 << (width // 2)  # `static_cast<T>(1) << (sizeof(T) / 2)`: sizeof is in bytes
    v1 = v0 // 2
    nmn = wrap(-mn, width * 8)
    return [(0, 0, False), (0, v0, False), (v0, 0, False), (v0, v0, False), (v1, v1, False), (v1 - 1, v1 - 1, False),
            (v1, v0, False), (v0, v1, False), (v0, v1 >> 1, False), (mx, mx, True), (mn, mn, True),
            (v0, -v0, False), (-v1, v1, False), (v1 - 1, -(v1 - 1), False), (v1, -v0, False), (v0, -v1, False),
            (v0, -(v1 >> 1), False), (mx, -mx, True), (mn, nmn, True),
            (-v0, -v0, False), (-v1, -v1, False), (-(v1 - 1), -(v1 - 1), False), (-v1, -v0, False), (-v0, -v1, False),
            (-v0, -(v1 >> 1), False), (-mx, -mx, True), (nmn, nmn, True)]


@pytest.mark.parametrize("width", [4, 8, 16])
def test_decimalv3_add_sub_mul(oracle, width):
    bits = width * 8
    for op, cases, f in ((K.EXPR_ADD, _add_cases(width), lambda a, b: a + b),
                         (K.EXPR_SUB, _sub_cases(width), lambda a, b: a - b),
                         (K.EXPR_MUL, _mul_cases(width), lambda a, b: a * b)):
        for a, b, expect_ov in cases:
            ov, z = oracle.checked_binop(width, op, a, b)
            assert ov == expect_ov, (width, op, a, b)
            if not expect_ov:  # the reference also pins the wrapped value; for non-overflow it is the exact one
                assert z == wrap(f(a, b), bits), (width, op, a, b)


def test_int128_mul_is_multi3_not_builtin(oracle):
    # int128_arithmetics_x86_64.h:184-229: |x|*|y| == 2^127 is flagged although -2^127 is representable
    ov, _ = oracle.checked_binop(16, K.EXPR_MUL, -(1 << 64), 1 << 63)
    assert ov is True
    ov, z = oracle.checked_binop(16, K.EXPR_MUL, -(1 << 63), (1 << 63) - 1)
    assert ov is False and z == -(1 << 63) * ((1 << 63) - 1)


# ---- be/test/exec/join_hash_map_test.cpp:914-1000 (JoinKeyHash) ------------------------------------------
def _occupancy(fn, keys, nb):
    b = np.fromiter((fn(int(k)) for k in keys), dtype=np.int64, count=len(keys))
    cnt = np.bincount(b, minlength=nb)
    return int(cnt.min()), int(cnt.max())


def test_join_key_hash_vectors(oracle):
    nb, lg = 1 << 16, 16
    L = oracle.lib
    h32 = lambda v: L.orc_join_key_hash32(v & 0xFFFFFFFF, lg)
    h64 = lambda v: L.orc_join_key_hash64(v, lg)
    assert _occupancy(h32, range(0, nb * 3 * 5, 3), nb) == (0, 11)
    assert _occupancy(h32, range(0, nb * 7 * 5, 7), nb) == (0, 14)
    assert _occupancy(h32, range(0, nb * 5), nb) == (4, 6)
    assert _occupancy(h64, range(0, nb * 3 * 5, 3), nb) == (3, 7)
    assert _occupancy(h64, range(0, nb * 7 * 5, 7), nb) == (4, 7)
    assert _occupancy(h64, range(0, nb * 5), nb) == (4, 6)
    assert L.orc_join_key_hash_bytes(b"abcd", 4, nb) == 2777932099 % nb


def test_calc_bucket_num(oracle):
    # :1002-1023
    L = oracle.lib
    assert L.orc_join_key_hash32(1, 2) == 2
    assert [L.orc_join_key_hash32(v, 2) for v in (1, 2, 3, 4)] == [2, 0, 3, 1]


def test_crc32c_matches_hardware(oracle):
    L = oracle.lib
    if not L.orc_has_hw_crc():
        pytest.skip("no SSE4.2")
    rng = np.random.default_rng(7)
    for _ in range(200):
        v = int(rng.integers(0, 2**63)) * 2 + int(rng.integers(0, 2))
        seed = int(rng.integers(0, 2**32))
        import struct
        sw = L.orc_crc_hash_64(struct.pack("<Q", v), 8, seed)
        # crc_hash_64 over exactly 8 bytes = two identical-word crc steps (util/hash.h:145-167)
        hw = L.orc_hw_crc32c_u64(L.orc_hw_crc32c_u64(seed, v), v)
        assert sw == hw


# ---- be/test/exprs/agg/aggregate_test.cpp -----------------------------------------------------------------
COL1 = list(range(1024)) + [100, 200]   # gen_input_column1 :104-113
COL2 = list(range(2000, 3000))          # gen_input_column2 :127-133


def _agg(oracle, fn, type_, values, precision=0, scale=0, nulls=None, phase=K.AGG_UPDATE_FINALIZE, threads=1):
    e = ExprBuilder()
    arg = e.slot(0, type_, precision, scale)
    plan = ScanAggPlan(exprs=e, aggs=[(fn, arg)], phase=phase)
    ch = K.Chunk(len(values), {0: K.Column.from_ints(type_, values, nulls, precision, scale)})
    return oracle.scan_agg(plan, ch, n_threads=threads, chunk_size=256)


def _two_phase(oracle, fn, type_, v1, v2, precision=0, scale=0):
    """update+serialize each input, then merge+finalize both intermediates (test_agg_function :224-261)."""
    inter = [_agg(oracle, fn, type_, v, precision, scale, phase=K.AGG_UPDATE_SERIALIZE) for v in (v1, v2)]
    c0, c1 = inter[0].columns[0], inter[1].columns[0]
    if K.is_binary(c0.type):
        blobs = K.column_pyvalues(c0) + K.column_pyvalues(c1)
        col = K.Column.from_strings(blobs, type_=K.TYPE_VARBINARY)
        col.precision, col.scale = c0.precision, c0.scale
    else:
        vals = K.column_pyvalues(c0) + K.column_pyvalues(c1)
        col = K.Column.from_ints(c0.type, vals, None, c0.precision, c0.scale)
    e = ExprBuilder()
    arg = e.slot(0, col.type, col.precision, col.scale)
    plan = ScanAggPlan(exprs=e, aggs=[(fn, arg)], phase=K.AGG_MERGE_FINALIZE)
    return K.column_pyvalues(oracle.scan_agg(plan, K.Chunk(2, {0: col})).columns[0])[0]


def _one(chunk):
    return K.column_pyvalues(chunk.columns[0])[0]


@pytest.mark.parametrize("type_", [K.TYPE_SMALLINT, K.TYPE_INT, K.TYPE_BIGINT, K.TYPE_LARGEINT, K.TYPE_DOUBLE])
def test_sum_count_avg_max_min(oracle, type_):
    # test_count :331-343, test_sum :345-367, test_avg :399-414, test_max :957-976, test_min :993-1012
    assert _one(_agg(oracle, K.AGG_COUNT, type_, COL1)) == 1026
    assert _one(_agg(oracle, K.AGG_COUNT, type_, COL2)) == 1000
    assert _two_phase(oracle, K.AGG_COUNT, type_, COL1, COL2) == 2026
    assert _one(_agg(oracle, K.AGG_SUM, type_, COL1)) == 524076
    assert _one(_agg(oracle, K.AGG_SUM, type_, COL2)) == 2499500
    assert _two_phase(oracle, K.AGG_SUM, type_, COL1, COL2) == 3023576
    assert _one(_agg(oracle, K.AGG_AVG, type_, COL1)) == 524076 / 1026.0
    assert _one(_agg(oracle, K.AGG_AVG, type_, COL2)) == 2499500 / 1000.0
    assert _two_phase(oracle, K.AGG_AVG, type_, COL1, COL2) == 3023576 / 2026.0
    assert _one(_agg(oracle, K.AGG_MAX, type_, COL1)) == 1023
    assert _one(_agg(oracle, K.AGG_MAX, type_, COL2)) == 2999
    assert _two_phase(oracle, K.AGG_MAX, type_, COL1, COL2) == 2999
    assert _one(_agg(oracle, K.AGG_MIN, type_, COL1)) == 0
    assert _one(_agg(oracle, K.AGG_MIN, type_, COL2)) == 2000
    assert _two_phase(oracle, K.AGG_MIN, type_, COL1, COL2) == 0


@pytest.mark.parametrize("type_,p,s", [(K.TYPE_DECIMAL32, 9, 9), (K.TYPE_DECIMAL64, 9, 3), (K.TYPE_DECIMAL128, 38, 15)])
def test_decimal_sum_avg(oracle, type_, p, s):
    # test_decimal_sum :369-397: every decimal width sums into DECIMAL128 with the same integers
    r = _agg(oracle, K.AGG_SUM, type_, COL1, p, s)
    assert r.columns[0].type == K.TYPE_DECIMAL128 and r.columns[0].scale == s
    assert _one(r) == 524076
    assert _one(_agg(oracle, K.AGG_SUM, type_, COL2, p, s)) == 2499500
    assert _two_phase(oracle, K.AGG_SUM, type_, COL1, COL2, p, s) == 3023576
    # test_decimal_avg :431-468: decimal_div_integer(sum, count, scale)
    ddi = oracle.decimal_div_integer
    assert _one(_agg(oracle, K.AGG_AVG, type_, COL1, p, s)) == ddi(524076, 1026, s)
    assert _one(_agg(oracle, K.AGG_AVG, type_, COL2, p, s)) == ddi(2499500, 1000, s)
    assert _two_phase(oracle, K.AGG_AVG, type_, COL1, COL2, p, s) == ddi(3023576, 2026, s)


def test_decimal_div_integer_is_half_away_from_zero(oracle):
    # arithmetic_operation.h:409-441 + decimalv3.h:86-100, checked against exact rational arithmetic
    from fractions import Fraction
    for s, count, scale in [(524076, 1026, 9), (2499500, 1000, 3), (3023576, 2026, 15), (-524076, 1026, 2), (7, 2, 0), (-7, 2, 0),
                            (1, 3, 2), (2, 3, 2), (-2, 3, 2)]:
        rs = scale + 6 if scale <= 6 else (12 if scale <= 12 else scale)
        q = Fraction(s * 10 ** (rs - scale), count)
        fl = q.numerator // q.denominator
        frac = q - fl
        expect = fl + (1 if (frac > Fraction(1, 2) or (frac == Fraction(1, 2) and q > 0)) else 0)
        assert oracle.decimal_div_integer(s, count, scale) == expect


def test_date_avg_pins_from_date(oracle):
    # test_avg :428-429 with gen_input_column1/2<DateValue> :193-212 — pins date::from_date for 44 dates
    L = oracle.lib
    d1 = [L.orc_from_date(2000 + j, 1, 1) for j in range(20)] + [L.orc_from_date(2000, 1, 1), L.orc_from_date(2001, 1, 1)]
    d2 = [L.orc_from_date(1000 + j, 1, 1) for j in range(20)] + [L.orc_from_date(2000, 1, 1), L.orc_from_date(1000, 1, 1)]
    assert _one(_agg(oracle, K.AGG_AVG, K.TYPE_DATE, d1)) == 2454716
    assert _one(_agg(oracle, K.AGG_AVG, K.TYPE_DATE, d2)) == 2106058
    assert (sum(d1) + sum(d2)) // 44 == 2280387
    assert L.orc_from_date(1970, 1, 1) == 2440588  # date::UNIX_EPOCH_JULIAN, runtime/time_types.h:177
    # max/min DATE :990-991
    assert _one(_agg(oracle, K.AGG_MAX, K.TYPE_DATE, d1)) == L.orc_from_date(2019, 1, 1)


def test_nullable_sum_count(oracle):
    # test_sum_nullable :1250-1297 / test_count_nullable :1299-1318: NULL rows are skipped; all-NULL sum is NULL
    vals = list(range(100))
    nulls = [1 if i % 3 == 0 else 0 for i in range(100)]
    expect = sum(v for v, n in zip(vals, nulls) if not n)
    assert _one(_agg(oracle, K.AGG_SUM, K.TYPE_INT, vals, nulls=nulls)) == expect
    assert _one(_agg(oracle, K.AGG_COUNT, K.TYPE_INT, vals, nulls=nulls)) == 100 - sum(nulls)
    assert _one(_agg(oracle, K.AGG_SUM, K.TYPE_INT, vals, nulls=[1] * 100)) is None
    assert _one(_agg(oracle, K.AGG_COUNT, K.TYPE_INT, vals, nulls=[1] * 100)) == 0
    # no-group-by over empty input: sum is NULL, count is 0 (aggregator.cpp:938-945)
    assert _one(_agg(oracle, K.AGG_SUM, K.TYPE_INT, [])) is None
    assert _one(_agg(oracle, K.AGG_COUNT, K.TYPE_INT, [])) == 0


def test_threads_do_not_change_results(oracle):
    for t in (1, 2, 5):
        assert _one(_agg(oracle, K.AGG_SUM, K.TYPE_BIGINT, COL1, threads=t)) == 524076


# ---- join tables: be/test/exec/join_hash_map_test.cpp:1126-1220, 1964-2030 --------------------------------
def test_join_build_chain_layout(oracle):
    import ctypes as C
    keys = np.arange(0, 10, dtype=np.int32)  # JoinBuildFunc test: keys 0..9 after the dummy row
    bs = C.c_uint32()
    oracle.lib.orc_join_build_raw(keys.ctypes.data, 4, 10, C.byref(bs), None, None)
    assert bs.value == oracle.lib.orc_calc_bucket_size(11) == 16  # NormalizeCapacity(11 + 10/4) + 1
    first = np.zeros(bs.value, np.uint32)
    nxt = np.zeros(11, np.uint32)
    oracle.lib.orc_join_build_raw(keys.ctypes.data, 4, 10, C.byref(bs), first.ctypes.data, nxt.ctypes.data)
    # every row is reachable exactly once by walking first[]/next[]; row 0 is the sentinel
    seen = []
    for b in range(bs.value):
        i = int(first[b])
        while i:
            assert oracle.lib.orc_join_key_hash32(int(keys[i - 1]), 4) == b
            seen.append(i)
            i = int(nxt[i])
    assert sorted(seen) == list(range(1, 11))


def test_join_inner_one_to_many_and_nulls(oracle):
    build = K.Chunk(6, {0: K.Column.from_ints(K.TYPE_INT, [1, 2, 2, 3, 5, 7], nulls=[0, 0, 0, 0, 1, 0]),
                        1: K.Column.from_ints(K.TYPE_BIGINT, [10, 20, 21, 30, 50, 70])})
    probe = K.Chunk(5, {2: K.Column.from_ints(K.TYPE_INT, [2, 3, 4, 5, 7], nulls=[0, 0, 0, 0, 1]),
                        3: K.Column.from_ints(K.TYPE_BIGINT, [200, 300, 400, 500, 700])})
    plan = JoinPlan([0], [2], build_out=[1], probe_out=[3])
    rows = K.chunk_rows(oracle.join(plan, build, probe))
    assert rows == [(200, 20), (200, 21), (300, 30)]  # NULL keys never match (tpp:62-72)
    # LEFT SEMI: one row per matching probe row; LEFT ANTI: the rest, the NULL probe key included (tpp:1349-1420, PlanNodes.thrift:710-713)
    semi = JoinPlan([0], [2], probe_out=[3], join_type=K.JOIN_LEFT_SEMI)
    anti = JoinPlan([0], [2], probe_out=[3], join_type=K.JOIN_LEFT_ANTI)
    assert K.chunk_rows(oracle.join(semi, build, probe), sort=False) == [(200,), (300,)]
    assert K.chunk_rows(oracle.join(anti, build, probe), sort=False) == [(400,), (500,), (700,)]


def test_fixed_size_key_packing(oracle):
    # CompileFixedSizeKeyColumn :1041-1056: (0,2)->8589934592, (1,3)->12884901889 — two int32 keys pack low-to-high
    build = K.Chunk(2, {0: K.Column.from_ints(K.TYPE_INT, [0, 1]), 1: K.Column.from_ints(K.TYPE_INT, [2, 3]),
                        9: K.Column.from_ints(K.TYPE_BIGINT, [8589934592, 12884901889])})
    probe = K.Chunk(2, {4: K.Column.from_ints(K.TYPE_BIGINT, [8589934592, 12884901889])})
    import ctypes as C
    cch, keep = build.to_c()
    out = np.zeros(2, dtype=K.I128)
    slots = (C.c_int32 * 2)(0, 1)
    assert oracle.lib.orc_pack_keys(C.byref(cch), slots, 2, C.c_void_p(out.ctypes.data)) == 8
    assert K.i128_to_pyints(out) == [8589934592, 12884901889]
    # and a join on the packed BIGINT finds exactly those rows
    plan = JoinPlan([9], [4], build_out=[0, 1], probe_out=[4])
    assert K.chunk_rows(oracle.join(plan, build, probe)) == [(8589934592, 0, 2), (12884901889, 1, 3)]


# ---- runtime filter (oracle/orc_rf.h) pinned by be/test/exprs/runtime_filter_test.cpp ----------------------------------------
def test_simd_block_filter_known_answers(oracle):
    # TestSimdBlockFilter :255-265: init(100), insert_hash(1, 18, ..., 188); every inserted hash tests true, every hash + 1 false
    L = oracle.lib
    f = L.orc_rf_create(100, 0)
    try:
        assert L.orc_rf_info(f, 0) == 4  # init(100): ceil(log2(100)) = 7 -> 2^(7-5) buckets of 32 bytes (runtime_filter.cpp:26-35)
        for i in range(1, 201, 17):
            L.orc_rf_insert_hash(f, i)
        for i in range(1, 201, 17):
            assert L.orc_rf_test_hash(f, i) == 1
            assert L.orc_rf_test_hash(f, i + 1) == 0
    finally:
        L.orc_rf_free(f)


def test_join_runtime_filter_known_answers(oracle):
    # TestJoinRuntimeFilter :342-380 (Int32RF): insert 0, 17, ..., 187 -> min 0, max 187; test_data(i) true, test_data(i + 1) false;
    # evaluating the column 0..200 keeps exactly the 12 inserted values; has_null only after insert_null
    from tests.oracle_lib import OracleRuntimeFilter
    rf = OracleRuntimeFilter(oracle, 100)
    try:
        vals = np.arange(0, 201, 17, dtype=np.int32)
        rf.insert_column(K.Column.from_ints(K.TYPE_INT, vals))
        info = rf.info()
        assert (info["min"], info["max"], info["has_null"], info["inserted"]) == (0, 187, False, 12)
        for i in vals:
            assert oracle.lib.orc_rf_test_data(rf.h, int(i)) == 1
            assert oracle.lib.orc_rf_test_data(rf.h, int(i) + 1) == 0
        sel = rf.evaluate(K.Column.from_ints(K.TYPE_INT, np.arange(0, 201, dtype=np.int32)))
        assert sel.sum() == 12 and list(np.nonzero(sel)[0]) == list(vals)
        # NULL probe rows follow has_null (runtime_filter.h:1396-1401)
        col = K.Column.from_ints(K.TYPE_INT, [17, 17, 18], nulls=[0, 1, 1])
        assert list(rf.evaluate(col)) == [1, 0, 0]
        rf.insert_column(K.Column.from_ints(K.TYPE_INT, [0], nulls=[1]))
        assert rf.info()["has_null"] and list(rf.evaluate(col)) == [1, 1, 1]
    finally:
        rf.close()


def test_runtime_filter_sizes_and_negative_keys(oracle):
    # SimdBlockFilter::init: log_num_buckets = max(1, ceil(log2(nums)) - 5); the hash of an integer is its sign-extended value
    L = oracle.lib
    for nums, buckets in ((0, 2), (1, 2), (64, 2), (65, 4), (1 << 20, 1 << 15), ((1 << 20) + 1, 1 << 16)):
        f = L.orc_rf_create(nums, 0)
        assert L.orc_rf_info(f, 0) == buckets
        L.orc_rf_free(f)
    from tests.oracle_lib import OracleRuntimeFilter
    rf = OracleRuntimeFilter(oracle, 1000)
    try:
        rf.insert_column(K.Column.from_ints(K.TYPE_BIGINT, [-5, 7, -(1 << 40)]))
        assert (rf.info()["min"], rf.info()["max"]) == (-(1 << 40), 7)
        assert list(rf.evaluate(K.Column.from_ints(K.TYPE_BIGINT, [-5, 7, -(1 << 40), 8, -(1 << 41)]))) == [1, 1, 1, 0, 0]
        # the same key inserted through a narrower column sets the same bits (sign extension)
        rf2 = OracleRuntimeFilter(oracle, 1000)
        rf2.insert_column(K.Column.from_ints(K.TYPE_INT, [-5, 7]))
        rf3 = OracleRuntimeFilter(oracle, 1000)
        rf3.insert_column(K.Column.from_ints(K.TYPE_BIGINT, [-5, 7]))
        assert (rf2.directory() == rf3.directory()).all()
        rf2.close()
        rf3.close()
    finally:
        rf.close()


def test_runtime_filter_with_the_mixing_hash(oracle):
    # the same source with ENABLE_PHMAP_MIX (phmap_utils.h:85-95): hash = hi + lo of key * 0xde5fb9d2630458e9 (= phase-1 phmap_mix8)
    from tests.oracle_lib import OracleRuntimeFilter
    rf = OracleRuntimeFilter(oracle, 100, mix=True)
    plain = OracleRuntimeFilter(oracle, 100)
    try:
        vals = np.arange(0, 201, 17, dtype=np.int32)
        col = K.Column.from_ints(K.TYPE_INT, vals)
        rf.insert_column(col)
        plain.insert_column(col)
        assert rf.evaluate(col).all() and (rf.directory() != plain.directory()).any()
        # equal to inserting the mixed hashes directly
        f = oracle.lib.orc_rf_create(100, 0)
        for v in vals:
            oracle.lib.orc_rf_insert_hash(f, oracle.lib.orc_phmap_mix8(int(v), 1))
        d = np.empty((4, 8), np.uint32)
        oracle.lib.orc_rf_directory(f, d.ctypes.data)
        oracle.lib.orc_rf_free(f)
        assert (d == rf.directory()).all()
        # sparse keys (8 of every 32, like TPC-H order keys): the identity hash leaves 3/4 of the buckets empty, the mix does not
        all_idx = np.random.default_rng(1).permutation(400_000)
        idx, other = np.sort(all_idx[:200_000]), np.sort(all_idx[200_000:300_000])  # inserted keys / keys of the same domain that are not
        keys = K.Column.from_ints(K.TYPE_INT, ((idx // 8) * 32 + idx % 8 + 1).astype(np.int32))
        a, b = OracleRuntimeFilter(oracle, 200_000), OracleRuntimeFilter(oracle, 200_000, mix=True)
        a.insert_column(keys)
        b.insert_column(keys)
        used = lambda f: int((f.directory().any(axis=1)).sum())
        assert used(a) * 3 < used(b)
        probe = K.Column.from_ints(K.TYPE_INT, ((other // 8) * 32 + other % 8 + 1).astype(np.int32))
        assert b.evaluate(probe).mean() < 0.05 < a.evaluate(probe).mean()
        a.close()
        b.close()
    finally:
        rf.close()
        plain.close()


# ---- be/test/storage/column_predicate_test.cpp (test_gt :1033, test_ge :1157, test_lt :1281, test_le :1405, test_is_null :1977,
#      test_not_null :2038): nullable INT column {1, 2, 3, NULL, 4}, literal "3"; evaluate() writes these selection bytes ----------
PREDICATE_VECTORS = [(K.PRED_GT, "0,0,0,0,1"), (K.PRED_GE, "0,0,1,0,1"), (K.PRED_LT, "1,1,0,0,0"), (K.PRED_LE, "1,1,1,0,0"),
                     (K.PRED_IS_NULL, "0,0,0,1,0"), (K.PRED_NOT_NULL, "1,1,1,0,1"),
                     # test_eq / test_ne's INT blocks use the same column shape; a NULL never compares equal or unequal
                     (K.PRED_EQ, "0,0,1,0,0"), (K.PRED_NE, "1,1,0,0,1")]


def predicate_vector_chunk(type_=K.TYPE_INT):
    nulls = np.array([0, 0, 0, 1, 0], np.uint8)
    return K.Chunk(5, {0: K.Column.from_ints(type_, [1, 2, 3, 0, 4], nulls), 1: K.Column.from_ints(K.TYPE_INT, np.arange(5, dtype=np.int32))})


def selection_of(kept_chunk, n):
    sel = ["0"] * n
    for r in K.column_pyvalues(kept_chunk.columns[1]):
        sel[r] = "1"
    return ",".join(sel)


@pytest.mark.parametrize("op,want", PREDICATE_VECTORS)
@pytest.mark.parametrize("type_", [K.TYPE_TINYINT, K.TYPE_SMALLINT, K.TYPE_INT, K.TYPE_BIGINT, K.TYPE_DATE])
def test_column_predicate_known_answers(oracle, op, want, type_):
    from starrocks_b200 import queries as Q
    got = oracle.scan_agg(Q.filter_plan([pred(0, op, type_, 3)], [1]), predicate_vector_chunk(type_))
    assert selection_of(got, 5) == want


def test_conjunction_is_evaluate_and(oracle):
    # column_predicate_test.cpp:1060-1075: evaluate_and on an all-ones selection equals evaluate; two predicates AND their selections
    from starrocks_b200 import queries as Q
    got = oracle.scan_agg(Q.filter_plan([pred(0, K.PRED_GE, K.TYPE_INT, 2), pred(0, K.PRED_LE, K.TYPE_INT, 3)], [1]), predicate_vector_chunk())
    assert selection_of(got, 5) == "0,1,1,0,0"


# ---- be/test/column/fixed_length_column_test.cpp:58-74 (filter i % 2 over 0..99 keeps 2k+1) and :130-163 (nullable column, NULL
#      where i % 3 != 0, filter keeps rows 50..99: row j is NULL iff (j + 50) % 3 != 0, else j + 50) ------------------------------
def filter_vector_chunks():
    v = np.arange(100, dtype=np.int32)
    plain = K.Chunk(100, {0: K.Column.from_ints(K.TYPE_INT, v), 1: K.Column.from_ints(K.TYPE_TINYINT, (v % 2).astype(np.int8))})
    nulls = (v % 3 != 0).astype(np.uint8)
    nullable = K.Chunk(100, {0: K.Column.from_ints(K.TYPE_INT, np.where(nulls, 0, v).astype(np.int32), nulls), 1: K.Column.from_ints(K.TYPE_INT, v)})
    return plain, nullable


def check_filter_vectors(run):
    from starrocks_b200 import queries as Q
    plain, nullable = filter_vector_chunks()
    got = run(Q.filter_plan([pred(1, K.PRED_EQ, K.TYPE_TINYINT, 1)], [0]), plain)
    assert got.n_rows == 50 and K.column_pyvalues(got.columns[0]) == [2 * k + 1 for k in range(50)]
    got = run(Q.filter_plan([pred(1, K.PRED_GE, K.TYPE_INT, 50)], [0]), nullable)
    assert got.n_rows == 50
    assert K.column_pyvalues(got.columns[0]) == [None if (j + 50) % 3 else j + 50 for j in range(50)]


def test_column_filter_known_answers(oracle):
    check_filter_vectors(lambda plan, ch: oracle.scan_agg(plan, ch))


# ---- be/test/exec/chunks_sorter_test.cpp: the three fixture chunks (:441-482) and the sorted orders the reference asserts ----------
def sorter_fixture():
    cust = [2, 12, 41, 54, 58, 71, 4, 16, 49, 55, 69, 6, 24, 52, 56, 70]
    nulls = [0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1]
    nation = ["JORDAN", "JORDAN", "IRAN", "EGYPT", "JORDAN", "", "EGYPT", "IRAN", "IRAN", "IRAN", "null0", "SAUDI ARABIA", "JORDAN", "IRAQ", "IRAN", "null1"]
    region = ["MIDDLE EAST"] * 5 + [""] + ["MIDDLE EAST"] * 4 + ["null0"] + ["MIDDLE EAST"] * 4 + ["null1"]
    mkt = ["AUTOMOBILE", "HOUSEHOLD", "HOUSEHOLD", "AUTOMOBILE", "HOUSEHOLD", "", "MACHINERY", "FURNITURE", "FURNITURE", "MACHINERY", "null0",
           "AUTOMOBILE", "MACHINERY", "HOUSEHOLD", "FURNITURE", "null1"]

    def strcol(v):
        c = K.Column.from_strings([x.encode() for x in v])
        c.nulls = np.array(nulls, np.uint8)
        return c
    return K.Chunk(16, {0: K.Column.from_ints(K.TYPE_INT, np.array(cust, np.int32)), 1: strcol(nation), 2: strcol(region), 3: strcol(mkt)})


CUST, NATION, REGION, MKT = 0, 1, 2, 3
SORT_VECTORS = [
    # (key slots, is_asc, nulls_first, expected cust_key order)            chunks_sorter_test.cpp
    ([REGION, CUST], [0, 1], [1, 1], [69, 70, 71, 2, 4, 6, 12, 16, 24, 41, 49, 52, 54, 55, 56, 58]),             # :799-837
    ([REGION, CUST], [1, 0], [0, 0], [58, 56, 55, 54, 52, 49, 41, 24, 16, 12, 6, 4, 2, 71, 70, 69]),             # :839-877
    ([REGION, NATION, CUST], [0, 1, 0], [1, 1, 1], [71, 70, 69, 54, 4, 56, 55, 49, 41, 16, 52, 58, 24, 12, 2, 6]),  # :879-920
    ([MKT, REGION, NATION, CUST], [0, 1, 0, 0], [0, 1, 1, 0], [24, 55, 4, 58, 12, 52, 41, 56, 49, 16, 6, 2, 54, 71, 70, 69]),  # :922-966
]


def cust_keys(ch):
    return K.column_pyvalues(ch.columns[0])


@pytest.mark.parametrize("keys,asc,nf,want", SORT_VECTORS)
def test_full_sort_known_answers(oracle, keys, asc, nf, want):
    assert cust_keys(oracle.sort(sorter_fixture(), keys, asc, nf)) == want


def test_topn_sort_with_limit(oracle):
    # :703-747: ORDER BY <column>, cust_key (ascending, NULLs first) LIMIT 1..15
    cases = [(CUST, [2, 4, 6, 12, 16, 24, 41, 49, 52, 54, 55, 56, 58, 69, 70, 71]),
             (NATION, [69, 70, 71, 4, 54, 16, 41, 49, 55, 56, 52, 2, 12, 24, 58]),
             (REGION, [69, 70, 71, 2, 4, 6, 12, 16, 24, 41, 49, 52, 54, 55, 56])]
    ch = sorter_fixture()
    for col, expected in cases:
        for limit in range(1, 16):
            assert cust_keys(oracle.sort(ch, [col, CUST], [1, 1], [1, 1], 0, limit)) == expected[:limit]


def test_topn_offset_and_limit(oracle):
    ch = sorter_fixture()
    # :969-1006: region DESC, nation, cust_key, NULLs first, OFFSET 2 LIMIT 7
    assert cust_keys(oracle.sort(ch, [REGION, NATION, CUST], [0, 1, 1], [1, 1, 1], 2, 7)) == [71, 4, 54, 16, 41, 49, 55]
    # :1008-1060: NULLs last, OFFSET 7, LIMIT 8..9; an offset past the end gives nothing
    for limit in (8, 9):
        assert cust_keys(oracle.sort(ch, [REGION, NATION, CUST], [0, 1, 1], [0, 0, 0], 7, limit)) == [52, 2, 12, 24, 58, 6, 69, 70, 71][:limit]
        assert oracle.sort(ch, [REGION, NATION, CUST], [0, 1, 1], [0, 0, 0], 100, limit).n_rows == 0


def test_sort_is_stable_and_carries_every_column(oracle):
    ch = sorter_fixture()
    out = oracle.sort(ch, [REGION], [1], [1])  # ties everywhere: arrival order within NULLs and within 'MIDDLE EAST'
    assert cust_keys(out) == [71, 69, 70, 2, 12, 41, 54, 58, 4, 16, 49, 55, 6, 24, 52, 56]
    rows_in = {r[0]: r for r in K.chunk_rows(ch)}
    assert all(rows_in[r[0]] == r for r in K.chunk_rows(out))


# ---- be/test/exec/agg_hash_map_test.cpp:232-292 (AggHashMapKeyNotFoundsTest): the streaming sink's lookups. Five key chunks; the first
#      is aggregated, the others run build_hash_map_with_selection (1) / ..._and_allocation (2); expected not-found flags :290 ---------
LOOKUP_DATA = [[1, 2, 1, 1], [1, 2, 1, 1], [1, 2, 3, 3], [4, 4, 4, 4], [5, 5, 5, 5]]
LOOKUP_OPS = [0, 1, 2, 1, 2]
LOOKUP_EXPECT = [[0, 0, 0, 0], [0, 0, 1, 0], [1, 1, 1, 1], [1, 0, 0, 0]]


@pytest.mark.parametrize("kind", ["int32", "nullable int32", "string", "nullable string", "int32 + int32 (fixed-size serialized key)"])
def test_agg_hash_map_not_found_selection(oracle, kind):
    def col(vals):
        nulls = np.zeros(len(vals), np.uint8) if "nullable" in kind else None
        if "string" in kind:
            return K.Column.from_strings([str(v).encode() for v in vals], nulls)
        return K.Column.from_ints(K.TYPE_INT, np.array(vals, np.int32), nulls)
    two = kind.startswith("int32 +")
    chunks = [K.Chunk(4, {0: col(v), **({1: col([7] * 4)} if two else {})}) for v in LOOKUP_DATA]
    e = ExprBuilder()
    keys = [e.slot(0, K.TYPE_VARCHAR if "string" in kind else K.TYPE_INT)] + ([e.slot(1, K.TYPE_INT)] if two else [])
    plan = ScanAggPlan(exprs=e, keys=keys, aggs=[(K.AGG_COUNT, None)], phase=K.AGG_UPDATE_SERIALIZE)
    assert oracle.agg_lookup_sequence(plan, chunks, LOOKUP_OPS) == LOOKUP_EXPECT


# ---- segment data pages: frame-of-reference coding (oracle/orc_page.cpp) -----------------------------------------------------------
def test_for_coding_round_trips_of_the_reference_tests(oracle):
    # be/test/util/frame_of_reference_coding_test.cpp:70-84 (half / one / two / two and a half frames), :98-115 (int64), :117-159
    for n in (64, 128, 256, 320):
        a = np.arange(n, dtype=np.int32)
        assert (oracle.for_decode(oracle.for_encode(a), np.int32) == a).all()
    a = np.arange(320, dtype=np.int64)
    assert (oracle.for_decode(oracle.for_encode(a), np.int64) == a).all()
    assert oracle.for_decode(oracle.for_encode(np.array([2019], np.int32)), np.int32).tolist() == [2019]
    assert oracle.for_decode(oracle.for_encode(np.array([2019, 2020], np.int32)), np.int32).tolist() == [2019, 2020]
    empty = oracle.for_encode(np.zeros(0, np.int32))
    assert len(empty) == 4 + 1 and len(oracle.for_decode(empty, np.int32)) == 0  # TestZeroValue


def test_for_page_sizes_and_bits_of_the_reference_tests(oracle):
    # be/test/storage/rowset/frame_of_reference_page_test.cpp:269-283: 0..127 ascending -> deltas of 1 bit: 4 + 16 + footer 7 = 27 bytes
    page = oracle.for_encode(np.arange(128, dtype=np.int32))
    assert len(page) == 27 and page[-7] == 1 and page[-6] == 1 and page[-5] == 128 and int.from_bytes(page[-4:], "little") == 128
    # :304-318: 128..1 descending -> value - min in 7 bits: 4 + 112 + 7 = 123 bytes
    page = oracle.for_encode((128 - np.arange(128)).astype(np.int32))
    assert len(page) == 123 and page[-7] == 0 and page[-6] == 7
    # :320-335 TestFindBitsOfInt
    assert [oracle.bits(0x06, 1), oracle.bits(0x0F, 1), oracle.bits(0x000100FF, 4), oracle.bits(0x1FFFFFFFF, 8), oracle.bits(0, 16)] == [3, 4, 17, 33, 0]


def test_for_page_layout_bit_for_bit(oracle):
    # frame_of_reference_coding.cpp:95-117: values packed MSB first; the example of the comment: 1, 2, 4, 8 in 4 bits -> 0001 0010 0100 1000
    page = oracle.for_encode(np.array([11, 3, 10, 5, 7, 18], np.int32))  # min 3, deltas 8 0 7 2 4 15 in 4 bits
    assert page[:4] == (3).to_bytes(4, "little") and page[4:7] == bytes([0x80, 0x72, 0x4F]) and page[7:9] == bytes([0, 4])
    rng = np.random.default_rng(5)
    for dt, lo, hi in ((np.int8, -60, 60), (np.int16, -9000, 9000), (np.int32, -2**30, 2**30), (np.int64, -2**61, 2**61)):
        for n in (1, 127, 128, 129, 1000, 4096):
            a = rng.integers(lo, hi, n).astype(dt)
            assert (oracle.for_decode(oracle.for_encode(a), dt) == a).all()
            s = np.sort(a)  # non-decreasing frames take the delta format
            p = oracle.for_encode(s)
            assert (oracle.for_decode(p, dt) == s).all() and p[-5 - 2 * ((n + 127) // 128)] == 1
    # the overflow format (2) keeps the values themselves: one frame whose range does not fit the type
    a = np.array([-2**31, 2**31 - 1, 0, 5], np.int32)
    p = oracle.for_encode(a)
    assert p[-7] == 2 and p[-6] == 32 and (oracle.for_decode(p, np.int32) == a).all()
    wide = K.to_i128_array([3 * 10**30, -7 * 10**29, 12345, 0, 10**37])
    got = oracle.for_decode(oracle.for_encode(wide), K.I128)
    assert (got["lo"] == wide["lo"]).all() and (got["hi"] == wide["hi"]).all()


def test_joins_against_an_empty_hash_table_like_the_reference(oracle):
    # be/test/exec/join_hash_map_test.cpp:2376-2388 EmptyHashMapTest: check_empty_hash_map(join type, 5 probe rows, expected rows, columns)
    from starrocks_b200.plan import JoinPlan
    build = K.Chunk(0, {0: K.Column.from_ints(K.TYPE_INT, np.zeros(0, np.int32)), 1: K.Column.from_ints(K.TYPE_BIGINT, np.zeros(0, np.int64))})
    probe = K.Chunk(5, {2: K.Column.from_ints(K.TYPE_INT, np.arange(5, dtype=np.int32)), 3: K.Column.from_ints(K.TYPE_BIGINT, np.arange(5))})
    expect = {K.JOIN_LEFT_OUTER: 5, K.JOIN_FULL_OUTER: 5, K.JOIN_LEFT_ANTI: 5, K.JOIN_INNER: 0, K.JOIN_LEFT_SEMI: 0, K.JOIN_RIGHT_SEMI: 0,
              K.JOIN_RIGHT_OUTER: 0, K.JOIN_RIGHT_ANTI: 0, K.JOIN_NULL_AWARE_LEFT_ANTI: 5}
    for jt, rows in expect.items():
        out = oracle.join(JoinPlan([0], [2], build_out=[1], probe_out=[2, 3], join_type=jt), build, probe)
        assert out.n_rows == rows, jt
    # a small hand-checked case: build keys 1, 2, 2, NULL; probe keys 2, 3, 1, NULL
    build = K.Chunk(4, {0: K.Column.from_ints(K.TYPE_INT, np.array([1, 2, 2, 9], np.int32), np.array([0, 0, 0, 1], np.uint8)),
                        1: K.Column.from_ints(K.TYPE_BIGINT, np.array([10, 20, 21, 90]))})
    probe = K.Chunk(4, {2: K.Column.from_ints(K.TYPE_INT, np.array([2, 3, 1, 7], np.int32), np.array([0, 0, 0, 1], np.uint8)),
                        3: K.Column.from_ints(K.TYPE_BIGINT, np.array([200, 300, 100, 700]))})
    rows = lambda jt: K.chunk_rows(oracle.join(JoinPlan([0], [2], build_out=[1], probe_out=[2, 3], join_type=jt), build, probe), sort=False)
    inner = [(2, 200, 21), (2, 200, 20), (1, 100, 10)]
    assert rows(K.JOIN_INNER) == inner
    assert rows(K.JOIN_LEFT_OUTER) == [(2, 200, 21), (2, 200, 20), (3, 300, None), (1, 100, 10), (None, 700, None)]
    assert rows(K.JOIN_RIGHT_OUTER) == inner + [(None, None, 90)]
    assert rows(K.JOIN_FULL_OUTER) == [(2, 200, 21), (2, 200, 20), (3, 300, None), (1, 100, 10), (None, 700, None), (None, None, 90)]
    assert rows(K.JOIN_RIGHT_SEMI) == [(10,), (20,), (21,)] and rows(K.JOIN_RIGHT_ANTI) == [(90,)]
    # NOT IN: LEFT ANTI keeps the NULL probe key, the NULL-aware type sees a NULL on the build side and keeps nothing
    # (hash_joiner.h:386-396); without that NULL it drops the NULL probe key (join_hash_map.tpp:1356-1383)
    assert rows(K.JOIN_LEFT_ANTI) == [(3, 300), (None, 700)] and rows(K.JOIN_NULL_AWARE_LEFT_ANTI) == []
    build = K.Chunk(3, {0: K.Column.from_ints(K.TYPE_INT, np.array([1, 2, 2], np.int32), np.zeros(3, np.uint8)),
                        1: K.Column.from_ints(K.TYPE_BIGINT, np.array([10, 20, 21]))})
    assert rows(K.JOIN_LEFT_ANTI) == [(3, 300), (None, 700)] and rows(K.JOIN_NULL_AWARE_LEFT_ANTI) == [(3, 300)]


def test_string_predicates_of_the_reference_tests(oracle):
    # be/test/storage/column_predicate_test.cpp:564-660: col = 'aa' over {aa, ab, ac} -> 1,0,0; nullable col = 'abc' over
    # {xxx, yyy, abc, NULL, abc} -> 0,0,1,0,1 (a NULL never satisfies a comparison)
    from starrocks_b200.plan import ScanAggPlan, pred
    ids = K.Column.from_ints(K.TYPE_INT, np.arange(5, dtype=np.int32))
    ch = K.Chunk(3, {0: K.Column.from_strings([b"aa", b"ab", b"ac"]), 1: K.Column.from_ints(K.TYPE_INT, np.arange(3, dtype=np.int32))})
    keep = lambda plan, c: [r[0] for r in K.chunk_rows(oracle.scan_agg(plan, c), sort=False)]
    assert keep(ScanAggPlan(preds=[pred(0, K.PRED_EQ, K.TYPE_VARCHAR, b"aa")], out_slots=[1]), ch) == [0]
    ch = K.Chunk(5, {0: K.Column.from_strings([b"xxx", b"yyy", b"abc", b"", b"abc"], nulls=np.array([0, 0, 0, 1, 0], np.uint8)), 1: ids})
    assert keep(ScanAggPlan(preds=[pred(0, K.PRED_EQ, K.TYPE_VARCHAR, b"abc")], out_slots=[1]), ch) == [2, 4]
    assert keep(ScanAggPlan(preds=[pred(0, K.PRED_NE, K.TYPE_VARCHAR, b"abc")], out_slots=[1]), ch) == [0, 1]
    # Slice::compare: memcmp over the common length, then the shorter string is smaller
    ch = K.Chunk(5, {0: K.Column.from_strings([b"", b"ab", b"abc", b"abd", b"b"]), 1: ids})
    assert keep(ScanAggPlan(preds=[pred(0, K.PRED_LT, K.TYPE_VARCHAR, b"abc")], out_slots=[1]), ch) == [0, 1]
    assert keep(ScanAggPlan(preds=[pred(0, K.PRED_GE, K.TYPE_VARCHAR, b"abc")], out_slots=[1]), ch) == [2, 3, 4]


def test_bitset_runtime_filter_of_the_reference_test(oracle):
    # be/test/exprs/runtime_filter_test.cpp:1409-1455 TestEvaluateBitsetFilter: even keys 0..198, min 0 / max 198
    from tests.oracle_lib import OracleBitsetFilter
    keys = K.Column.from_ints(K.TYPE_INT, np.arange(0, 200, 2, dtype=np.int32))
    probe = np.array([-1, 5, 10, 15, 20, 25, 200], np.int32)
    f = OracleBitsetFilter(oracle, 0, 198).insert_column(keys)
    assert f.evaluate(K.Column.from_ints(K.TYPE_INT, probe)).tolist() == [0, 0, 1, 0, 1, 0, 0]
    with_nulls = K.Column.from_ints(K.TYPE_INT, np.concatenate([probe, [0, 0]]).astype(np.int32), np.array([0] * 7 + [1, 1], np.uint8))
    assert f.evaluate(with_nulls).tolist() == [0, 0, 1, 0, 1, 0, 0, 0, 0]
    f.close()
    f = OracleBitsetFilter(oracle, 0, 198).insert_column(K.Column.from_ints(K.TYPE_INT, np.concatenate([np.arange(0, 200, 2), [7]]).astype(np.int32),
                                                                            np.array([0] * 100 + [1], np.uint8)))  # insert_null()
    assert f.evaluate(K.Column.from_ints(K.TYPE_INT, probe)).tolist() == [0, 0, 1, 0, 1, 0, 0]
    assert f.evaluate(with_nulls).tolist() == [0, 0, 1, 0, 1, 0, 0, 1, 1]
    assert len(f.data()) == 25 and f.data()[0] == 0b01010101  # 199 values, eight per byte, bit (x - min) % 8
    f.close()
    # try_create_runtime_bitset_filter (runtime_filter_bank.cpp:160-200): bitset bytes against one byte per build row, the L2 and the L3 cache
    ch = lambda mn, mx, rows, l2, l3: OracleBitsetFilter.chosen(oracle, mn, mx, rows, l2, l3)
    assert ch(0, 799, 100, 0, 0) and not ch(0, 800, 100, 0, 0)            # 100 bytes of bitset for 100 rows; 101 bytes is too many
    assert ch(0, 3199, 100, 1 << 20, 0) and not ch(0, 3200, 100, 1 << 20, 0)  # up to 4x when it fits the L2
    assert not ch(0, 3199, 100, 399, 0) and ch(0, 1599, 100, 0, 1 << 25) and not ch(0, 1600, 100, 0, 1 << 25)  # 2x for the L3
    assert not ch(5, 4, 100, 1 << 20, 1 << 25)                            # empty build side: min > max


def test_inner_join_other_conjuncts_by_hand(oracle):
    # HashJoiner::_process_other_conjunct for INNER joins (hash_joiner.cpp:488-507): the joined rows that satisfy the conjuncts; NULL fails
    from starrocks_b200.plan import JoinPlan
    build = K.Chunk(4, {0: K.Column.from_ints(K.TYPE_INT, np.array([1, 2, 2, 3], np.int32)),
                        1: K.Column.from_ints(K.TYPE_BIGINT, np.array([10, 20, 30, 40]), np.array([0, 0, 0, 1], np.uint8))})
    probe = K.Chunk(4, {2: K.Column.from_ints(K.TYPE_INT, np.array([2, 1, 3, 2], np.int32)), 3: K.Column.from_ints(K.TYPE_BIGINT, np.array([25, 5, 99, 35]))})
    rows = lambda conj: sorted(K.chunk_rows(oracle.join(JoinPlan([0], [2], build_out=[1], probe_out=[2, 3], conjuncts=conj), build, probe), sort=False))
    assert rows([]) == sorted([(2, 25, 20), (2, 25, 30), (1, 5, 10), (3, 99, None), (2, 35, 20), (2, 35, 30)], key=repr) or len(rows([])) == 6
    assert rows([(3, K.PRED_GT, 1)]) == [(2, 25, 20), (2, 35, 20), (2, 35, 30)]          # p.v > b.v; the NULL b.v never passes
    assert rows([(3, K.PRED_GT, 1), (3, K.PRED_NE, 1)]) == [(2, 25, 20), (2, 35, 20), (2, 35, 30)]
    assert rows([(1, K.PRED_GE, 3)]) == [(1, 5, 10), (2, 25, 30)]                           # b.v >= p.v
