"""Streaming pre-aggregation sink — the oracle's restatement of AggregateStreamingSinkOperator
(be/src/exec/pipeline/aggregate/aggregate_streaming_sink_operator.cpp:73-307) and of AggrAutoContext (be/src/exec/aggregator.h:172-193).
CPU only. The reference holds no known-answer test for the state machine ("parity unpinned" for which rows are passed through);
what IS observable — and checked here — is (a) the intermediate row a passed-through input row becomes
(convert_to_serialize_format), (b) that phase 2 over the sink's output gives the one-phase result whatever the mode did, and
(c) the transitions the constants imply on inputs with very low and very high reduction."""
import numpy as np
import pytest

from starrocks_b200 import capi as K
from starrocks_b200 import queries as Q
from starrocks_b200.plan import ExprBuilder, ScanAggPlan, pred

AUTO, FORCE_STREAMING, FORCE_PREAGGREGATION = 0, 1, 2  # TStreamingPreaggregationMode, gensrc/thrift/PlanNodes.thrift:689-694


def hc_chunk(keys, vals, vnulls=None):
    return K.Chunk(len(keys), {Q.HC_K: K.Column.from_ints(K.TYPE_BIGINT, np.asarray(keys, np.int64)),
                               Q.HC_V: K.Column.from_ints(K.TYPE_BIGINT, np.asarray(vals, np.int64), vnulls)})


def one_phase(oracle, ch):
    return K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), ch))


def merged(oracle, inter):
    return K.chunk_rows(oracle.scan_agg(Q.highcard_merge_plan(), inter))


def test_force_streaming_emits_one_intermediate_row_per_input_row(oracle):
    rng = np.random.default_rng(1)
    n = 10_000
    keys, vals = rng.integers(0, 50, n), rng.integers(-1000, 1000, n)
    vnulls = (rng.random(n) < 0.1).astype(np.uint8)
    ch = hc_chunk(keys, vals, vnulls)
    inter, st = oracle.scan_agg_streaming(Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE), ch, FORCE_STREAMING)
    assert inter.n_rows == n and st["rows_streamed"] == n and st["groups"] == 0
    rows = K.chunk_rows(inter, sort=False)
    # sum.h / count.h convert_to_serialize_format: the value itself (NULL stays NULL), count(*) = 1; arrival order is kept
    assert rows == [(int(k), None if nl else int(v), 1) for k, v, nl in zip(keys, vals, vnulls)]
    assert merged(oracle, inter) == one_phase(oracle, ch)


def test_force_preaggregation_is_the_blocking_first_phase(oracle):
    rng = np.random.default_rng(2)
    ch = hc_chunk(rng.integers(0, 300, 20_000), rng.integers(0, 100, 20_000))
    plan = Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE)
    inter, st = oracle.scan_agg_streaming(plan, ch, FORCE_PREAGGREGATION)
    assert st["rows_streamed"] == 0 and st["groups"] == inter.n_rows == 300
    assert K.chunk_rows(inter) == K.chunk_rows(oracle.scan_agg(plan, ch))


def test_auto_keeps_aggregating_when_the_reduction_is_high(oracle):
    rng = np.random.default_rng(3)
    n = 400_000  # ~98 chunks of 4096 rows over 2000 keys: the table never needs to expand after the first chunks
    ch = hc_chunk(rng.integers(0, 2000, n), rng.integers(0, 100, n))
    inter, st = oracle.scan_agg_streaming(Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE), ch, AUTO)
    assert st["rows_streamed"] == 0 and inter.n_rows == 2000 and st["final_state"] == "init_preagg"
    assert merged(oracle, inter) == one_phase(oracle, ch)


def test_auto_passes_through_when_nothing_reduces(oracle):
    n = 4096 * 400  # all keys distinct: reduction 1.0
    ch = hc_chunk(np.arange(n), np.ones(n, np.int64))
    inter, st = oracle.scan_agg_streaming(Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE), ch, AUTO)
    # The trajectory the reference's constants give (400 chunks of 4096 distinct keys):
    #   chunk 0 INIT_PREAGG (empty table: always aggregate); chunk 1 INIT_PREAGG: the table must grow, it holds >= 256 KB, so
    #   STREAMING_HT_MIN_REDUCTION asks for a reduction > 1.1 and sees 1.0 -> ADJUST (same chunk); 5 chunks of 0 hits (StableLimit)
    #   -> PASS_THROUGH until pass_through_count > continuous_limit (100) = 96 chunks -> FORCE_PREAGG for ForcePreaggLimit + 1 = 4
    #   chunks -> ADJUST 5 -> PASS_THROUGH with the doubled limit (196 chunks) -> FORCE_PREAGG 4 -> ADJUST 5 -> PASS_THROUGH 84.
    c = st["chunks"]
    assert (c["init_preagg"], c["adjust"], c["pass_through"], c["force_preagg"], c["preagg"], c["selective_preagg"]) == (2, 14, 376, 8, 0, 0)
    assert st["final_state"] == "pass_through"
    assert st["rows_streamed"] > n // 2 and st["rows_streamed"] + st["groups"] == n == inter.n_rows
    assert merged(oracle, inter) == one_phase(oracle, ch)


@pytest.mark.parametrize("mode", [AUTO, FORCE_STREAMING, FORCE_PREAGGREGATION])
def test_phase_two_result_does_not_depend_on_the_mode(oracle, mode):
    rng = np.random.default_rng(4)
    # three regimes back to back: heavy repetition, all distinct, moderate repetition — AUTO visits several states
    keys = np.concatenate([rng.integers(0, 100, 100_000), 10**6 + np.arange(300_000), rng.integers(0, 150_000, 200_000)])
    vals = rng.integers(-10**12, 10**12, len(keys))
    vnulls = (rng.random(len(keys)) < 0.05).astype(np.uint8)
    ch = hc_chunk(keys, vals, vnulls)
    inter, st = oracle.scan_agg_streaming(Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE), ch, mode)
    assert st["rows_streamed"] + st["groups"] == inter.n_rows
    if mode == AUTO:
        assert 0 < st["rows_streamed"] < len(keys) and st["chunks"]["adjust"] > 0
    assert merged(oracle, inter) == one_phase(oracle, ch)


def test_streaming_with_predicates_avg_and_decimal_keys(oracle):
    """every aggregate kind the sink can pass through: SUM / AVG (intermediate blob {sum, count}) / COUNT(col) / MIN / MAX over a
    DECIMAL64 column with NULLs, two key columns, a scan predicate in front"""
    rng = np.random.default_rng(5)
    n = 60_000
    nulls = (rng.random(n) < 0.1).astype(np.uint8)
    ch = K.Chunk(n, {0: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 40, n).astype(np.int32)),
                     1: K.Column.from_ints(K.TYPE_DATE, rng.integers(2450000, 2450030, n).astype(np.int32)),
                     2: K.Column.from_ints(K.TYPE_DECIMAL64, rng.integers(-10**9, 10**9, n), nulls, 15, 2),
                     3: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 10, n).astype(np.int32))})

    def plan(phase, slots=(0, 1, 2), merge=False):
        e = ExprBuilder()
        if merge:
            k0, k1 = e.slot(0, K.TYPE_INT), e.slot(1, K.TYPE_DATE)
            args = [e.slot(2, K.TYPE_DECIMAL128, 38, 2), e.slot(3, K.TYPE_VARBINARY, K.TYPE_DECIMAL64, 2), e.slot(4, K.TYPE_BIGINT),
                    e.slot(5, K.TYPE_DECIMAL64, 15, 2), e.slot(6, K.TYPE_DECIMAL64, 15, 2)]
            return ScanAggPlan(exprs=e, keys=[k0, k1], aggs=list(zip([K.AGG_SUM, K.AGG_AVG, K.AGG_COUNT, K.AGG_MIN, K.AGG_MAX], args)), phase=phase)
        k0, k1, m = e.slot(0, K.TYPE_INT), e.slot(1, K.TYPE_DATE), e.slot(2, K.TYPE_DECIMAL64, 15, 2)
        return ScanAggPlan(preds=[pred(3, K.PRED_LT, K.TYPE_INT, 7)], exprs=e, keys=[k0, k1],
                           aggs=[(K.AGG_SUM, m), (K.AGG_AVG, m), (K.AGG_COUNT, m), (K.AGG_MIN, m), (K.AGG_MAX, m)], phase=phase)

    want = K.chunk_rows(oracle.scan_agg(plan(K.AGG_UPDATE_FINALIZE), ch))
    for mode in (AUTO, FORCE_STREAMING, FORCE_PREAGGREGATION):
        inter, st = oracle.scan_agg_streaming(plan(K.AGG_UPDATE_SERIALIZE), ch, mode, chunk_size=1000)
        got = K.chunk_rows(oracle.scan_agg(plan(K.AGG_MERGE_FINALIZE, merge=True), inter))
        assert got == want, mode


def test_device_auto_policy_replayed_on_the_cpu(tmp_path):
    """the device sink's AUTO policy is host-only code (starrocks_b200/csrc/srb_stream_policy.h); tests/cpp/test_stream_policy.cpp
    replays it against a model table. All-distinct keys: one aggregated unit, StableLimit selective units, then pass-through with
    a short FORCE_PREAGG / ADJUST excursion every continuous_limit units — the trajectory the oracle's restatement of the reference
    shows on the same input (test_auto_passes_through_when_nothing_reduces). A small key domain: everything is aggregated."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "test_stream_policy")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-o", exe, os.path.join(root, "tests", "cpp", "test_stream_policy.cpp")])

    def run(*args):
        return [int(x) for x in subprocess.check_output([exe] + [str(a) for a in args]).split()]

    # units, rows per unit, distinct, "L2" bytes -> units aggregated / passed through / selective, rows passed, groups, final state
    assert run(400, 4096, 1, 256 * 1024) == [9, 376, 15, 391 * 4096, 9 * 4096, 2]  # 1 INIT + 2 x 4 FORCE_PREAGG; 3 x 5 ADJUST; PASS_THROUGH
    assert run(100, 4096, 0, 256 * 1024) == [100, 0, 0, 0, 2000, 0]
    # B200 scale: 600 M all-distinct rows in 2^24-row units against the 96 MB threshold: one unit builds a table, 587 M rows pass through
    assert run(36, 1 << 24, 1, 96 << 20) == [1, 30, 5, 35 << 24, 1 << 24, 2]
