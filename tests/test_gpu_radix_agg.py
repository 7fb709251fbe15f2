"""Radix-partitioned aggregation (starrocks_b200/csrc/srb_radix_agg.cuh) of the high-cardinality GROUP BY k -> SUM(v), COUNT(*), phase 1 and the
phase-2 merge, against the oracle's aggregation (oracle/orc_agg.cpp, pinned to the reference's aggregate_test / agg_hash_map_test vectors)
and against the table-based path. The path is opt-in (SRB_RADIX=1): correct, but slower than the table path on a B200 as measured.
Groups are compared order-normalised and bit-exact."""
import numpy as np
import pytest

from starrocks_b200 import capi as K
from starrocks_b200 import queries as Q
from starrocks_b200 import tpch_synth as T

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from starrocks_b200 import ops
    c = ops.Context(device=0)
    yield c
    c.close()


def sorted_cols(ch):
    """the chunk's columns as numpy arrays, rows sorted by the (unique) key in slot 0; NULL values normalised to 0"""
    order = np.argsort(np.asarray(ch.columns[0].data), kind="stable")
    out = []
    for c in ch.columns.values():
        d = np.asarray(c.data)[order]
        nl = np.asarray(c.nulls)[order] if c.nulls is not None else np.zeros(ch.n_rows, np.uint8)
        out += [np.where(nl != 0, 0, d), nl]
    return out


def same_groups(a, b):
    return a.n_rows == b.n_rows and list(a.columns) == list(b.columns) and all(bool((x == y).all()) for x, y in zip(sorted_cols(a), sorted_cols(b)))


def gpu_agg(ctx, plan, chunks, device=True):
    from starrocks_b200 import ops
    op = ops.ScanAggOperator(ctx, plan)
    ptrs = []
    try:
        for ch in chunks:
            if device:
                ch, pp = ctx.device_chunk(ch)
                ptrs += pp
            op.push_chunk(ch)
        op.set_finishing()
        out = op.pull_device()
        got = ctx.fetch_chunk(out) if out is not None else K.Chunk(0)
        return got, op.counters()["kernel_launches"], op.profile()
    finally:
        op.close()
        for p in ptrs:
            ctx.device_free(p)


@pytest.mark.parametrize("n,nk", [(4_500_000, 750_000), (20_000_000, 3_300_000)])
def test_radix_path_matches_the_oracle_and_the_table_path(ctx, oracle, monkeypatch, n, nk):
    monkeypatch.setenv("SRB_RADIX", "1")
    monkeypatch.setenv("SRB_RADIX_MIN_ROWS", str(1 << 22))
    ch = T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=nk)
    want = oracle.scan_agg(Q.highcard_plan(), ch, n_threads=8)
    got, launches, _ = gpu_agg(ctx, Q.highcard_plan(group_hint=nk), [ch])
    assert launches <= 4  # scatter + aggregate (+ key unpacking): no table was built, no emit walk
    assert same_groups(got, want)
    monkeypatch.setenv("SRB_NO_RADIX", "1")
    table, launches_t, _ = gpu_agg(ctx, Q.highcard_plan(group_hint=nk), [ch])
    assert launches_t > launches and same_groups(table, want)


def test_radix_path_with_wide_values_and_negative_keys(ctx, oracle, monkeypatch):
    # keys over the whole BIGINT range, values whose sums wrap around 2^64 (SUM(BIGINT) wraps in the reference: sum.h:58-63, int64 state)
    monkeypatch.setenv("SRB_RADIX", "1")
    monkeypatch.setenv("SRB_RADIX_MIN_ROWS", str(1 << 22))
    rng = np.random.default_rng(91)
    n = 4_300_000
    keys = rng.integers(-2**63, 2**63 - 1, 700_000)
    k = keys[rng.integers(0, len(keys), n)]
    v = rng.integers(-2**62, 2**62, n)
    ch = K.Chunk(n, {Q.HC_K: K.Column.from_ints(K.TYPE_BIGINT, k), Q.HC_V: K.Column.from_ints(K.TYPE_BIGINT, v)})
    want = oracle.scan_agg(Q.highcard_plan(), ch, n_threads=8)
    got, launches, _ = gpu_agg(ctx, Q.highcard_plan(group_hint=700_000), [ch])
    assert launches <= 4 and same_groups(got, want)


def test_buckets_with_more_keys_than_register_groups_take_the_slow_path(ctx, oracle, monkeypatch):
    # all-distinct keys: ~4400 keys per partition = ~8.6 per bucket, so many buckets hold more than the 8 register groups
    monkeypatch.setenv("SRB_RADIX", "1")
    monkeypatch.setenv("SRB_RADIX_MIN_ROWS", str(1 << 22))
    n = 4_500_000
    rng = np.random.default_rng(92)
    ch = K.Chunk(n, {Q.HC_K: K.Column.from_ints(K.TYPE_BIGINT, rng.permutation(n).astype(np.int64) * 7919),
                     Q.HC_V: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(0, 1 << 20, n))})
    want = oracle.scan_agg(Q.highcard_plan(), ch, n_threads=8)
    got, launches, _ = gpu_agg(ctx, Q.highcard_plan(group_hint=n), [ch])
    assert launches <= 4 and got.n_rows == n and same_groups(got, want)


def test_a_hot_key_overflows_its_partition_and_the_table_path_takes_over(ctx, oracle, monkeypatch):
    monkeypatch.setenv("SRB_RADIX", "1")
    monkeypatch.setenv("SRB_RADIX_MIN_ROWS", str(1 << 22))
    n = 4_400_000
    ch = T.host_chunk("highcard", 5, n, ["k", "v"], n_keys=600_000)
    ch.columns[Q.HC_K].data[::3] = 424242  # a third of the rows share one key
    want = oracle.scan_agg(Q.highcard_plan(), ch, n_threads=8)
    got, launches, _ = gpu_agg(ctx, Q.highcard_plan(group_hint=600_000), [ch])
    assert launches > 4 and same_groups(got, want)


def test_a_second_batch_merges_with_the_radix_groups_through_the_table(ctx, oracle, monkeypatch):
    monkeypatch.setenv("SRB_RADIX", "1")
    monkeypatch.setenv("SRB_RADIX_MIN_ROWS", str(1 << 22))
    n1, n2, nk = 4_500_000, 1_200_000, 800_000
    a = T.host_chunk("highcard", 0, n1, ["k", "v"], n_keys=nk)
    b = T.host_chunk("highcard", n1, n2, ["k", "v"], n_keys=nk)
    both = T.host_chunk("highcard", 0, n1 + n2, ["k", "v"], n_keys=nk)
    want = oracle.scan_agg(Q.highcard_plan(), both, n_threads=8)
    got, _, _ = gpu_agg(ctx, Q.highcard_plan(group_hint=nk), [a, b])
    assert same_groups(got, want)
    got, _, _ = gpu_agg(ctx, Q.highcard_plan(group_hint=nk), [a, b, ops_slice(a, 100_000)])
    three = oracle.scan_agg(Q.highcard_plan(), concat([both, ops_slice(a, 100_000)]), n_threads=8)
    assert same_groups(got, three)


def ops_slice(ch, n):
    from starrocks_b200 import ops
    return ops.slice_chunk(ch, 0, n)


def concat(chunks):
    from starrocks_b200 import distributed as D
    return D.concat_chunks(chunks)


def test_merge_phase_through_the_radix_path(ctx, oracle, monkeypatch):
    """phase 2 of the two-phase GROUP BY: (k, partial sum, partial count) tuples of several phase-1 operators (what a rank receives from the
    shuffle) merged by the radix path = one-phase aggregation over all rows"""
    monkeypatch.setenv("SRB_RADIX", "1")
    monkeypatch.setenv("SRB_RADIX_MIN_ROWS", str(1 << 22))
    from starrocks_b200 import ops
    n, nk, pieces = 12_000_000, 2_000_000, 4
    ch = T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=nk)
    want = oracle.scan_agg(Q.highcard_plan(), ch, n_threads=8)
    inters = []
    for i in range(pieces):
        part = ops.slice_chunk(ch, i * n // pieces, (i + 1) * n // pieces)
        inters.append(oracle.scan_agg(Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE), part, n_threads=8))
    tuples = concat(inters)
    assert tuples.n_rows >= 1 << 22
    for c in tuples.columns.values():  # the static merge shape takes null-free tuples (SUM states of non-NULL values)
        assert c.nulls is None or not c.nulls.any()
        c.nulls = None
    got, launches, _ = gpu_agg(ctx, Q.highcard_merge_plan(group_hint=nk), [tuples])
    assert launches <= 4 and same_groups(got, want)
    monkeypatch.setenv("SRB_NO_RADIX", "1")
    table, launches_t, _ = gpu_agg(ctx, Q.highcard_merge_plan(group_hint=nk), [tuples])
    assert launches_t > launches and same_groups(table, want)
