"""srb_partition_run_noalloc (include/srb/srb_c_api.h): the shuffle partitioner for callers that cannot allocate. It has no counterpart in
the reference; its results are held to the existing entry point's (srb_partition_run, itself checked against the oracle's restatement of
the shuffle hash in tests/test_gpu_exchange.py): same counts, and the same rows per destination (any order inside a destination, as there).
The file name sorts last on purpose: a failure here must not hide the parity tests."""
import ctypes as C

import numpy as np
import pytest

from starrocks_b200 import capi as K

pytestmark = pytest.mark.gpu

MAX_ROWS = 1 << 20  # SRB_PARTITION_NOALLOC_MAX_ROWS


@pytest.fixture(scope="module")
def ctx():
    from starrocks_b200 import ops
    c = ops.Context(device=0)
    yield c
    c.close()


def _chunk(rng, n, with_nulls):
    nulls = (rng.random(n) < 0.1).astype(np.uint8) if with_nulls else None
    return K.Chunk(n, {7: K.Column.from_ints(K.TYPE_BIGINT, rng.integers(-2**40, 2**40, n), nulls),
                       3: K.Column.from_ints(K.TYPE_INT, rng.integers(0, 5000, n).astype(np.int32)),
                       9: K.Column.from_ints(K.TYPE_DECIMAL128, [int(x) for x in rng.integers(-2**62, 2**62, n)], None, 27, 4)})


def _by_destination(ctx, grouped_dev, counts):
    from starrocks_b200 import ops
    got = ctx.fetch_chunk(grouped_dev)
    out, pos = [], 0
    for c in counts:
        out.append(sorted(K.chunk_rows(ops.slice_chunk(got, pos, pos + c), sort=False), key=repr))
        pos += c
    return out


def _run_noalloc(ctx, dch, key_slots, world, out_bufs, counts_dev):
    """calls the entry point through the C ABI with caller-owned device buffers; returns (status, device Chunk, counts)"""
    from starrocks_b200 import ops
    L = ops.lib()
    n = dch.n_rows
    out = K.Chunk(n)
    for slot, c in dch.columns.items():
        d, nl = out_bufs[slot]
        out.columns[slot] = K.Column(c.type, d, nl if c.nulls is not None else None, None, c.precision, c.scale, K.MEM_DEVICE)
    cin, keep_in = dch.to_c()
    cout, keep_out = out.to_c()
    ks = (C.c_int32 * len(key_slots))(*key_slots)
    st = L.srb_partition_run_noalloc(0, None, world, C.byref(cin), len(key_slots), ks, C.byref(cout), counts_dev)
    if st != K.OK:
        return st, None, None
    counts = np.zeros(world, np.int64)
    ops._check(L.srb_memcpy_d2h(ctx.h, counts.ctypes.data, counts_dev, world * 8))  # a blocking copy on the default stream: orders behind the kernels
    return st, out, counts.tolist()


@pytest.mark.parametrize("world", [2, 8])
def test_noalloc_partitioner_matches_the_allocating_one_on_two_calls_in_a_row(ctx, world):
    from starrocks_b200 import ops
    rng = np.random.default_rng(1234 + world)
    cap = 300_000
    # the caller's buffers, allocated once up front (the entry point itself allocates nothing)
    out_bufs = {7: (ctx.device_alloc(cap * 8), ctx.device_alloc(cap)), 3: (ctx.device_alloc(cap * 4), ctx.device_alloc(cap)),
                9: (ctx.device_alloc(cap * 16), ctx.device_alloc(cap))}
    counts_dev = ctx.device_alloc(64 * 8)
    part = ops.Partitioner(ctx, world)
    try:
        # two calls in a row share the static scratch: the second (other size, other keys, no null maps, two partition columns)
        # must not see the first one's totals / cursors / destination bytes
        for n, with_nulls, keys in ((cap, True, [7]), (123_457, False, [3, 7]), (1, True, [7]), (0, False, [3])):
            ch = _chunk(rng, n, with_nulls)
            dch, ptrs = ctx.device_chunk(ch)
            try:
                want_dev, want_counts = part.run(dch, keys)
                want = _by_destination(ctx, want_dev, want_counts)
                st, got_dev, got_counts = _run_noalloc(ctx, dch, keys, world, out_bufs, counts_dev)
                assert st == K.OK, ops.lib().srb_last_error()
                assert got_counts == want_counts and sum(got_counts) == n
                assert _by_destination(ctx, got_dev, got_counts) == want
            finally:
                for p in ptrs:
                    ctx.device_free(p)
    finally:
        part.close()
        for d, nl in out_bufs.values():
            ctx.device_free(d)
            ctx.device_free(nl)
        ctx.device_free(counts_dev)


def test_noalloc_partitioner_rejects_chunks_beyond_its_static_scratch(ctx):
    from starrocks_b200 import ops
    n = MAX_ROWS + 1
    col = ctx.device_alloc(n * 4)
    outc = ctx.device_alloc(n * 4)
    counts_dev = ctx.device_alloc(64 * 8)
    try:
        dch = K.Chunk(n, {0: K.Column(K.TYPE_INT, col, None, None, 0, 0, K.MEM_DEVICE)})
        st, _, _ = _run_noalloc(ctx, dch, [0], 4, {0: (outc, None)}, counts_dev)
        assert st == K.INVALID_ARGUMENT
        msg = ops.lib().srb_last_error().decode()
        assert "SRB_PARTITION_NOALLOC_MAX_ROWS" in msg and str(MAX_ROWS) in msg
    finally:
        for p in (col, outc, counts_dev):
            ctx.device_free(p)
