"""world_size-2 tests of the multi-rank host logic (starrocks_b200/distributed.py) on CPU: torch.distributed with the gloo
backend moves the intermediate rows, the oracle plays the local engine (the product has no CPU engine — this checks the
sharding / two-phase / exchange plumbing that bench.py runs around the GPU operators)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from starrocks_b200 import capi as K, distributed as D, queries as Q, tpch_synth as T  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_range_covers_every_row_once():
    for n in (0, 1, 7, 600_037_902):
        for world in (1, 2, 3, 8):
            prev = 0
            for r in range(world):
                b, e = D.shard_range(n, r, world)
                assert b == prev and e >= b
                prev = e
            assert prev == n
            sizes = [D.shard_range(n, r, world)[1] - D.shard_range(n, r, world)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        D.shard_range(10, 2, 2)


def test_group_by_destination_is_a_stable_counting_sort():
    ch = K.Chunk(6, {0: K.Column.from_ints(K.TYPE_BIGINT, [10, 11, 12, 13, 14, 15]),
                     1: K.Column.from_strings([b"a", b"bb", b"", b"dddd", b"e", b"ff"], nulls=[0, 0, 1, 0, 0, 0])})
    g, counts = D.group_by_destination(ch, np.array([1, 0, 1, 0, 2, 1]), 3)
    assert counts == [2, 3, 1]
    assert K.column_pyvalues(g.columns[0]) == [11, 13, 10, 12, 15, 14]
    assert K.column_pyvalues(g.columns[1]) == [b"bb", b"dddd", b"a", None, b"ff", b"e"]
    back = D.concat_chunks([g, g])
    assert back.n_rows == 12 and K.column_pyvalues(back.columns[1])[6:] == K.column_pyvalues(g.columns[1])
    with pytest.raises(ValueError):
        D.group_by_destination(ch, np.array([0, 0, 0, 0, 0, 3]), 3)


def _shuffle_dest(oracle, col: K.Column, world: int) -> np.ndarray:
    """destination of every row by the reference's shuffle hash of the key value (exchange/shuffler.h:70-85)"""
    w = K.type_width(col.type)
    raw = np.ascontiguousarray(col.data).view(np.uint8).reshape(-1, w)
    return np.array([oracle.lib.orc_shuffle_channel(raw[i].tobytes(), w, world) for i in range(raw.shape[0])], dtype=np.int64)


def _worker(rank, world, port, case, n_rows, out_q):
    import torch.distributed as dist
    from oracle_lib import Oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        oracle = Oracle()
        ex = D.ObjectExchange(dist, rank, world)
        b, e = D.shard_range(n_rows, rank, world)
        if case == "q1":
            shard = T.host_chunk("lineitem", b, e - b, T.LINEITEM_Q1)
            got = D.two_phase_gather(lambda ch: oracle.scan_agg(Q.q1_plan(phase=K.AGG_UPDATE_SERIALIZE), ch),
                                     lambda inter: oracle.scan_agg(Q.q1_merge_plan(inter), inter), ex, shard)
        else:
            shard = T.host_chunk("highcard", b, e - b, ["k", "v"], n_keys=5000)
            got = D.two_phase_shuffle(lambda ch: oracle.scan_agg(Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE), ch),
                                      lambda inter: D.group_by_destination(inter, _shuffle_dest(oracle, inter.columns[0], world), world),
                                      lambda recv: oracle.scan_agg(Q.highcard_merge_plan(), recv), ex, shard)
        out_q.put((rank, K.chunk_rows(got)))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def _run(case, n_rows, world=2):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, n_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = {}
    for _ in range(world):
        r, rows = q.get(timeout=240)
        res[r] = rows
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return res


def test_two_rank_q1_gather_merge_equals_one_rank(oracle):
    n = 40_000
    res = _run("q1", n)
    whole = T.host_chunk("lineitem", 0, n, T.LINEITEM_Q1)
    want = K.chunk_rows(oracle.scan_agg(Q.q1_plan(), whole))
    assert res[0] == want and res[1] == want  # every rank holds the merged result
    assert len(want) == 4


def test_two_rank_highcard_shuffle_merge_equals_one_rank(oracle):
    n = 30_000
    res = _run("highcard", n)
    whole = T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=5000)
    want = K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), whole))
    keys0 = {r[0] for r in res[0]}
    keys1 = {r[0] for r in res[1]}
    assert not (keys0 & keys1)  # a group ends on exactly one rank
    assert len(keys0) > 1000 and len(keys1) > 1000  # and the shuffle hash spreads them
    assert sorted(res[0] + res[1]) == want


def test_four_ranks_gather_and_shuffle(oracle):
    """the same host logic at world size 4 (the scaling run goes to 8 ranks): ragged shards, every rank ends with the merged Q1 result;
    the shuffle leaves every group on exactly one of the four ranks"""
    n = 40_003
    res = _run("q1", n, world=4)
    want = K.chunk_rows(oracle.scan_agg(Q.q1_plan(), T.host_chunk("lineitem", 0, n, T.LINEITEM_Q1)))
    assert all(res[r] == want for r in range(4))
    n = 30_001
    res = _run("highcard", n, world=4)
    want = K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=5000)))
    key_sets = [{row[0] for row in res[r]} for r in range(4)]
    assert sum(len(s) for s in key_sets) == len(set().union(*key_sets)) == len(want)
    assert all(len(s) > 500 for s in key_sets)
    assert sorted(sum((res[r] for r in range(4)), [])) == want
