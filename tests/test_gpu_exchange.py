"""GPU tests of the exchange step (SURVEY §8 a12 / e): device hash partitioning against the oracle's shuffle hash, the NCCL
all-to-all / all-gather of libsrb, and — when two GPUs are visible — the whole two-phase high-cardinality aggregation."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from starrocks_b200 import capi as K, distributed as D, queries as Q, tpch_synth as T  # noqa: E402
from tests import oracle_chains as OC  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    from starrocks_b200 import ops
    c = ops.Context(device=0)
    yield c
    c.close()


def _dest(oracle, col, world):
    w = K.type_width(col.type)
    raw = np.ascontiguousarray(col.data).view(np.uint8).reshape(-1, w)
    return np.array([oracle.lib.orc_shuffle_channel(raw[i].tobytes(), w, world) for i in range(raw.shape[0])], dtype=np.int64)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_partitioner_routes_rows_like_the_shuffle_hash(ctx, oracle, world):
    from starrocks_b200 import ops
    n = 50_000
    ch = T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=7000)
    dch, ptrs = ctx.device_chunk(ch)
    part = ops.Partitioner(ctx, world)
    try:
        grouped, counts = part.run(dch, [Q.HC_K])
        got = ctx.fetch_chunk(grouped)
        dest = _dest(oracle, ch.columns[Q.HC_K], world)
        assert counts == np.bincount(dest, minlength=world).tolist() and sum(counts) == n
        pos = 0
        want_rows = K.chunk_rows(ch, sort=False)
        for d in range(world):  # destination d's rows are contiguous and are exactly the rows that hash to d (any order inside)
            piece = sorted(K.chunk_rows(ops.slice_chunk(got, pos, pos + counts[d]), sort=False))
            assert piece == sorted(r for r, dd in zip(want_rows, dest) if dd == d)
            pos += counts[d]
    finally:
        part.close()
        for p in ptrs:
            ctx.device_free(p)


def test_set_finishing_merged_on_a_single_rank_communicator(ctx, oracle):
    """the one-call two-phase aggregation with world size 1: the gather returns the rank's own table, nothing is merged in, the result is
    the local result (MIN / MAX / AVG / DOUBLE states included)"""
    from starrocks_b200 import ops
    from starrocks_b200.plan import ExprBuilder, ScanAggPlan
    box = {}
    comm = ops.Communicator(ctx, lambda b: box.__setitem__("id", b), lambda: box["id"])
    try:
        ch = T.host_chunk("lineitem", 3, 300_000, T.LINEITEM_Q1)
        for plan in (Q.q1_plan(), Q.q6_plan()):
            op = ops.ScanAggOperator(ctx, plan)
            op.push_chunk(ch)
            op.set_finishing_merged(comm)
            assert K.chunk_rows(op.pull_all()) == K.chunk_rows(oracle.scan_agg(plan, ch))
            op.close()
        e = ExprBuilder()
        k = e.slot(Q.L_RETURNFLAG_CODE, K.TYPE_INT)
        qty = e.slot(Q.L_QUANTITY, K.TYPE_DECIMAL64, 15, 2)
        plan = ScanAggPlan(exprs=e, keys=[k], aggs=[(K.AGG_MIN, qty), (K.AGG_MAX, qty), (K.AGG_AVG, qty), (K.AGG_COUNT, None)], group_hint=8)
        op = ops.ScanAggOperator(ctx, plan)
        op.push_chunk(ch)
        op.set_finishing_merged(comm)
        assert K.chunk_rows(op.pull_all()) == K.chunk_rows(oracle.scan_agg(plan, ch))
        op.close()
        with pytest.raises(ops.SrbError):  # a serialize-phase plan does not finalise
            op = ops.ScanAggOperator(ctx, Q.q1_plan(phase=K.AGG_UPDATE_SERIALIZE))
            try:
                op.push_chunk(ch)
                op.set_finishing_merged(comm)
            finally:
                op.close()
    finally:
        comm.close()


def test_comm_abort_cancels_the_communicator(ctx):
    """srb_comm_abort (ncclCommAbort): afterwards every collective on the communicator fails with CANCELLED instead of blocking"""
    from starrocks_b200 import ops
    box = {}
    comm = ops.Communicator(ctx, lambda b: box.__setitem__("id", b), lambda: box["id"])
    ch = T.host_chunk("highcard", 0, 1000, ["k", "v"], n_keys=10)
    dch, ptrs = ctx.device_chunk(ch)
    try:
        comm.check()
        assert comm.all_gather(dch).n_rows == 1000
        comm.abort()
        comm.abort()  # idempotent
        for call in (lambda: comm.all_gather(dch), lambda: comm.shuffle(dch, [Q.HC_K]), lambda: comm.all_to_all(dch, [1000]), comm.check):
            with pytest.raises(ops.SrbError) as ei:
                call()
            assert ei.value.code == K.CANCELLED
    finally:
        comm.close()
        for p in ptrs:
            ctx.device_free(p)


def test_single_rank_communicator_round_trip(ctx):
    from starrocks_b200 import ops
    box = {}
    comm = ops.Communicator(ctx, lambda b: box.__setitem__("id", b), lambda: box["id"])
    ch = T.host_chunk("highcard", 0, 10_000, ["k", "v"], n_keys=100)
    dch, ptrs = ctx.device_chunk(ch)
    try:
        recv, rc = comm.all_to_all(dch, [ch.n_rows])
        assert rc == [ch.n_rows]
        assert K.chunk_rows(ctx.fetch_chunk(recv), sort=False) == K.chunk_rows(ch, sort=False)
        allg = comm.all_gather(dch)
        assert K.chunk_rows(ctx.fetch_chunk(allg), sort=False) == K.chunk_rows(ch, sort=False)
        # the fused shuffle on one rank: every row routes to rank 0 and lands in the receive arena (any order)
        for rows in (10_000, 3, 10_000):  # the arena is reused, then regrown
            part_in = ops.slice_chunk(ch, 0, rows)
            d2, p2 = ctx.device_chunk(part_in)
            got, rc2 = comm.shuffle(d2, [Q.HC_K])
            assert rc2 == [rows] and K.chunk_rows(ctx.fetch_chunk(got)) == K.chunk_rows(part_in)
            for p in p2:
                ctx.device_free(p)
        big = T.host_chunk("lineitem", 0, 300_000, ["l_orderkey", "l_extendedprice", "l_shipdate"])
        d3, p3 = ctx.device_chunk(big)
        got, rc3 = comm.shuffle(d3, [Q.L_ORDERKEY])
        assert rc3 == [big.n_rows] and K.chunk_rows(ctx.fetch_chunk(got)) == K.chunk_rows(big) and comm.launches() >= 8
        for p in p3:
            ctx.device_free(p)
        # host chunks of any column kind (the intermediate rows of a low-cardinality plan: DECIMAL128 sums, {sum,count} blobs, nullable)
        li = T.host_chunk("lineitem", 0, 50_000, T.LINEITEM_Q1)
        op = ops.ScanAggOperator(ctx, Q.q1_plan(phase=K.AGG_UPDATE_SERIALIZE))
        op.push_chunk(li)
        op.set_finishing()
        inter = op.pull_all()
        op.close()
        assert any(K.is_binary(c.type) for c in inter.columns.values())
        got = comm.all_gather(inter)
        assert K.chunk_rows(got, sort=False) == K.chunk_rows(inter, sort=False)
        empty = ops.slice_chunk(inter, 0, 0)
        assert comm.all_gather(empty).n_rows == 0
        big = D.concat_chunks([inter] * 400)  # > the 16 KB fixed slot: the exactly sized second collective
        assert K.chunk_rows(comm.all_gather(big), sort=False) == K.chunk_rows(big, sort=False)
    finally:
        comm.close()
        for p in ptrs:
            ctx.device_free(p)


Q3_COLS = {"customer": ["c_custkey", "c_mktsegment_code"], "orders": ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"],
           "lineitem": ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"]}


def _q3_tables(n_c, n_o, n_l, rank=0, world=1):
    """this rank's row ranges of the three tables (range-partitioned scans)"""
    out = []
    for name, n in (("customer", n_c), ("orders", n_o), ("lineitem", n_l)):
        b, e = D.shard_range(n, rank, world)
        out.append(T.host_chunk(name, b, e - b, Q3_COLS[name], **({"n_customers": n_c} if name == "orders" else {})))
    return out


def test_q3_shuffle_plan_on_a_single_rank_communicator(ctx, oracle):
    """the multi-GPU Q3 plan (broadcast of the customer build side, hash partition + all-to-all of both sides of the orders
    join) through a world-size-1 communicator: same code path as N > 1, result equal to the oracle chain"""
    from starrocks_b200 import ops, pipelines as P
    cust, orders, line = _q3_tables(3000, 30000, 120000)
    want = K.chunk_rows(OC.q3_oracle(oracle, cust, orders, line))
    assert len(want) > 50
    box = {}
    comm = ops.Communicator(ctx, lambda b: box.__setitem__("id", b), lambda: box["id"])
    part = ops.Partitioner(ctx, 1)
    q = P.TpchQ3(ctx, comm, part)
    rows = {}
    try:
        agg = q.run(cust, orders, line, rows=rows)
        got = K.chunk_rows(agg.pull_all())
        assert q.launches() > 0 and rows["j2_out"] >= len(want)
    finally:
        q.close()
        part.close()
        comm.close()
    assert got == want
    q = P.TpchQ3(ctx)  # the local plan gives the same rows
    try:
        assert K.chunk_rows(q.run(cust, orders, line).pull_all()) == want
    finally:
        q.close()


# ---- two GPUs: phase 1 -> device partition -> NCCL all-to-all -> phase 2 ------------------------------------------------------
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_rows, n_keys, out_q):
    import torch
    import torch.distributed as dist
    from starrocks_b200 import ops
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)  # only carries the NCCL id; the data moves through libsrb
    ctx = ops.Context(device=rank, rank=rank, world_size=world)
    try:
        def put(b):
            dist.broadcast_object_list([b], src=0)

        def get():
            box = [None]
            dist.broadcast_object_list(box, src=0)
            return box[0]

        comm = ops.Communicator(ctx, put, get)
        part = ops.Partitioner(ctx, world)
        b, e = D.shard_range(n_rows, rank, world)
        shard = T.host_chunk("highcard", b, e - b, ["k", "v"], n_keys=n_keys)
        dch, ptrs = ctx.device_chunk(shard)
        op1 = ops.ScanAggOperator(ctx, Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE))
        op1.push_chunk(dch)
        op1.set_finishing()
        inter = op1.pull_device()
        grouped, counts = part.run(inter, [0])
        recv, rcounts = comm.all_to_all(grouped, counts)
        op2 = ops.ScanAggOperator(ctx, Q.highcard_merge_plan())
        op2.push_chunk(recv)
        op2.set_finishing()
        rows = K.chunk_rows(op2.pull_all())
        # the fused shuffle (scatter into the peers' receive arenas) delivers the same rows from the same sources
        recv_nccl = K.chunk_rows(ctx.fetch_chunk(recv))
        for _ in range(2):
            recv2, rcounts2 = comm.shuffle(inter, [0])
            assert rcounts2 == rcounts and K.chunk_rows(ctx.fetch_chunk(recv2)) == recv_nccl
        # low-cardinality plan: phase 1 on the rank's lineitem rows, host gather of the <= 4 intermediate rows, phase 2 everywhere
        lb, le = D.shard_range(n_rows, rank, world)
        li = T.host_chunk("lineitem", lb, le - lb, T.LINEITEM_Q1)
        if rank == 1:
            li = ops.slice_chunk(li, 0, 0)  # a rank without rows still takes part in the gather
        op3 = ops.ScanAggOperator(ctx, Q.q1_plan(phase=K.AGG_UPDATE_SERIALIZE))
        op3.push_chunk(li)
        op3.set_finishing()
        inter3 = op3.pull_all()
        everyone = comm.all_gather(inter3)
        op4 = ops.ScanAggOperator(ctx, Q.q1_merge_plan(inter3))
        op4.push_chunk(everyone)
        op4.set_finishing()
        q1_rows = K.chunk_rows(op4.pull_all())
        # the same two-phase plan as ONE call: the ranks' group tables are gathered and merged on the device (srb_op_set_finishing_merged);
        # a rank without rows takes part with an empty table, and a no-GROUP-BY plan merges its single state block
        merged = []
        for plan, cols in ((Q.q1_plan(), T.LINEITEM_Q1), (Q.q6_plan(), T.LINEITEM_Q6)):
            piece = T.host_chunk("lineitem", lb, le - lb, cols)
            if rank == 1 and cols is T.LINEITEM_Q1:
                piece = ops.slice_chunk(piece, 0, 0)
            op5 = ops.ScanAggOperator(ctx, plan)
            op5.push_chunk(piece)
            op5.set_finishing_merged(comm)
            merged.append(K.chunk_rows(op5.pull_all()))
            op5.close()
        out_q.put((rank, (rows, q1_rows, merged)))
        for o in (op1, op2, op3, op4):
            o.close()
        part.close()
        comm.close()
        for p in ptrs:
            ctx.device_free(p)
        dist.barrier()
    finally:
        ctx.close()
        dist.destroy_process_group()


def test_two_gpu_highcard_two_phase(oracle):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    n, nk, world = 400_000, 50_000, 2
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_worker, args=(r, world, port, n, nk, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    whole = T.host_chunk("highcard", 0, n, ["k", "v"], n_keys=nk)
    want = K.chunk_rows(oracle.scan_agg(Q.highcard_plan(), whole))
    hc = {r: res[r][0] for r in res}
    assert not ({r[0] for r in hc[0]} & {r[0] for r in hc[1]})
    assert sorted(hc[0] + hc[1]) == want
    b0, e0 = D.shard_range(n, 0, world)
    want_q1 = K.chunk_rows(oracle.scan_agg(Q.q1_plan(), T.host_chunk("lineitem", b0, e0 - b0, T.LINEITEM_Q1)))
    assert res[0][1] == want_q1 and res[1][1] == want_q1
    want_q6 = K.chunk_rows(oracle.scan_agg(Q.q6_plan(), T.host_chunk("lineitem", 0, n, T.LINEITEM_Q6)))
    for r in range(world):
        assert res[r][2][0] == want_q1 and res[r][2][1] == want_q6


def _join_worker(rank, world, port, sizes, out_q):
    """Q3 (shuffle join) and SSB Q4.1 (broadcast join) on this rank's row ranges"""
    import torch
    import torch.distributed as dist
    from starrocks_b200 import ops, pipelines as P
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ctx = ops.Context(device=rank, rank=rank, world_size=world)
    try:
        def put(b):
            dist.broadcast_object_list([b], src=0)

        def get():
            box = [None]
            dist.broadcast_object_list(box, src=0)
            return box[0]

        comm = ops.Communicator(ctx, put, get)
        part = ops.Partitioner(ctx, world)
        cust, orders, line = _q3_tables(*sizes["q3"], rank=rank, world=world)
        q3 = P.TpchQ3(ctx, comm, part)
        q3_rows = K.chunk_rows(q3.run(cust, orders, line).pull_all())
        q3.close()
        # SSB Q4.1: every rank filters its slice of each dimension, the slices are all-gathered (broadcast join), lineorder is
        # range-partitioned; phase 1 per rank, the <= 35 intermediate rows are gathered and merged on every rank
        n_lo, n_c, n_s, n_p = sizes["ssb"]
        dims = {}
        for name, n in (("ssb_customer", n_c), ("ssb_supplier", n_s), ("ssb_part", n_p), ("ssb_dates", T.SSB_DAYS)):
            b, e = D.shard_range(n, rank, world)
            dims[name] = T.host_chunk(name, b, e - b, P.SSB_DIM_COLS[name])
        b, e = D.shard_range(n_lo, rank, world)
        lo = T.host_chunk("lineorder", b, e - b, T.LINEORDER_Q41, n_customers=n_c, n_suppliers=n_s, n_parts=n_p)
        ssb = P.SsbQ41(ctx, replicate=comm.all_gather).build(dims)
        agg = ssb.probe(lo, phase=K.AGG_UPDATE_SERIALIZE)
        inter = agg.pull_all()
        agg.close()
        everyone = comm.all_gather(inter)
        op2 = ops.ScanAggOperator(ctx, Q.ssb_q41_merge_plan())
        op2.push_chunk(everyone)
        op2.set_finishing()
        ssb_rows = K.chunk_rows(op2.pull_all())
        op2.close()
        ssb.close()
        # global runtime filter: every rank builds a partial filter over its slice of the keys, sized for all of them; after the
        # merge every rank holds the OR (SimdBlockFilter::merge) and the combined min / max / has_null
        from starrocks_b200.plan import JoinPlan
        all_keys = sizes["rf_keys"]
        b, e = D.shard_range(len(all_keys), rank, world)
        nulls = np.zeros(e - b, np.uint8)
        nulls[0] = rank  # only rank 1 inserts a NULL
        j = ops.HashJoinOperator(ctx, JoinPlan([0], [1]))
        j.push_chunk(K.Chunk(e - b, {0: K.Column.from_ints(K.TYPE_INT, all_keys[b:e], nulls=nulls)}))
        j.set_finishing()
        rf = comm.merge_runtime_filter(j.build_runtime_filter(0, len(all_keys)))
        rf_state = (rf.info(), rf.directory())
        rf.close()
        j.close()
        out_q.put((rank, (q3_rows, ssb_rows, rf_state)))
        part.close()
        comm.close()
        dist.barrier()
    finally:
        ctx.close()
        dist.destroy_process_group()


def test_two_gpu_shuffle_join_and_broadcast_join(oracle):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    from starrocks_b200 import pipelines as P
    world = 2
    sizes = {"q3": (6000, 60000, 240000), "ssb": (300_000, 30_000, 2_000, 200_000),
             "rf_keys": np.random.default_rng(3).integers(-(1 << 28), 1 << 28, 50_001).astype(np.int32)}
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = _free_port()
    procs = [mpc.Process(target=_join_worker, args=(r, world, port, sizes, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=300) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cust, orders, line = _q3_tables(*sizes["q3"])
    want = K.chunk_rows(OC.q3_oracle(oracle, cust, orders, line))
    keys = [{r[:3] for r in res[k][0]} for k in range(world)]
    assert not (keys[0] & keys[1])  # the order key decides the rank: groups are disjoint
    assert sorted(res[0][0] + res[1][0]) == want
    n_lo, n_c, n_s, n_p = sizes["ssb"]
    kw = dict(n_customers=n_c, n_suppliers=n_s, n_parts=n_p)
    dims = {name: T.host_chunk(name, 0, n, P.SSB_DIM_COLS[name]) for name, n in
            (("ssb_customer", n_c), ("ssb_supplier", n_s), ("ssb_part", n_p), ("ssb_dates", T.SSB_DAYS))}
    want_ssb = K.chunk_rows(OC.ssb_q41_oracle(oracle, dims, T.host_chunk("lineorder", 0, n_lo, T.LINEORDER_Q41, **kw)))
    assert res[0][1] == want_ssb and res[1][1] == want_ssb
    from tests.oracle_lib import OracleRuntimeFilter
    keys = sizes["rf_keys"]
    nulls = np.zeros(len(keys), np.uint8)
    nulls[D.shard_range(len(keys), 1, world)[0]] = 1
    want_rf = OracleRuntimeFilter(oracle, len(keys)).insert_column(K.Column.from_ints(K.TYPE_INT, keys, nulls=nulls))
    for r in range(world):
        info, directory = res[r][2]
        assert info == want_rf.info() and info["has_null"] and (directory == want_rf.directory()).all()
    want_rf.close()
