#!/usr/bin/env python
"""bench.py — the BE pipeline engine's scan -> filter -> hash-aggregate / hash-join path through libsrb.so on N B200s, one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--query q1|q6|highcard|q3|ssb] [--scaling weak|strong]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

The line's headline (`metric`, `value`, `roofline`, `e2e`, `cpu_baseline`) is TPC-H SF100 Q1 (BASELINE.json configs[1]): a step = one pass of
the hot path over the rank's resident lineitem shard (600,037,902 rows of the seven columns Q1 reads, 44 B/row with dictionary-coded keys):
operator create -> push the device-resident range -> set_finishing -> pull the 4 result rows; at N > 1 the step includes the gather of the
ranks' <= 4 intermediate rows and the phase-2 merge on every rank. `value` = rows of all ranks / time with inputs already in HBM; `e2e` = the
same pipeline fed from pinned HOST buffers through the same C ABI; `roofline` = algorithmic bytes of the dominant kernel / its CUDA-event
duration vs the measured HBM copy peak; `cpu_baseline` / `--impl reference` = the CPU restatement of the reference's pipeline (oracle/; the
real BE cannot be built here) on the box's host cores over a bounded sample of the same table.

The default invocation also runs the other four configs of BASELINE.json — Q6, the high-cardinality GROUP BY, Q3 (shuffle join) and SSB Q4.1
(broadcast star join) — with fewer steps each and reports them under `queries.{q6,highcard,q3,ssb}` (ms_per_step, roofline, e2e, and at N = 1
cpu_baseline), plus `queries.q1_strong` at N > 1 (SF100 split over the ranks). Every workload carries `parity`: at every N the result of the
SAME multi-GPU plan (exchange included: gather + merge, hash shuffle, broadcast, runtime-filter merge, final TopN) over a bounded sample of the
same counter-based tables is compared with the oracle on rank 0; Q1 and Q6 additionally compare the FULL timed result (`parity_full`): every
rank runs the oracle's first phase over its whole resident shard, in slices, and rank 0 merges the ranks' intermediates through the oracle's
merge phase. The oracle is only the checker / the CPU baseline here, never on the measured GPU path.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time
import traceback

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SF100_LINEITEM = 600_037_902
HIGHCARD_ROWS, HIGHCARD_KEYS = 600_000_000, 100_000_000
SUB_QUERIES = ["q6", "highcard", "q3", "ssb"]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="srb")
    ap.add_argument("--query", default="q1", choices=["q1", "q6", "highcard", "q3", "ssb"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="scan queries at N > 1: every rank its own SF100 shard (weak) or SF100 split over the ranks (strong)")
    ap.add_argument("--rows", type=int, default=0, help="rows per rank (default: the SF100 table)")
    ap.add_argument("--e2e-rows", type=int, default=256 << 20, help="rows per end-to-end step fed as encoded pages (the decoded-column variant uses at most 64 Mi)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="CPU time budget of every cpu_baseline (1 warm-up + best of 5 on a sample sized for it)")
    ap.add_argument("--sub-steps", type=int, default=5, help="timed steps of every workload under `queries`")
    ap.add_argument("--parity-rows", type=int, default=8 << 20, help="rows (fact-table rows) of the bounded parity sample, over all ranks")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-sub", action="store_true", help="headline only: skip the workloads under `queries`")
    ap.add_argument("--phase-times", action="store_true", help="add the per-stage host times of a step to the JSON line")
    ap.add_argument("--gather", choices=["device", "nccl", "gloo", "torch"], default="device",
                    help="N > 1, low-cardinality plans: device = the ranks' aggregate states gathered and merged on the device in one call "
                         "(srb_op_set_finishing_merged); nccl / gloo / torch = serialize -> host rows gathered -> merge operator")
    return ap.parse_args()


def workload(query: str):
    from starrocks_b200 import queries as Q, tpch_synth as T
    if query == "q1":
        return dict(name="TPC-H SF100 Q1: lineitem filter + GROUP BY (4 groups) + 8 DECIMAL aggregates", table="lineitem", cols=T.LINEITEM_Q1,
                    plan=Q.q1_plan, bytes_per_row=44, rows=SF100_LINEITEM, kw={}, metric="TPC-H SF100 Q1 rows/s")
    if query == "q6":
        return dict(name="TPC-H SF100 Q6: lineitem range filter + SUM(DECIMAL128)", table="lineitem", cols=T.LINEITEM_Q6, plan=Q.q6_plan,
                    bytes_per_row=28, rows=SF100_LINEITEM, kw={}, metric="TPC-H SF100 Q6 rows/s")
    return dict(name="GROUP BY 100M distinct BIGINT keys: SUM, COUNT over 600M rows", table="highcard", cols=["k", "v"],
                plan=lambda **kw: Q.highcard_plan(**dict({"group_hint": HIGHCARD_KEYS}, **kw)), bytes_per_row=16, rows=HIGHCARD_ROWS,
                kw={"n_keys": HIGHCARD_KEYS}, metric="high-cardinality GROUP BY rows/s")


# ---- clocks sampled DURING the timed region -----------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.sm_max = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8), "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.05)

    def finish(self):
        self.stop_flag = True
        self.join(timeout=2)
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons)}


# ---- the CPU legs: oracle/ as the timed baseline and as the checker (the only places bench.py touches oracle/) --------------------------
def _oracle():
    from tests.oracle_lib import Oracle
    return Oracle()


def _best_of(run, n: int = 5):
    """1 warm-up + best of n -> (seconds, last result)"""
    out = run()
    best = None
    for _ in range(n):
        t0 = time.perf_counter()
        out = run()
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best, out


def cpu_scan_baseline(w, seconds: float, threads: int, phase=None):
    """oracle/ (kind "port") with `threads` host threads on a sample sized so that 1 warm-up + 5 runs take about `seconds`; T = 1 beside it"""
    from starrocks_b200 import tpch_synth as T
    orc = _oracle()
    probe_rows = 1 << 20
    plan = w["plan"]
    ch = T.host_chunk(w["table"], 0, probe_rows, w["cols"], **w["kw"])
    orc.scan_agg(plan(), ch, n_threads=threads)
    t0 = time.perf_counter()
    orc.scan_agg(plan(), ch, n_threads=threads)
    rate = probe_rows / (time.perf_counter() - t0)
    rows = int(min(max(rate * seconds / 7.0, probe_rows), 64 << 20))
    ch = T.host_chunk(w["table"], 0, rows, w["cols"], **w["kw"])
    best, _ = _best_of(lambda: orc.scan_agg(plan(), ch, n_threads=threads))
    r1 = max(probe_rows, rows // max(1, threads))
    ch1 = T.host_chunk(w["table"], 0, r1, w["cols"], **w["kw"])
    t0 = time.perf_counter()
    orc.scan_agg(plan(), ch1, n_threads=1)
    one = r1 / (time.perf_counter() - t0)
    return {"value": rows / best, "unit": "rows/s", "cores": threads, "kind": "port", "value_1_thread": one,
            "value_from_pages": cpu_from_pages(orc, w, ch, threads),
            "sample": f"first {rows} rows of the same synthetic table, 4096-row chunks, {threads} threads with private tables + merge, "
                      f"1 warm-up + best of 5; value_1_thread: one run over {r1} rows; value_from_pages: the same sample fed as frame-of-reference "
                      f"pages (decode with {threads} workers + the pipeline, best of 2), the input of the GPU arm's e2e"}


def cpu_from_pages(orc, w, ch, threads: int, runs: int = 2):
    """the CPU pipeline fed like the GPU arm's e2e: the chunk's columns as frame-of-reference pages (built untimed by the oracle's encoder),
    every run = decode all pages (`threads` workers over the pages) + scan / aggregate -> rows/s (best of `runs`)"""
    import numpy as np
    from starrocks_b200 import capi as K
    page_rows = 1 << 16
    enc = {}
    for slot, c in ch.columns.items():
        a = np.asarray(c.data)
        parts, offs, rows = [], [0], [0]
        for s0 in range(0, len(a), page_rows):
            p = orc.for_encode(a[s0:s0 + page_rows])
            parts.append(p)
            offs.append(offs[-1] + len(p))
            rows.append(min(len(a), s0 + page_rows))
        enc[slot] = (np.frombuffer(b"".join(parts) + b"\0" * 16, np.uint8), np.array(offs, np.int64), np.array(rows, np.int64), a.dtype, c)
    best = None
    for _ in range(runs):
        t0 = time.perf_counter()
        dec = K.Chunk(ch.n_rows)
        for slot, (data, offs, rows, dt, c) in enc.items():
            dec.columns[slot] = K.Column.from_ints(c.type, orc.for_decode_pages(data, offs, rows, dt, threads), None, c.precision, c.scale)
        orc.scan_agg(w["plan"](), dec, n_threads=threads)
        dt_ = time.perf_counter() - t0
        best = dt_ if best is None else min(best, dt_)
    return ch.n_rows / best


JOIN_METRIC = {"q3": "TPC-H SF100 Q3 rows/s (lineitem rows)", "ssb": "SSB SF100 Q4.1 rows/s (lineorder rows)"}
JOIN_NAME = {"q3": "TPC-H SF100 Q3: 3 filters, customer |><| orders |><| lineitem hash joins (build/probe/gather), GROUP BY 3 keys, ORDER BY revenue DESC, o_orderdate LIMIT 10",
             "ssb": "SSB SF100 Q4.1: 3 dimension filters, 4 hash joins (broadcast build sides), GROUP BY d_year, c_nation, ORDER BY d_year, c_nation"}
Q3_COLS = {"customer": ["c_custkey", "c_mktsegment_code"], "orders": ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"],
           "lineitem": ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"]}
SSB_TOP = ([0, 1], [1, 1], [1, 1], 1024)  # ORDER BY d_year, c_nation (Q4.1.sql): key slots, asc, nulls first, limit >= the <= 35 groups


def join_sizes(query: str, n_fact: int):
    """table -> rows for a fact table of n_fact rows (the SF100 ratios)"""
    if query == "q3":
        return {"customer": max(30, n_fact // 40), "orders": n_fact // 4, "lineitem": n_fact}
    from starrocks_b200 import tpch_synth as T
    return {"ssb_customer": max(100, n_fact // 200), "ssb_supplier": max(50, n_fact // 3000), "ssb_part": max(100, n_fact // 430), "ssb_dates": T.SSB_DAYS,
            "lineorder": n_fact}


def join_cols_kw(query: str, sizes):
    from starrocks_b200 import pipelines as P, tpch_synth as T
    if query == "q3":
        return Q3_COLS, {"orders": {"n_customers": sizes["customer"]}}
    cols = dict(P.SSB_DIM_COLS, lineorder=T.LINEORDER_Q41)
    return cols, {"lineorder": dict(n_customers=sizes["ssb_customer"], n_suppliers=sizes["ssb_supplier"], n_parts=sizes["ssb_part"])}


def join_alg_bytes(query: str, sizes) -> int:
    """every column a table contributes to the query, counted once in full"""
    if query == "q3":
        return sizes["customer"] * 5 + sizes["orders"] * 16 + sizes["lineitem"] * 24
    return sizes["lineorder"] * 24 + sizes["ssb_customer"] * 12 + sizes["ssb_supplier"] * 8 + sizes["ssb_part"] * 8 + sizes["ssb_dates"] * 8


def join_host_tables(query: str, n_fact: int):
    from starrocks_b200 import tpch_synth as T
    sizes = join_sizes(query, n_fact)
    cols, kw = join_cols_kw(query, sizes)
    return {t: T.host_chunk(t, 0, n, cols[t], **kw.get(t, {})) for t, n in sizes.items()}


def join_oracle_chain(orc, query: str, tb, threads: int):
    """-> (groups, final ORDER BY rows) of the oracle's operator chain"""
    from tests import oracle_chains as OC
    if query == "q3":
        groups = OC.q3_oracle(orc, tb["customer"], tb["orders"], tb["lineitem"], n_threads=threads)
        return groups, OC.q3_top_oracle(orc, groups)
    groups = OC.ssb_q41_oracle(orc, tb, tb["lineorder"], n_threads=threads)
    return groups, orc.sort(groups, SSB_TOP[0], SSB_TOP[1], SSB_TOP[2], 0, SSB_TOP[3])


def cpu_join_baseline(query: str, seconds: float, threads: int):
    """the oracle's operator chain with `threads` workers in every operator (scan/filter and aggregation with private tables, probe drivers
    over one shared join table, the build in one driver as the reference's) on a sample sized for `seconds`; T = 1 beside it"""
    orc = _oracle()
    probe_rows = 1 << 20
    tb = join_host_tables(query, probe_rows)
    join_oracle_chain(orc, query, tb, threads)
    t0 = time.perf_counter()
    join_oracle_chain(orc, query, tb, threads)
    rate = probe_rows / (time.perf_counter() - t0)
    rows = int(min(max(rate * seconds / 7.0, probe_rows), 48 << 20))
    tb = join_host_tables(query, rows)
    best, _ = _best_of(lambda: join_oracle_chain(orc, query, tb, threads))
    r1 = max(probe_rows, rows // max(1, threads))
    tb1 = join_host_tables(query, r1)
    t0 = time.perf_counter()
    join_oracle_chain(orc, query, tb1, 1)
    one = r1 / (time.perf_counter() - t0)
    return {"value": rows / best, "unit": "rows/s", "cores": threads, "kind": "port", "value_1_thread": one,
            "sample": f"the first {rows} fact-table rows (dimension tables scaled with them) through the oracle's operator chain, {threads} workers per "
                      f"operator (probe drivers share one join table), 1 warm-up + best of 5; value_1_thread: one run over {r1} fact rows"}


def base_config(w_name: str, rows_per_gpu, bytes_per_row, world: int, scaling: str):
    return {"workload": w_name, "rows_per_gpu": rows_per_gpu, "bytes_per_row": bytes_per_row, "l2": "inputs (>= 14 GB per step) exceed the 126 MB L2",
            "parallelism": f"range-partitioned x{world} ({scaling})"}


def run_reference(args):
    """--impl reference: the reference's CPU pipeline for this path (oracle port; the BE itself cannot be built here), all host threads,
    every step a bounded sample of the arm's workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = int(os.environ.get("WORLD_SIZE", "1"))
    threads = os.cpu_count() or 1
    per_step = max(4.0, min(20.0, 100.0 / max(1, args.steps + args.warmup)))
    orc = _oracle()
    if args.query in ("q3", "ssb"):
        probe_rows = 1 << 20
        tb = join_host_tables(args.query, probe_rows)
        t0 = time.perf_counter()
        join_oracle_chain(orc, args.query, tb, threads)
        rate = probe_rows / (time.perf_counter() - t0)
        rows = int(min(max(rate * per_step, probe_rows), 48 << 20))
        tb = join_host_tables(args.query, rows)
        run = lambda: join_oracle_chain(orc, args.query, tb, threads)
        name, metric, scaling, bpr = JOIN_NAME[args.query], JOIN_METRIC[args.query], "strong", 24
        rows_cfg = join_sizes(args.query, args.rows or SF100_LINEITEM)
        sample = f"each step: the first {rows} fact-table rows (dimension tables scaled with them), {threads} workers per operator"
    else:
        from starrocks_b200 import tpch_synth as T
        w = workload(args.query)
        probe_rows = 1 << 20
        ch = T.host_chunk(w["table"], 0, probe_rows, w["cols"], **w["kw"])
        t0 = time.perf_counter()
        orc.scan_agg(w["plan"](), ch, n_threads=threads)
        rate = probe_rows / (time.perf_counter() - t0)
        rows = int(min(max(rate * per_step, probe_rows), 64 << 20))
        ch = T.host_chunk(w["table"], 0, rows, w["cols"], **w["kw"])
        run = lambda: orc.scan_agg(w["plan"](), ch, n_threads=threads)
        from_pages = cpu_from_pages(orc, w, ch, threads)
        name, metric, scaling, bpr = w["name"], w["metric"], args.scaling, w["bytes_per_row"]
        rows_cfg = args.rows or w["rows"]
        sample = f"each step: first {rows} rows of the same synthetic table, 4096-row chunks, {threads} threads"
    for _ in range(args.warmup):
        run()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run()
    dt = time.perf_counter() - t0
    value = rows * args.steps / dt
    cb = {"value": value, "unit": "rows/s", "cores": threads, "kind": "port", "sample": sample}
    if args.query not in ("q3", "ssb"):
        cb["value_from_pages"] = from_pages
        cb["sample"] += ("; `value` times the pipeline over DECODED host columns (the easier job: the GPU arm's e2e starts from encoded pages); "
                         "value_from_pages = the same sample fed as frame-of-reference pages, decode included")
    print(json.dumps({"impl": "reference", "metric": metric, "value": value, "unit": "rows/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "int128", "data": "synthetic",
                      "config": base_config(name, rows_cfg, bpr, world, scaling), "rows_per_step": rows,
                      "cpu_baseline": cb, "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ---- the GPU arm: environment ------------------------------------------------------------------------------------------------------
class Env:
    pass


def setup_env(args) -> Env:
    import torch
    import torch.distributed as dist
    from starrocks_b200 import distributed as D, ops
    env = Env()
    env.torch, env.dist = torch, dist
    env.rank = int(os.environ.get("RANK", "0"))
    env.world = int(os.environ.get("WORLD_SIZE", "1"))
    env.local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(env.local)
    env.dev = torch.device("cuda", env.local)
    if env.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=env.dev)
    env.ctx = ops.Context(device=env.local, rank=env.rank, world_size=env.world)
    env.comm = env.part = env.host_ex = None
    if env.world > 1:
        # libsrb's own NCCL communicator moves the rows; torch.distributed only carries the 128-byte NCCL id (and the harness's own scalars)
        def put(b):
            dist.broadcast_object_list([b], src=0)

        def get():
            box = [None]
            dist.broadcast_object_list(box, src=0)
            return box[0]

        env.comm = ops.Communicator(env.ctx, put, get)
        env.part = ops.Partitioner(env.ctx, env.world)
        if args.gather in ("device", "nccl"):
            env.host_ex = env.comm
        else:
            env.host_ex = D.ObjectExchange(dist, env.rank, env.world, group=dist.new_group(backend="gloo") if args.gather == "gloo" else None)
    env.cores = os.cpu_count() or 1
    env.peaks = {}
    try:
        env.peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    env.peak = float(env.peaks.get("hbm_gbs", 6650.0))
    env.peak_source = "MEASURED_PEAKS.json hbm_gbs (measured copy bandwidth)" if env.peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    return env


def barrier(env):
    if env.world > 1:
        env.dist.barrier()
    env.torch.cuda.synchronize()


def max_over_ranks(env, vals):
    t = env.torch.tensor(list(vals), dtype=env.torch.float64, device=env.dev)
    if env.world > 1:
        env.dist.all_reduce(t, op=env.dist.ReduceOp.MAX)
    return t.tolist()


def sum_over_ranks(env, val: int) -> int:
    t = env.torch.tensor([val], dtype=env.torch.int64, device=env.dev)
    if env.world > 1:
        env.dist.all_reduce(t)
    return int(t.item())


def gather_objects(env, obj):
    """-> list of every rank's object on rank 0 (None elsewhere)"""
    if env.world == 1:
        return [obj]
    box = [None] * env.world if env.rank == 0 else None
    env.dist.gather_object(obj, box, dst=0)
    return box


def all_true(env, flag) -> bool:
    t = env.torch.tensor([1 if flag else 0], dtype=env.torch.int64, device=env.dev)
    if env.world > 1:
        env.dist.all_reduce(t, op=env.dist.ReduceOp.MIN)
    return bool(t.item())


def free_device(env):
    import gc
    gc.collect()
    env.torch.cuda.empty_cache()
    env.ctx.trim()


def rows_key(chunk, sort=True):
    from starrocks_b200 import capi as K
    return K.chunk_rows(chunk, sort)


def big_chunks_equal(a, b) -> bool:
    """order-normalised equality of two host chunks of fixed-width columns, in numpy (10^6..10^8 rows)"""
    import numpy as np
    from starrocks_b200 import capi as K
    if a.n_rows != b.n_rows or list(a.columns) != list(b.columns):
        return False

    def cols(ch):
        out = []
        for c in ch.columns.values():
            d = np.asarray(c.data)
            nl = np.asarray(c.nulls) if c.nulls is not None else np.zeros(ch.n_rows, np.uint8)
            parts = [d["lo"], d["hi"]] if d.dtype == K.I128 else [d]
            for p in parts:
                out.append(np.where(nl != 0, np.zeros(1, p.dtype), p))
            out.append(nl)
        order = np.lexsort(out[::-1])
        return [x[order] for x in out]

    return all(bool((x == y).all()) for x, y in zip(cols(a), cols(b)))


# ---- scan -> filter -> aggregate queries (Q1, Q6, high-cardinality GROUP BY) --------------------------------------------------------
def merge_plan(query: str, inter, world: int):
    from starrocks_b200 import queries as Q
    if query == "q1":
        return Q.q1_merge_plan(inter)
    if query == "q6":
        return Q.q6_merge_plan(inter)
    return Q.highcard_merge_plan(group_hint=max(1024, HIGHCARD_KEYS // world))


_PLANS = {}


def scan_step(env, query: str, w, chunk, marks=None, collect=None, device_result=False, group_hint=None):
    """one pass of the hot path over this rank's shard; at N > 1 it includes the exchange and the phase-2 merge (starrocks_b200/distributed.py:
    gather for the low-cardinality plans, hash shuffle for the high-cardinality one) -> (result chunk or None, result rows on this rank)"""
    from starrocks_b200 import capi as K, ops, queries as Q
    ctx, comm = env.ctx, env.comm
    two_phase = env.world > 1

    def mark():
        if marks is not None:
            marks.append(time.perf_counter())

    mark()
    kw = {} if group_hint is None else {"group_hint": group_hint}
    merged_finish = two_phase and query in ("q1", "q6") and env.gather_device
    # the plan is what the FE hands the BE once per query: built once, not once per step
    pkey = (query, two_phase and not merged_finish, group_hint)
    plan = _PLANS.get(pkey)
    if plan is None:
        plan = _PLANS[pkey] = w["plan"](phase=K.AGG_UPDATE_SERIALIZE, **kw) if two_phase and not merged_finish else w["plan"](**kw)
    op = ops.ScanAggOperator(ctx, plan)
    op2 = None
    try:
        op.push_chunk(chunk)
        if merged_finish:
            # phase 1 -> gather of the ranks' aggregate states -> merge -> finalize in ONE call, on the device (srb_op_set_finishing_merged)
            mark()
            op.set_finishing_merged(comm)
            two_phase = False
        else:
            op.set_finishing()
        mark()
        if two_phase:
            if query in ("q1", "q6"):
                # <= 4 intermediate rows per rank (AVG states are {sum,count} blobs): gathered, then merged on every rank
                inter = op.pull_all()
                mark()
                recv = env.host_ex.all_gather(inter)
                mark()
                op2 = ops.ScanAggOperator(ctx, merge_plan(query, inter, env.world))
            else:
                inter = op.pull_device()
                mark()
                if os.environ.get("SRB_NO_P2P"):  # A/B: device partition, then NCCL send/recv of the grouped rows
                    grouped, counts = env.part.run(inter, [0])
                    recv, _ = comm.all_to_all(grouped, counts)
                else:  # partition scatter fused with the exchange: rows go straight into the peers' receive buffers
                    recv, _ = comm.shuffle(inter, [0])
                mark()
                hint = max(1024, (group_hint or HIGHCARD_KEYS) // env.world)
                op2 = ops.ScanAggOperator(ctx, Q.highcard_merge_plan(group_hint=hint))
            op2.push_chunk(recv)
            op2.set_finishing()
            mark()
            last = op2
        else:
            last = op
        if query == "highcard":
            out = last.pull_device()  # ~10^8 result rows stay on the device (24 B/row); only the row count is read
            n_out = out.n_rows if out is not None else 0
            if device_result and out is not None:
                out = ctx.fetch_chunk(out)
            elif not device_result:
                out = None
        else:
            out = last.pull_all()
            n_out = out.n_rows
        mark()
        if collect is not None:
            c, p = op.counters(), op.profile()
            if op2 is not None:
                c = dict(c, kernel_launches=c["kernel_launches"] + op2.counters()["kernel_launches"])
                p = dict(p, phase2=op2.profile(), phase2_rows_in=op2.counters()["rows_in"])
            collect.append((c, p))
        return out, n_out
    finally:
        op.close()
        if op2 is not None:
            op2.close()


def shape_name(stats):
    shapes = {p["shape"] for _, p in stats}
    static = sum(p["static_launches"] for _, p in stats)
    if static and 100 in shapes:
        return f"run-time compiled shape (TMA-staged static kernel, NVRTC), {static} static launches"
    return f"static shape {sorted(shapes)} (TMA-staged), {static} static launches" if static else "generic interpreter"


def view_rows(env, w, tensors, a: int, n: int):
    """device Chunk over rows [a, a+n) of resident column tensors"""
    from starrocks_b200 import capi as K, tpch_synth as T
    ch = K.Chunk(n)
    for name in w["cols"]:
        slot, t, p, s = T._SLOT[name]
        x = tensors[name]
        ch.columns[slot] = K.Column(t, x.data_ptr() + a * x.element_size(), None, None, p, s, K.MEM_DEVICE)
    return ch


def scan_e2e(env, args, query, w, chunk, rows, warmup, steps):
    """Host buffers -> C ABI -> result on the host. Pinned host copies of a slice of the shard are pushed as SRB_MEM_PINNED chunks; the library
    stages them into HBM in slices overlapping copies with the kernel; the result rows are read back."""
    import ctypes as C
    from starrocks_b200 import capi as K, ops
    ctx = env.ctx
    n = int(min(args.e2e_rows, rows, 64 << 20))
    host = K.Chunk(n)
    pins, h2d = [], 0
    for slot, c in chunk.columns.items():
        wbytes = K.type_width(c.type)
        p = C.c_void_p()
        ops._check(ctx.L.srb_pinned_alloc(ctx.h, n * wbytes, C.byref(p)))
        ops._check(ctx.L.srb_memcpy_d2h(ctx.h, p, c.data, n * wbytes))
        pins.append(p)
        host.columns[slot] = K.Column(c.type, p.value, None, None, c.precision, c.scale, K.MEM_PINNED)
        h2d += n * wbytes

    def step():
        op = ops.ScanAggOperator(ctx, w["plan"]())
        op.push_chunk(host)
        op.set_finishing()
        out = op.pull_all() if query != "highcard" else op.pull_device()
        op.close()
        return out

    try:
        for _ in range(max(1, warmup - 1)):
            out = step()
        barrier(env)
        steps = max(2, min(steps, 5))
        t0 = time.perf_counter()
        for _ in range(steps):
            out = step()
        barrier(env)
        dt = max_over_ranks(env, [time.perf_counter() - t0])[0]
    finally:
        for p in pins:
            ctx.L.srb_pinned_free(ctx.h, p)
    d2h = sum(len(K.column_pyvalues(c)) * max(K.type_width(c.type), 24) for c in out.columns.values()) if query != "highcard" else 8
    return {"value": n * env.world * steps / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "rows_per_step": n,
            "steps": steps, "note": "pinned host columns pushed through srb_op_push_chunk; staged H2D in 4 Mi-row slices overlapped with the kernel"
                                    + ("; the ~10^7 result groups stay on the device, the row count is read" if query == "highcard" else "")}


def scan_e2e_pages(env, args, query, w, tensors, row0, rows, warmup, steps):
    """The same query fed the way the BE's scan is fed: ENCODED data pages (frame-of-reference, the reference's FOR_ENCODING) in pinned host
    memory -> srb_op_push_pages (H2D copy of the page bytes, decode on the device, scan, aggregate) -> result on the host. The pages are built
    (untimed) from the resident synthetic columns by the torch page builder, byte-identical to the reference's ForEncoder output."""
    import ctypes as C
    from starrocks_b200 import capi as K, ops, pages as PG, tpch_synth as T
    torch, ctx = env.torch, env.ctx
    n = int(min(args.e2e_rows, rows))
    # 64 Ki values per page, 32 Mi rows per page set: a column's encoded bytes of a set (~40 MB for an 8-byte column of Q1) cross PCIe as ONE
    # copy, and copies below ~30 MB do not reach the link's rate (measured: 7 x 9.5 MB 46.7 GB/s, 7 x 38 MB 55.2 GB/s; tools/h2d_probe.py)
    page_rows, set_pages = 1 << 16, int(os.environ.get("SRB_E2E_SET_PAGES", "512"))
    cols, pins, h2d = {}, [], 0
    try:
        for name in w["cols"]:
            slot, t, p, s = T._SLOT[name]
            enc, pages = PG.for_encode_column(tensors[name][row0:row0 + n], page_rows)
            nb = int(enc.numel())
            ptr = C.c_void_p()
            ops._check(ctx.L.srb_pinned_alloc(ctx.h, nb + 64, C.byref(ptr)))
            pins.append(ptr)
            ops._check(ctx.L.srb_memcpy_d2h(ctx.h, ptr, enc.data_ptr(), nb))
            cols[slot] = PG.PageColumn(t, ptr.value, nb, pages, PG.ENC_FOR, p, s, K.MEM_PINNED)
            h2d += nb
            del enc
        torch.cuda.empty_cache()
        n_pages = len(next(iter(cols.values())).pages)
        sets = [PG.prepare({slot: PG.slice_pages(c, a, min(a + set_pages, n_pages)) for slot, c in cols.items()}) for a in range(0, n_pages, set_pages)]

        def step():
            op = ops.ScanAggOperator(ctx, w["plan"]())
            try:
                for ps in sets:
                    op.push_pages_prepared(ps)
                op.set_finishing()
                return op.pull_all() if query != "highcard" else op.pull_device()
            finally:
                op.close()

        for _ in range(max(1, warmup - 1)):
            out = step()
        barrier(env)
        steps = max(2, min(steps, 5))
        t0 = time.perf_counter()
        for _ in range(steps):
            out = step()
        barrier(env)
        dt = max_over_ranks(env, [time.perf_counter() - t0])[0]
    finally:
        for p in pins:
            ctx.L.srb_pinned_free(ctx.h, p)
    d2h = sum(len(K.column_pyvalues(c)) * max(K.type_width(c.type), 24) for c in out.columns.values()) if query != "highcard" else 8
    return {"value": n * env.world * steps / dt, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "rows_per_step": n, "steps": steps,
            "encoded_bytes_per_row": h2d / n, "decoded_bytes_per_row": w["bytes_per_row"],
            "note": "frame-of-reference data pages (the reference's FOR_ENCODING, 64 Ki values per page) in pinned host memory pushed through "
                    "srb_op_push_pages in page sets of 32 Mi rows: H2D copy of the encoded bytes, decode on the device, scan, aggregate; the copy of a "
                    "page set overlaps the kernels of the previous one" + ("; the ~10^7 result groups stay on the device, the row count is read" if query == "highcard" else "")}, out


def scan_parity_sample(env, args, query, w) -> bool:
    """the SAME multi-GPU plan (phase 1 -> exchange -> phase 2) over a bounded sample split over the ranks, against the oracle on rank 0"""
    from starrocks_b200 import distributed as D, tpch_synth as T
    S = args.parity_rows
    kw = dict(w["kw"])
    hint = None
    if query == "highcard":
        kw["n_keys"] = max(1000, S // 6)  # the full config's six rows per key
        hint = kw["n_keys"]
    b, e = D.shard_range(S, env.rank, env.world)
    shard = T.host_chunk(w["table"], b, e - b, w["cols"], **kw)
    dch, ptrs = env.ctx.device_chunk(shard)
    try:
        got, _ = scan_step(env, query, w, dch, device_result=True, group_hint=hint)
    finally:
        for p in ptrs:
            env.ctx.device_free(p)
    if query == "highcard":  # every group lives on exactly one rank after the shuffle: rank 0 concatenates
        parts = gather_objects(env, got)
        if env.rank == 0:
            got = D.concat_chunks([p for p in parts if p is not None and p.n_rows])
    ok = True
    if env.rank == 0:
        plan = w["plan"](group_hint=hint) if hint else w["plan"]()
        want = _oracle().scan_agg(plan, T.host_chunk(w["table"], 0, S, w["cols"], **kw), n_threads=env.cores)
        ok = big_chunks_equal(got, want) if query == "highcard" else rows_key(got) == rows_key(want)
    return all_true(env, ok)


def scan_parity_full(env, query, w, tensors, row0: int, rows: int, final) -> dict:
    """the timed step's own result against the oracle over EVERY resident row: each rank runs the oracle's first phase over its shard in
    32 Mi-row slices (copied out of HBM through pinned buffers), rank 0 merges all intermediates through the oracle's merge phase"""
    from starrocks_b200 import capi as K, distributed as D, tpch_synth as T
    torch = env.torch
    t0 = time.perf_counter()
    orc = _oracle()
    threads = max(1, env.cores // env.world)
    piece = 32 << 20
    stage = {name: torch.empty(min(piece, rows), dtype=tensors[name].dtype, pin_memory=True) for name in w["cols"]}
    inter = []
    for a in range(0, rows, piece):
        m = min(piece, rows - a)
        ch = K.Chunk(m)
        for name in w["cols"]:
            stage[name][:m].copy_(tensors[name][row0 + a:row0 + a + m], non_blocking=True)
        torch.cuda.synchronize()
        for name in w["cols"]:
            slot, t, p, s = T._SLOT[name]
            ch.columns[slot] = K.Column.from_ints(t, stage[name][:m].numpy(), None, p, s)
        inter.append(orc.scan_agg(w["plan"](phase=K.AGG_UPDATE_SERIALIZE), ch, n_threads=threads))
    mine = D.concat_chunks(inter)
    parts = gather_objects(env, mine)
    ok = True
    if env.rank == 0:
        allp = D.concat_chunks(parts)
        want = orc.scan_agg(merge_plan(query, allp, env.world), allp)
        ok = rows_key(final) == rows_key(want)
    ok = all_true(env, ok)
    return {"ok": ok, "rows_checked": rows * env.world, "seconds": round(time.perf_counter() - t0, 2),
            "how": f"oracle phase 1 over every rank's whole shard ({threads} threads per rank, 32 Mi-row slices), merged by the oracle's merge phase on rank 0"}


def bench_scan(env, args, query: str, steps: int, warmup: int, scaling: str, tensors=None, headline=False, label=None):
    """-> (result dict, resident tensors). `tensors` = already resident lineitem columns to reuse (Q6 after Q1, the strong-scaling pass)."""
    from starrocks_b200 import tpch_synth as T
    torch = env.torch
    w = workload(query)
    full = args.rows or w["rows"]
    world, rank = env.world, env.rank
    if tensors is None:  # resident shard of the weak layout: rows [rank*full, (rank+1)*full) of the synthetic table
        _, tensors = T.device_chunk(w["table"], rank * full, full, w["cols"], env.dev, **w["kw"])
        torch.cuda.synchronize()
    if scaling == "strong" and world > 1:
        from starrocks_b200 import distributed as D
        b, e = D.shard_range(full, rank, world)
        row0, rows = 0, e - b  # the rank's first full/world resident rows stand for its range of the one SF100 table
    else:
        row0, rows = 0, full
    chunk = view_rows(env, w, tensors, row0, rows)
    phase_ms = [] if args.phase_times else None

    def step(collect=None):
        marks = [] if phase_ms is not None else None
        r = scan_step(env, query, w, chunk, marks=marks, collect=collect)
        if marks is not None:
            phase_ms.append([(marks[i + 1] - marks[i]) * 1e3 for i in range(len(marks) - 1)])
        return r

    for _ in range(warmup):
        step()
    sampler = ClockSampler(env.local)
    stats = []
    barrier(env)
    sampler.start()
    alloc0 = env.ctx.alloc_stats()
    t0 = time.perf_counter()
    for _ in range(steps):
        out, n_out = step(stats)
    barrier(env)
    dt = time.perf_counter() - t0
    alloc1 = env.ctx.alloc_stats()
    clocks = sampler.finish()
    # device-side time of the timed steps: CUDA events recorded by the library on the operator's own stream
    kernel_ms = sum(p["kernel_ms"] for _, p in stats)
    batches = sum(p["batches"] for _, p in stats)
    launches = sum(c["kernel_launches"] for c, _ in stats)
    dt_max, kernel_s_max = max_over_ranks(env, [dt, kernel_ms / 1e3])
    total_rows = sum_over_ranks(env, rows)
    if query == "highcard":
        n_out = sum_over_ranks(env, n_out)  # groups over all ranks (every group lives on exactly one rank after the shuffle)
    result = {}
    per_batch_rows = rows * steps / max(1, batches)
    achieved = w["bytes_per_row"] * per_batch_rows / (kernel_s_max / max(1, batches)) / 1e9 if kernel_s_max > 0 else 0.0
    if query == "highcard":  # 16 B/row in + 24 B per group out, over the whole step (the path is several kernels)
        alg = rows * 16 + (n_out // world) * 24
        achieved = alg / (kernel_s_max / steps) / 1e9 if kernel_s_max > 0 else 0.0
    # DRAM bytes per launch from the committed `ncu --set full` capture of this kernel (profiles/traffic.json, written by
    # tools/ncu_summary.py): measured bytes per row of the captured launch x the rows of an average launch here
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(query)
        if tj and tj.get("rows_per_launch"):
            traffic = tj["dram_bytes_per_launch"] / tj["rows_per_launch"] * per_batch_rows
            traffic_src = f"{tj['report']}: {tj['dram_bytes_per_launch'] / tj['rows_per_launch']:.2f} B/row x rows per launch"
    except Exception:
        pass
    result = {
        "metric": w["metric"], "value": total_rows * steps / dt_max, "unit": "rows/s", "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": dt_max / steps * 1e3, "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "int128" if query != "highcard" else "int64",
        "data": "synthetic", "config": base_config(w["name"], rows, w["bytes_per_row"], world, scaling), "kernel": shape_name(stats),
        "scan_gb_per_s": total_rows * steps * w["bytes_per_row"] / dt_max / 1e9, "gpu_launches": launches,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": env.peak, "unit": "GB/s", "frac": achieved / env.peak if env.peak else None,
                     "traffic": traffic, "traffic_source": traffic_src, "peak_source": env.peak_source,
                     "launch": (f"{int(per_batch_rows)} rows x {w['bytes_per_row']} B per launch, {kernel_s_max / max(1, batches) * 1e3:.3f} ms avg (CUDA events on the op stream)"
                                if query != "highcard" else f"whole phase-1 operator per step: {rows} rows x 16 B in + {n_out // world} groups x 24 B out, "
                                f"{kernel_s_max / steps * 1e3:.3f} ms of kernels (CUDA events on the op stream)")},
        "clocks": clocks, "result_rows": n_out,
        # the timed steps should live off the context's buffer pool: driver allocator calls inside them, and the time they took (rank 0)
        "allocator_in_timed_steps_rank0": {k: alloc1[k] - alloc0[k] for k in ("n_malloc", "n_free", "malloc_us", "free_us")} | {"pooled_bytes": alloc1["pooled_bytes"]},
    }
    if stats and "phase2" in stats[-1][1]:
        result["phase2_profile_rank0"] = dict(stats[-1][1]["phase2"], rows_in=stats[-1][1]["phase2_rows_in"])
    if phase_ms:
        timed = phase_ms[-steps:]
        result["phase_ms_rank0"] = [sum(t[i] for t in timed) / len(timed) for i in range(len(timed[0]))]
    if not args.no_e2e:
        # headline e2e: the scan fed with encoded pages (what crosses PCIe is what a segment holds); beside it the same query fed with
        # already decoded host columns (srb_op_push_chunk), as round 1 measured it
        result["e2e"], e2e_out = scan_e2e_pages(env, args, query, w, tensors, row0, rows, warmup, steps)
        result["e2e_decoded_columns"] = scan_e2e(env, args, query, w, chunk, rows, warmup, steps)
        if query != "highcard" and not args.no_parity:  # the page-fed result over the e2e slice against the resident-column result of the same rows
            n_e = result["e2e"]["rows_per_step"]
            ref_out, _ = scan_step(env, query, w, view_rows(env, w, tensors, row0, n_e)) if env.world == 1 else (None, 0)
            if ref_out is not None:
                result["e2e"]["same_result_as_resident_columns"] = rows_key(e2e_out) == rows_key(ref_out)
    if not args.no_parity:
        result["parity"] = scan_parity_sample(env, args, query, w)
        result["parity_sample"] = f"{args.parity_rows} rows of the same generator split over {world} rank(s), the same plan (exchange + merge included) vs the oracle"
        if query in ("q1", "q6"):
            result["parity_full"] = scan_parity_full(env, query, w, tensors, row0, rows, out)
            result["parity"] = result["parity"] and result["parity_full"]["ok"]
        result["parity"] = result["parity"] and result.get("e2e", {}).get("same_result_as_resident_columns", True)
    if rank == 0 and world == 1 and not args.no_cpu:
        result["cpu_baseline"] = cpu_scan_baseline(w, args.cpu_seconds, env.cores)
    return result, tensors


def bench_q1_variants(env, args, tensors, steps: int, warmup: int):
    """Q1 as a real BE would hand it over — none of these matches a built-in kernel shape: VARCHAR(1) group keys as BinaryColumns
    (AggHashMapWithSerializedKey, agg_hash_map.h:652-898; 46 B/row in SURVEY 8d), nullable measures (null maps that do hold NULLs), one
    extra MIN / MAX, two more aggregates. Each gets the static TMA-staged kernel compiled for it at run time (csrc/srb_jit.h; the first step
    starts the compilation, srb_jit_wait() waits for it, the timed steps run the compiled kernel). Every variant is also run with SRB_JIT=off
    first: same rows, and the interpreter's (k_scan_agg_generic) time for the record."""
    import numpy as np
    from starrocks_b200 import capi as K, ops, queries as Q, tpch_synth as T
    from starrocks_b200.plan import ExprBuilder, ScanAggPlan, pred
    torch, ctx = env.torch, env.ctx
    w = workload("q1")
    n = int(min(w["rows"], 1 << 27))  # 134 M rows per launch: far beyond L2, a fifth of the table
    base = view_rows(env, w, tensors, 0, n)
    out = {}

    def run(name, plan_fn, chunk, bytes_per_row):
        torch.cuda.synchronize()  # the chunk's columns were just written on torch's stream; the library launches on its own

        def step():
            op = ops.ScanAggOperator(ctx, plan_fn())
            try:
                op.push_chunk(chunk)
                op.set_finishing()
                return op.pull_all(), op.profile()
            finally:
                op.close()
        old = os.environ.get("SRB_JIT")
        os.environ["SRB_JIT"] = "off"
        try:
            for _ in range(2):
                res_i, prof_i = step()
        finally:
            if old is None:
                os.environ.pop("SRB_JIT")
            else:
                os.environ["SRB_JIT"] = old
        t_c = time.perf_counter()
        step()
        ops.jit_wait()  # the kernel this plan's shape compiles to (if it is covered) is ready for the next operator
        compile_wait = time.perf_counter() - t_c
        for _ in range(max(1, warmup - 1)):
            step()
        env.torch.cuda.synchronize()
        t0 = time.perf_counter()
        ms = 0.0
        for _ in range(steps):
            res, prof = step()
            ms += prof["kernel_ms"]
        env.torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        ach = bytes_per_row * n / (ms / steps / 1e3) / 1e9 if ms else 0.0
        out[name] = {"ms_per_step": dt / steps * 1e3, "kernel_ms": ms / steps, "rows": n, "bytes_per_row": bytes_per_row, "kernel": shape_name([({}, prof)]),
                     "roofline": {"bound": "hbm", "achieved": ach, "peak": env.peak, "unit": "GB/s", "frac": ach / env.peak}, "result_rows": res.n_rows,
                     "interpreter_kernel_ms": prof_i["kernel_ms"], "same_rows_as_interpreter": rows_key(res) == rows_key(res_i),
                     "first_step_and_compile_s": round(compile_wait, 2)}
        return res

    # (a) VARCHAR(1) keys: bytes + u32 offsets on the device, decoded from the dictionary codes
    keep = []
    bin_chunk = K.Chunk(n, dict(base.columns))
    for slot, code_name, names in ((Q.L_RETURNFLAG, "l_returnflag_code", T.RETURNFLAGS), (Q.L_LINESTATUS, "l_linestatus_code", T.LINESTATUS)):
        lut = torch.tensor(list(b"".join(names)), dtype=torch.uint8, device=env.dev)
        data = torch.cat([lut[tensors[code_name][:n].long()], torch.zeros(16, dtype=torch.uint8, device=env.dev)])
        offs = torch.arange(n + 1, dtype=torch.int32, device=env.dev)
        keep += [data, offs]
        bin_chunk.columns[slot] = K.Column(K.TYPE_VARCHAR, data.data_ptr(), None, offs.data_ptr(), 0, 0, K.MEM_DEVICE, n)
    for s_ in (Q.L_RETURNFLAG_CODE, Q.L_LINESTATUS_CODE):
        bin_chunk.columns.pop(s_, None)
    a = run("q1_binary_keys", lambda: Q.q1_plan(dict_keys=False), bin_chunk, 46)
    # (b) nullable measures: 1 % NULLs in l_quantity and l_discount
    nul_chunk = K.Chunk(n, dict(base.columns))
    for slot in (Q.L_QUANTITY, Q.L_DISCOUNT):
        nm = (torch.rand(n, device=env.dev) < 0.01).to(torch.uint8)
        keep.append(nm)
        c = base.columns[slot]
        nul_chunk.columns[slot] = K.Column(c.type, c.data, nm.data_ptr(), None, c.precision, c.scale, K.MEM_DEVICE)
    run("q1_nullable_measures", Q.q1_plan, nul_chunk, 46)

    # (c) one more aggregate: MIN and MAX of l_extendedprice beside Q1's eight
    run("q1_plus_min_max", lambda: Q.q1_plan(with_min_max=True), base, 44)
    # (d) two more aggregates, one of them COUNT(x)
    run("q1_two_more_aggregates", lambda: Q.q1_plan(more_aggs=True), base, 44)
    out["jit"] = ops.jit_stats()
    # the binary-key variant must give Q1's numbers: same groups, keys as strings
    ref, _ = scan_step(env, "q1", w, base) if env.world == 1 else (None, 0)
    if ref is not None:
        strip = lambda rows: sorted(tuple(r[2:]) for r in rows)
        out["q1_binary_keys"]["same_aggregates_as_dict_keys"] = strip(rows_key(a)) == strip(rows_key(ref))
    del keep
    return out


# ---- the join pipelines (Q3 shuffle join, SSB Q4.1 broadcast star join) ---------------------------------------------------------
def join_step(env, query: str, tb, use_comm=True, marks=None, rows_out=None, want_groups=False):
    """one pass of the pipeline over the rank's shards -> (final ORDER BY rows (host), groups on this rank, kernel launches, fetched groups)"""
    from starrocks_b200 import capi as K, ops, pipelines as P, queries as Q
    ctx = env.ctx
    comm = env.comm if use_comm else None
    part = env.part if use_comm else None
    torch = env.torch

    def mark():
        if marks is not None:
            torch.cuda.synchronize()
            marks.append(time.perf_counter())

    mark()
    groups_host = None
    if query == "q3":
        q = P.TpchQ3(ctx, comm, part)
        try:
            agg = q.run(tb["customer"], tb["orders"], tb["lineitem"], mark=mark if marks is not None else None, rows=rows_out)
            groups = agg.pull_device()  # ~14 M groups stay on the device, partitioned by the order key
            n_groups = groups.n_rows if groups is not None else 0
            if want_groups:
                groups_host = ctx.fetch_chunk(groups) if groups is not None else K.Chunk(0)
            top = q.top(agg, groups=groups)  # ORDER BY revenue DESC, o_orderdate LIMIT 10 (+ gather of the ranks' top rows + final TopN)
            mark()
            return top, n_groups, q.launches(), groups_host
        finally:
            q.close()
    q = P.SsbQ41(ctx, replicate=comm.all_gather if comm is not None else None).build(tb)
    agg = op2 = top = None
    try:
        mark()
        stage_rows = []
        agg = q.probe(tb["lineorder"], phase=K.AGG_UPDATE_SERIALIZE if comm is not None else K.AGG_UPDATE_FINALIZE, stage_rows=stage_rows)
        out = agg.pull_all()
        launches = sum(o.counters()["kernel_launches"] for o in q.joins + q.keep + [agg]) + getattr(q, "scan_launches", 0)
        if comm is not None:  # <= 35 intermediate rows per rank: gathered and merged on every rank
            op2 = ops.ScanAggOperator(ctx, Q.ssb_q41_merge_plan())
            op2.push_chunk(comm.all_gather(out))
            op2.set_finishing()
            out = op2.pull_all()
            launches += op2.counters()["kernel_launches"]
        top = ops.TopNOperator(ctx, SSB_TOP[0], SSB_TOP[1], SSB_TOP[2], SSB_TOP[3])  # ORDER BY d_year, c_nation
        if out.n_rows:
            top.push_chunk(out)
        top.set_finishing()
        final = top.pull_all()
        launches += top.counters()["kernel_launches"]
        mark()
        if rows_out is not None:
            rows_out.update(zip((["after the scan's runtime filters"] if len(stage_rows) == 5 else []) + ["after supplier", "after customer", "after part", "after dates"],
                                stage_rows))
        return final, out.n_rows, launches, out
    finally:
        for o in (agg, op2, top):
            if o is not None:
                o.close()
        q.close()


def join_parity_sample(env, args, query: str) -> bool:
    """the multi-GPU plan (broadcast / shuffle / runtime-filter merge / gather + merge / final TopN) over a bounded sample split over the
    ranks, against the oracle's operator chain on rank 0: the groups as a set, the ORDER BY rows by their sort keys"""
    from starrocks_b200 import distributed as D, tpch_synth as T
    n_fact = max(1 << 20, args.parity_rows // 2)
    sizes = join_sizes(query, n_fact)
    cols, kw = join_cols_kw(query, sizes)
    tb, ptrs = {}, []
    for t, n in sizes.items():
        b, e = D.shard_range(n, env.rank, env.world)
        tb[t], pp = env.ctx.device_chunk(T.host_chunk(t, b, e - b, cols[t], **kw.get(t, {})))
        ptrs += pp
    try:
        top, _, _, groups = join_step(env, query, tb, want_groups=True)
    finally:
        for p in ptrs:
            env.ctx.device_free(p)
    parts = gather_objects(env, groups if query == "q3" else None)
    ok = True
    if env.rank == 0:
        want_groups, want_top = join_oracle_chain(_oracle(), query, join_host_tables(query, n_fact), env.cores)
        got_groups = D.concat_chunks([p for p in parts if p is not None and p.n_rows]) if query == "q3" else groups
        if query == "q3":  # ties on (revenue, o_orderdate) may be broken differently: compare the sort keys row by row, the groups as a set
            key = lambda rows: [(r[3], r[1]) for r in rows]
            ok = big_chunks_equal(got_groups, want_groups) and key(rows_key(top, False)) == key(rows_key(want_top, False))
        else:
            ok = rows_key(got_groups) == rows_key(want_groups) and rows_key(top, False) == rows_key(want_top, False)
    return all_true(env, ok)


def bench_join(env, args, query: str, steps: int, warmup: int):
    import ctypes as C
    from starrocks_b200 import capi as K, distributed as D, ops, pipelines as P, tpch_synth as T
    torch = env.torch
    world, rank, ctx = env.world, env.rank, env.ctx
    n_fact = args.rows or SF100_LINEITEM
    sizes = join_sizes(query, n_fact)
    cols, kw = join_cols_kw(query, sizes)
    # resident shards: this rank's contiguous row range of every table (strong scaling: the SF100 tables are split over the ranks)
    tables, keep = {}, []
    for t, n in sizes.items():
        b, e = D.shard_range(n, rank, world)
        tables[t], k = T.device_chunk(t, b, e - b, cols[t], env.dev, **kw.get(t, {}))
        keep.append(k)
    torch.cuda.synchronize()
    for _ in range(warmup):
        join_step(env, query, tables)
    sampler = ClockSampler(env.local)
    barrier(env)
    sampler.start()
    t0 = time.perf_counter()
    launches = 0
    for _ in range(steps):
        top, n_groups, l, _ = join_step(env, query, tables)
        launches += l
    barrier(env)
    dt = time.perf_counter() - t0
    clocks = sampler.finish()
    dt_max = max_over_ranks(env, [dt])[0]
    if query == "q3":
        n_groups = sum_over_ranks(env, n_groups)
    marks, stage_rows = [], {}
    join_step(env, query, tables, marks=marks, rows_out=stage_rows)  # one extra pass, synchronised after every stage, for the breakdown (outside the timed region)
    names = P.TpchQ3.STAGES + ["pull + TopN"] if query == "q3" else ["filter (+ broadcast) + build 4 dimensions", "4 probes + group by (+ gather, merge) + ORDER BY"]
    alg = join_alg_bytes(query, sizes)
    achieved = alg * steps / dt_max / 1e9 / world
    result = {
        "metric": JOIN_METRIC[query], "value": n_fact * steps / dt_max, "unit": "rows/s", "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": dt_max / steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int128" if query == "q3" else "int64",
        "data": "synthetic",
        "config": dict(base_config(JOIN_NAME[query], sizes, 24, world, "strong"),
                       exchange=("broadcast customer build side, hash shuffle (peer-memory scatter over NVLink) of both sides of the orders join, merged runtime filter, "
                                 "gather of the ranks' top rows" if query == "q3" else "dimension build sides all-gathered (broadcast join), intermediate rows gathered + merged") if world > 1 else "none (1 GPU)"),
        "gpu_launches": launches, "result_rows": n_groups, "top_rows": top.n_rows, "rows_between_operators_rank0": stage_rows,
        "stages_ms_rank0": dict(zip(names, [(marks[i + 1] - marks[i]) * 1e3 for i in range(len(marks) - 1)])),
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": env.peak, "unit": "GB/s", "frac": achieved / env.peak, "traffic": None, "peak_source": env.peak_source,
                     "note": "whole pipeline per GPU: algorithmic input bytes of this rank's shards (SURVEY 8d) over the step time; the joins are random-access "
                             "bound, not streaming (per-kernel figures: DESIGN.md section 4, profiles/)"},
        "clocks": clocks,
    }
    # ---- e2e: the same pipeline fed from pinned HOST chunks of a bounded slice, result pulled to the host --------------------------
    if not args.no_e2e:
        n_e = int(min(args.e2e_rows, n_fact // world))
        es = join_sizes(query, n_e)
        ecols, ekw = join_cols_kw(query, es)
        host, pins, h2d = {}, [], 0
        try:
            from starrocks_b200 import pages as PG
            paged = {"q3": ("orders", "lineitem"), "ssb": ("lineorder",)}[query]  # the big tables arrive as encoded pages, the dimensions as chunks
            for tname, n in es.items():
                dch, k = T.device_chunk(tname, 0, n, ecols[tname], env.dev, **ekw.get(tname, {}))
                if tname in paged and not os.environ.get("SRB_E2E_NO_PAGES"):
                    pcols = {}
                    for cname in ecols[tname]:
                        slot, t, pr, sc = T._SLOT[cname]
                        enc, pages = PG.for_encode_column(k[cname], 1 << 16)
                        nb = int(enc.numel())
                        p = C.c_void_p()
                        ops._check(ctx.L.srb_pinned_alloc(ctx.h, nb + 64, C.byref(p)))
                        ops._check(ctx.L.srb_memcpy_d2h(ctx.h, p, enc.data_ptr(), nb))
                        pins.append(p)
                        pcols[slot] = PG.PageColumn(t, p.value, nb, pages, PG.ENC_FOR, pr, sc, K.MEM_PINNED)
                        h2d += nb
                        del enc
                    host[tname] = PG.PageTable(pcols)
                    del dch, k
                    continue
                hc = K.Chunk(n)
                for slot, c in dch.columns.items():
                    wb = K.type_width(c.type)
                    p = C.c_void_p()
                    ops._check(ctx.L.srb_pinned_alloc(ctx.h, max(1, n * wb), C.byref(p)))
                    ops._check(ctx.L.srb_memcpy_d2h(ctx.h, p, c.data, n * wb))
                    pins.append(p)
                    hc.columns[slot] = K.Column(c.type, p.value, None, None, c.precision, c.scale, K.MEM_PINNED)
                    h2d += n * wb
                host[tname] = hc
                del dch, k
            for _ in range(2):  # the e2e slice runs the local plan on every rank (N independent host-fed pipelines)
                join_step(env, query, host, use_comm=False)
            barrier(env)
            esteps = max(2, min(steps, 5))
            t0 = time.perf_counter()
            for _ in range(esteps):
                e_top, _, _, _ = join_step(env, query, host, use_comm=False)
            barrier(env)
            dte = max_over_ranks(env, [time.perf_counter() - t0])[0]
        finally:
            for p in pins:
                ctx.L.srb_pinned_free(ctx.h, p)
        result["e2e"] = {"value": n_e * world * esteps / dte, "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": e_top.n_rows * 32,
                         "rows_per_step": n_e, "steps": esteps,
                         "note": "the fact tables (Q3: orders, lineitem; SSB: lineorder) pushed as frame-of-reference pages through srb_op_push_pages, the small "
                                 "tables as pinned host chunks through srb_op_push_chunk (H2D copies inside the timed region); the ORDER BY rows are read back" + (", the ~10^5..10^7 groups stay on the device" if query == "q3" else "")}
    del tables, keep
    free_device(env)
    if not args.no_parity:
        result["parity"] = join_parity_sample(env, args, query)
        result["parity_sample"] = (f"{max(1 << 20, args.parity_rows // 2)} fact rows (dimensions scaled) split over {world} rank(s), the same plan (exchange, runtime filter "
                                   "merge, TopN included) vs the oracle's operator chain")
    if rank == 0 and world == 1 and not args.no_cpu:
        result["cpu_baseline"] = cpu_join_baseline(query, args.cpu_seconds, env.cores)
    return result


# ---- main ----------------------------------------------------------------------------------------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    env = setup_env(args)
    env.gather_device = args.gather == "device"
    t_start = time.perf_counter()
    try:
        tensors = None
        if args.query in ("q3", "ssb"):
            result = bench_join(env, args, args.query, args.steps, args.warmup)
        else:
            result, tensors = bench_scan(env, args, args.query, args.steps, args.warmup, args.scaling, headline=True)
        if args.query == "q1" and not args.no_sub:
            queries = {}
            sub_w = min(args.warmup, 3)

            def sub(name, fn):
                t0 = time.perf_counter()
                try:
                    queries[name] = fn()
                except Exception as ex:  # a failing workload must not take the headline with it
                    queries[name] = {"error": f"{type(ex).__name__}: {ex}", "trace": traceback.format_exc(limit=4)}
                if isinstance(queries[name], dict):
                    queries[name]["wall_s"] = round(time.perf_counter() - t0, 1)

            if env.world > 1 and args.scaling == "weak":  # the other curve: SF100 split over the ranks
                sub("q1_strong", lambda: bench_scan(env, args, "q1", args.sub_steps, sub_w, "strong", tensors=tensors)[0])
            sub("q6", lambda: bench_scan(env, args, "q6", args.sub_steps, sub_w, args.scaling, tensors=tensors)[0])  # Q6 reads four of Q1's resident columns
            if env.world == 1:
                sub("q1_interpreter_variants", lambda: bench_q1_variants(env, args, tensors, 3, 2))
            tensors = None
            free_device(env)
            sub("highcard", lambda: bench_scan(env, args, "highcard", args.sub_steps, sub_w, args.scaling)[0])
            free_device(env)
            sub("q3", lambda: bench_join(env, args, "q3", args.sub_steps, sub_w))
            free_device(env)
            sub("ssb", lambda: bench_join(env, args, "ssb", args.sub_steps, sub_w))
            result["queries"] = queries
            result["parity_all"] = bool(result.get("parity", True)) and all(q.get("parity", False) for k, q in queries.items() if k != "q1_interpreter_variants")
        result["bench_wall_s"] = round(time.perf_counter() - t_start, 1)
        if env.rank == 0:
            print(json.dumps(result))
    finally:
        try:
            if env.comm is not None:
                env.part.close()
                env.comm.close()
            env.ctx.close()
            if env.world > 1:
                env.dist.destroy_process_group()
        except Exception:
            pass


if __name__ == "__main__":
    main()
