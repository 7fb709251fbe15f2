#!/usr/bin/env python
"""bench.py — TPC-H SF100 Q1 operator pipeline (scan -> filter -> low-cardinality GROUP BY with 8 DECIMAL aggregates)
through libsrb.so on N B200s, one process per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--query q1|q6|highcard|q3|ssb]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

A step = one pass of the hot path over the rank's resident lineitem shard (600,037,902 rows of the seven columns Q1
reads, 44 B/row with dictionary-coded keys): operator create -> push the device-resident range -> set_finishing -> pull
the 4 result rows. `value` = rows of all ranks / time, inputs already in HBM. `e2e` = the same pipeline fed from pinned
HOST buffers through the same C ABI (host->device copies and the result read inside the timed region) on a bounded
slice. `roofline` = algorithmic bytes of the dominant kernel / its CUDA-event duration vs the measured HBM copy peak.
`cpu_baseline` / `--impl reference` = the CPU restatement of the reference's pipeline (oracle/, the real BE cannot be
built here) on the box's host cores over a bounded sample of the same table.
Shards scale weakly (every rank scans its own 600 M-row range of the synthetic table); the only exchange of Q1 is the
gather of <= 4 intermediate state rows per rank, merged by a phase-2 (merge-finalize) operator on every rank INSIDE the
timed step. `--query highcard` at N > 1 runs phase 1 -> device hash partition -> NCCL all-to-all -> phase 2.
`--query q3` / `--query ssb` run the join pipelines of starrocks_b200/pipelines.py over the SF100 tables, range-partitioned
over the ranks (strong scaling: the configs are "SF100 ... across 8 B200"): Q3 as a shuffle join (customer build side
all-gathered, both sides of the orders join hash-partitioned and exchanged all-to-all), SSB Q4.1 as a broadcast join
(filtered dimension slices all-gathered, <= 35 intermediate rows gathered and merged).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SF100_LINEITEM = 600_037_902
HIGHCARD_ROWS, HIGHCARD_KEYS = 600_000_000, 100_000_000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="srb")
    ap.add_argument("--query", default="q1", choices=["q1", "q6", "highcard", "q3", "ssb"])
    ap.add_argument("--rows", type=int, default=0, help="rows per rank (default: the SF100 table)")
    ap.add_argument("--e2e-rows", type=int, default=64 << 20)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--phase-times", action="store_true", help="add the per-stage host times of a step to the JSON line")
    ap.add_argument("--gather", choices=["nccl", "gloo", "torch"], default="nccl", help="N > 1, low-cardinality plans: backend of the intermediate-row all-gather")
    return ap.parse_args()


def workload(query: str):
    from starrocks_b200 import queries as Q, tpch_synth as T
    if query == "q1":
        return dict(name="TPC-H SF100 Q1: lineitem filter + GROUP BY (4 groups) + 8 DECIMAL aggregates", table="lineitem", cols=T.LINEITEM_Q1,
                    plan=Q.q1_plan, bytes_per_row=44, rows=SF100_LINEITEM, kw={})
    if query == "q6":
        return dict(name="TPC-H SF100 Q6: lineitem range filter + SUM(DECIMAL128)", table="lineitem", cols=T.LINEITEM_Q6, plan=Q.q6_plan,
                    bytes_per_row=28, rows=SF100_LINEITEM, kw={})
    return dict(name="GROUP BY 100M distinct BIGINT keys: SUM, COUNT over 600M rows", table="highcard", cols=["k", "v"],
                plan=lambda **kw: Q.highcard_plan(group_hint=HIGHCARD_KEYS, **kw), bytes_per_row=16, rows=HIGHCARD_ROWS, kw={"n_keys": HIGHCARD_KEYS})


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

    def result(self):
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max, "reasons": sorted(self.reasons)}


# ---- CPU restatement of the reference pipeline on a bounded sample ------------------------------------------------------------
def cpu_baseline(w, seconds: float, threads: int):
    """Times oracle/ (kind "port") with `threads` host threads on a sample sized to take about `seconds`."""
    from starrocks_b200 import tpch_synth as T
    from tests.oracle_lib import Oracle
    orc = Oracle()
    probe_rows = 1 << 20
    ch = T.host_chunk(w["table"], 0, probe_rows, w["cols"], **w["kw"])
    t0 = time.perf_counter()
    orc.scan_agg(w["plan"](), ch, n_threads=threads)
    rate = probe_rows / (time.perf_counter() - t0)
    rows = int(min(max(rate * seconds, probe_rows), 64 << 20))
    ch = T.host_chunk(w["table"], 0, rows, w["cols"], **w["kw"])
    best = None
    for _ in range(2):
        t0 = time.perf_counter()
        out = orc.scan_agg(w["plan"](), ch, n_threads=threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return {"value": rows / best, "unit": "rows/s", "cores": threads, "kind": "port",
            "sample": f"first {rows} rows of the same synthetic table, 4096-row chunks, {threads} threads with private tables + merge, best of 2"}, out, ch


def run_reference(args):
    """--impl reference: the reference's CPU pipeline for this path (oracle port; the BE itself cannot be built here)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    per_step = max(4.0, min(20.0, 100.0 / max(1, args.steps + args.warmup)))
    if args.query in ("q3", "ssb"):
        name, run, rows = join_cpu_sample(args.query, per_step, threads)
        for _ in range(args.warmup):
            run()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            run()
        dt = time.perf_counter() - t0
        value = rows * args.steps / dt
        cb = {"value": value, "unit": "rows/s", "cores": threads, "kind": "port",
              "sample": f"each step: the first {rows} fact-table rows (dimension tables scaled with them), filters on {threads} threads, joins on 1"}
        print(json.dumps({"impl": "reference", "metric": JOIN_METRIC[args.query], "value": value, "unit": "rows/s", "n_gpus": args.gpus,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
                          "vs_baseline": None, "dtype": "int128", "data": "synthetic", "config": {"workload": name, "rows_per_step": rows},
                          "cpu_baseline": cb, "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    w = workload(args.query)
    from starrocks_b200 import tpch_synth as T
    from tests.oracle_lib import Oracle
    orc = Oracle()
    probe_rows = 1 << 20
    ch = T.host_chunk(w["table"], 0, probe_rows, w["cols"], **w["kw"])
    t0 = time.perf_counter()
    orc.scan_agg(w["plan"](), ch, n_threads=threads)
    rate = probe_rows / (time.perf_counter() - t0)
    rows = int(min(max(rate * per_step, probe_rows), 64 << 20))
    ch = T.host_chunk(w["table"], 0, rows, w["cols"], **w["kw"])
    for _ in range(args.warmup):
        orc.scan_agg(w["plan"](), ch, n_threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        orc.scan_agg(w["plan"](), ch, n_threads=threads)
    dt = time.perf_counter() - t0
    value = rows * args.steps / dt
    cb = {"value": value, "unit": "rows/s", "cores": threads, "kind": "port",
          "sample": f"each step: first {rows} rows of the same synthetic table, 4096-row chunks, {threads} threads"}
    print(json.dumps({"impl": "reference", "metric": f"TPC-H SF100 {args.query.upper()} rows/s" if args.query != "highcard" else "high-cardinality GROUP BY rows/s", "value": value, "unit": "rows/s", "n_gpus": args.gpus,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "int128", "data": "synthetic", "config": {"workload": w["name"], "rows_per_step": rows},
                      "cpu_baseline": cb, "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ---- the join pipelines (Q3 shuffle join, SSB Q4.1 broadcast star join) ---------------------------------------------------------
JOIN_METRIC = {"q3": "TPC-H SF100 Q3 rows/s (lineitem rows)", "ssb": "SSB SF100 Q4.1 rows/s (lineorder rows)"}
JOIN_NAME = {"q3": "TPC-H SF100 Q3: 3 filters, customer |><| orders |><| lineitem hash joins (build/probe/gather), GROUP BY 3 keys",
             "ssb": "SSB SF100 Q4.1: 3 dimension filters, 4 hash joins (broadcast build sides), GROUP BY d_year, c_nation"}
Q3_COLS = {"customer": ["c_custkey", "c_mktsegment_code"], "orders": ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"],
           "lineitem": ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"]}


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


def join_cpu_sample(query: str, seconds: float, threads: int):
    """-> (workload name, run() -> result chunk, fact rows): the oracle chain on a sample sized to take about `seconds`"""
    from starrocks_b200 import pipelines as P, tpch_synth as T
    from tests.oracle_lib import Oracle
    orc = Oracle()

    def tables(n_fact):
        sizes = join_sizes(query, n_fact)
        cols, kw = join_cols_kw(query, sizes)
        return {t: T.host_chunk(t, 0, n, cols[t], **kw.get(t, {})) for t, n in sizes.items()}

    def run_on(tb):
        if query == "q3":
            return P.q3_oracle(orc, tb["customer"], tb["orders"], tb["lineitem"], n_threads=threads)
        return P.ssb_q41_oracle(orc, tb, tb["lineorder"], n_threads=threads)

    probe_rows = 1 << 20
    tb = tables(probe_rows)
    t0 = time.perf_counter()
    run_on(tb)
    rate = probe_rows / (time.perf_counter() - t0)
    rows = int(min(max(rate * seconds, probe_rows), 48 << 20))
    tb = tables(rows)
    return JOIN_NAME[query], (lambda: run_on(tb)), rows


def run_join_query(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from starrocks_b200 import capi as K, distributed as D, ops, pipelines as P, queries as Q, tpch_synth as T

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    query = args.query
    n_fact = args.rows or SF100_LINEITEM
    sizes = join_sizes(query, n_fact)
    cols, kw = join_cols_kw(query, sizes)
    ctx = ops.Context(device=local, rank=rank, world_size=world)
    # resident shards: this rank's contiguous row range of every table (strong scaling: the SF100 tables are split over the ranks)
    tables, keep = {}, []
    for t, n in sizes.items():
        b, e = D.shard_range(n, rank, world)
        tables[t], k = T.device_chunk(t, b, e - b, cols[t], dev, **kw.get(t, {}))
        keep.append(k)
    torch.cuda.synchronize()
    comm = part = None
    if world > 1:
        def put(b):
            dist.broadcast_object_list([b], src=0)

        def get():
            box = [None]
            dist.broadcast_object_list(box, src=0)
            return box[0]

        comm = ops.Communicator(ctx, put, get)
        part = ops.Partitioner(ctx, world)

    def step(tb, marks=None, rows_out=None):
        """one pass of the pipeline over the rank's shards -> (result rows on this rank, kernel launches)"""
        def mark():
            if marks is not None:
                torch.cuda.synchronize()
                marks.append(time.perf_counter())

        mark()
        if query == "q3":
            q = P.TpchQ3(ctx, comm, part)
            agg = q.run(tb["customer"], tb["orders"], tb["lineitem"], mark=mark if marks is not None else None, rows=rows_out)
            out = agg.pull_device()  # ~14 M groups stay on the device, partitioned by the order key; only the row count is read
            n_out = out.n_rows if out is not None else 0
            mark()
            launches = q.launches()
            q.close()
            return n_out, launches
        q = P.SsbQ41(ctx, replicate=comm.all_gather if comm is not None else None).build(tb)
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
            op2.close()
        mark()
        if rows_out is not None:
            rows_out.update(zip((["after the scan's runtime filters"] if len(stage_rows) == 5 else []) + ["after supplier", "after customer", "after part", "after dates"],
                                stage_rows))
        agg.close()
        q.close()
        return out.n_rows, launches

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step(tables)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    launches = 0
    for _ in range(args.steps):
        n_out, l = step(tables)
        launches += l
    barrier()
    dt = time.perf_counter() - t0
    sampler.stop_flag = True
    sampler.join(timeout=2)
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    tot = torch.tensor([n_out], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if query == "q3":
            dist.all_reduce(tot)
    dt_max = t.item()
    marks, stage_rows = [], {}
    step(tables, marks, stage_rows)  # one extra pass, synchronised after every stage, for the breakdown (outside the timed region)
    names = P.TpchQ3.STAGES + ["pull"] if query == "q3" else ["filter (+ broadcast) + build 4 dimensions", "4 probes + group by (+ gather, merge)"]
    result = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        alg = join_alg_bytes(query, sizes)
        achieved = alg * args.steps / dt_max / 1e9 / world
        result = {
            "metric": JOIN_METRIC[query], "value": n_fact * args.steps / dt_max, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt_max / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "int128" if query == "q3" else "int64", "data": "synthetic",
            "config": {"workload": JOIN_NAME[query], "rows": sizes, "l2": "inputs (>= 14 GB per step over all ranks) exceed the 126 MB L2",
                       "parallelism": (f"range-partitioned scans x{world}; " + ("broadcast customer build side, hash shuffle (NCCL all-to-all) of both sides of the orders join"
                                       if query == "q3" else "dimension build sides all-gathered (broadcast join), intermediate rows gathered + merged")) if world > 1 else "1 GPU"},
            "gpu_launches": launches, "result_rows": int(tot.item()), "rows_between_operators_rank0": stage_rows,
            "stages_ms_rank0": dict(zip(names, [(marks[i + 1] - marks[i]) * 1e3 for i in range(len(marks) - 1)])),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "note": "whole pipeline per GPU: algorithmic input bytes of this rank's shards (SURVEY 8d) over the step time; "
                                 "the joins are random-access bound, not streaming"},
            "clocks": sampler.result(),
        }
    # ---- e2e: the same pipeline fed from pinned HOST chunks of a bounded slice, result pulled to the host --------------------------
    if not args.no_e2e:
        n_e = int(min(args.e2e_rows, n_fact // world))
        es = join_sizes(query, n_e)
        ecols, ekw = join_cols_kw(query, es)
        host, pins, h2d = {}, [], 0
        for tname, n in es.items():
            dch, k = T.device_chunk(tname, 0, n, ecols[tname], dev, **ekw.get(tname, {}))
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
        e_comm, e_part = comm, part
        comm = part = None  # the e2e slice runs the local plan on every rank (N independent host-fed pipelines)
        for _ in range(2):
            step(host)
        barrier()
        esteps = max(2, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(esteps):
            e_out, _ = step(host)
        barrier()
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        comm, part = e_comm, e_part
        for p in pins:
            ctx.L.srb_pinned_free(ctx.h, p)
        if rank == 0:
            result["e2e"] = {"value": n_e * world * esteps / t.item(), "unit": "rows/s", "h2d_bytes_per_step": h2d,
                             "d2h_bytes_per_step": 8 if query == "q3" else e_out * 16, "rows_per_step": n_e, "steps": esteps,
                             "note": "every table pushed as pinned host chunks through srb_op_push_chunk / srb_join_probe_push_chunk (staged H2D copies inside "
                                     "the timed region)" + ("; the ~10^5..10^7 result groups stay on the device, the row count is read" if query == "q3" else "")}
    if rank == 0 and world == 1 and not args.no_cpu:
        name, run, rows = join_cpu_sample(query, args.cpu_seconds, os.cpu_count() or 1)
        t0 = time.perf_counter()
        cpu_out = run()
        dtc = time.perf_counter() - t0
        result["cpu_baseline"] = {"value": rows / dtc, "unit": "rows/s", "cores": os.cpu_count() or 1, "kind": "port",
                                  "sample": f"the first {rows} fact-table rows (dimension tables scaled with them) through the oracle chain: filters on all threads, joins on 1"}
        # parity of the GPU pipeline on the same sample
        ss = join_sizes(query, rows)
        scols, skw = join_cols_kw(query, ss)
        tb = {tn: T.host_chunk(tn, 0, n, scols[tn], **skw.get(tn, {})) for tn, n in ss.items()}
        if query == "q3":
            q = P.TpchQ3(ctx)
            got = q.run(tb["customer"], tb["orders"], tb["lineitem"]).pull_all()
            q.close()
        else:
            q = P.SsbQ41(ctx).build(tb)
            agg = q.probe(tb["lineorder"])
            got = agg.pull_all()
            agg.close()
            q.close()
        result["parity_on_cpu_sample"] = K.chunk_rows(got) == K.chunk_rows(cpu_out)
    if rank == 0:
        print(json.dumps(result))
    if comm is not None:
        part.close()
        comm.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


# ---- the GPU arm ----------------------------------------------------------------------------------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
        return
    if args.query in ("q3", "ssb"):
        run_join_query(args)
        return
    import torch
    import torch.distributed as dist
    from starrocks_b200 import capi as K, distributed as D, ops, tpch_synth as T

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    w = workload(args.query)
    rows = args.rows or w["rows"]
    ctx = ops.Context(device=local, rank=rank, world_size=world)

    # resident shard: rows [rank*rows, (rank+1)*rows) of the synthetic table (weak scaling)
    chunk, keep = T.device_chunk(w["table"], rank * rows, rows, w["cols"], dev, **w["kw"])
    torch.cuda.synchronize()

    comm = part = None
    if world > 1:
        # libsrb's own NCCL communicator moves the intermediate rows; torch.distributed only carries the 128-byte NCCL id
        def put(b):
            dist.broadcast_object_list([b], src=0)

        def get():
            box = [None]
            dist.broadcast_object_list(box, src=0)
            return box[0]

        comm = ops.Communicator(ctx, put, get)
        part = ops.Partitioner(ctx, world)
        # the <= 4 intermediate rows of a low-cardinality plan: one fixed-slot ncclAllGather inside the library (srb_comm_all_gather on
        # host chunks); --gather gloo / torch route them through torch.distributed object collectives instead
        host_ex = comm if args.gather == "nccl" else D.ObjectExchange(dist, rank, world, group=dist.new_group(backend="gloo") if args.gather == "gloo" else None)

    phase_ms = [] if args.phase_times else None  # host wall time between the step's stages (rank 0; a diagnostic, not the metric)

    def step(collect=None):
        """one pass of the hot path over this rank's shard; at N > 1 it includes the exchange and the phase-2 merge
        (starrocks_b200/distributed.py: gather for the low-cardinality plans, hash shuffle for the high-cardinality one)"""
        from starrocks_b200 import queries as Q
        two_phase = world > 1 and args.query in ("q1", "highcard")
        marks = [time.perf_counter()] if phase_ms is not None else None

        def mark():
            if marks is not None:
                marks.append(time.perf_counter())

        plan = w["plan"](phase=K.AGG_UPDATE_SERIALIZE) if two_phase else w["plan"]()
        op = ops.ScanAggOperator(ctx, plan)
        op.push_chunk(chunk)
        op.set_finishing()
        mark()
        c, p = None, None
        op2 = None
        if two_phase:
            if args.query == "q1":
                # <= 4 intermediate rows (AVG states are {sum,count} blobs): a host object gather is the whole exchange
                inter = op.pull_all()
                mark()
                recv = host_ex.all_gather(inter)
                mark()
                op2 = ops.ScanAggOperator(ctx, Q.q1_merge_plan(inter))
            else:
                inter = op.pull_device()
                mark()
                if os.environ.get("SRB_NO_P2P"):  # A/B: device partition, then NCCL send/recv of the grouped rows
                    grouped, counts = part.run(inter, [0])
                    recv, _ = comm.all_to_all(grouped, counts)
                else:  # partition scatter fused with the exchange: rows go straight into the peers' receive buffers
                    recv, _ = comm.shuffle(inter, [0])
                mark()
                op2 = ops.ScanAggOperator(ctx, Q.highcard_merge_plan(group_hint=HIGHCARD_KEYS // world))
            op2.push_chunk(recv)
            op2.set_finishing()
            mark()
            out = op2.pull_all() if args.query == "q1" else op2.pull_device()
        else:
            out = op.pull_all() if args.query != "highcard" else op.pull_device()
        mark()
        if marks is not None:
            phase_ms.append([(marks[i + 1] - marks[i]) * 1e3 for i in range(len(marks) - 1)])
        n_out = out.n_rows
        c, p = op.counters(), op.profile()
        if op2 is not None:
            c2 = op2.counters()
            c = dict(c, kernel_launches=c["kernel_launches"] + c2["kernel_launches"])
        if collect is not None:
            collect.append((c, p))
        if args.query == "highcard":
            out = None  # ~10^8 result rows stay on the device (24 B/row); only the row count is read
        op.close()
        if op2 is not None:
            op2.close()
        return out, n_out

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    stats = []
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out, n_out = step(stats)
    barrier()
    dt = time.perf_counter() - t0
    sampler.stop_flag = True
    sampler.join(timeout=2)
    # device-side time of the timed steps: CUDA events recorded by the library on the operator's own stream
    kernel_ms = sum(p["kernel_ms"] for _, p in stats)
    batches = sum(p["batches"] for _, p in stats)
    launches = sum(c["kernel_launches"] for c, _ in stats)
    t = torch.tensor([dt, kernel_ms / 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt_max, kernel_s_max = t.tolist()

    final = out  # at N > 1 every rank already holds the merged Q1 result (gather + phase 2 ran inside the step)
    if world > 1 and args.query == "highcard":
        tot = torch.tensor([n_out], dtype=torch.int64, device=dev)
        dist.all_reduce(tot)
        n_out = int(tot.item())  # groups over all ranks (every group lives on exactly one rank after the shuffle)

    result = None
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        per_batch_rows = rows * args.steps / max(1, batches)
        achieved = w["bytes_per_row"] * per_batch_rows / (kernel_s_max / max(1, batches)) / 1e9 if kernel_s_max > 0 else 0.0
        # DRAM bytes per launch from the committed `ncu --set full` capture of this kernel (profiles/traffic.json, written by
        # tools/ncu_summary.py): measured bytes per row of the captured launch x the rows of an average launch here
        traffic, traffic_src = None, None
        try:
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.query)
            if tj and tj.get("rows_per_launch"):
                traffic = tj["dram_bytes_per_launch"] / tj["rows_per_launch"] * per_batch_rows
                traffic_src = f"{tj['report']}: {tj['dram_bytes_per_launch'] / tj['rows_per_launch']:.2f} B/row x rows per launch"
        except Exception:
            pass
        result = {
            "metric": f"TPC-H SF100 {args.query.upper()} rows/s" if args.query != "highcard" else "high-cardinality GROUP BY rows/s",
            "value": rows * world * args.steps / dt_max, "unit": "rows/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt_max / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int128",
            "data": "synthetic",
            "config": {"workload": w["name"], "rows_per_gpu": rows, "bytes_per_row": w["bytes_per_row"], "l2": "inputs (>= 16 GB per step) exceed the 126 MB L2",
                       "parallelism": f"range-partitioned x{world}", "kernel": ops_shape_name(stats)},
            "scan_gb_per_s": rows * world * args.steps * w["bytes_per_row"] / dt_max / 1e9,
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak if peak else None,
                         "traffic": traffic, "traffic_source": traffic_src, "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650",
                         "launch": f"{int(per_batch_rows)} rows x {w['bytes_per_row']} B per launch, {kernel_s_max / max(1, batches) * 1e3:.3f} ms avg (CUDA events on the op stream)"},
            "clocks": sampler.result(),
            "result_rows": n_out,
        }
        if phase_ms:
            timed = phase_ms[-args.steps:]
            result["phase_ms_rank0"] = [sum(t[i] for t in timed) / len(timed) for i in range(len(timed[0]))]
    # ---- e2e: the same pipeline fed from pinned host buffers through the C ABI ------------------------------------------------
    if not args.no_e2e:
        e2e = run_e2e(args, ctx, w, chunk, rows, dev, rank, world)
        if rank == 0:
            result["e2e"] = e2e
    # ---- CPU baseline + parity of the GPU result on the sampled range (rank 0, N = 1) ----------------------------------------------
    if rank == 0 and world == 1 and not args.no_cpu:
        cb, cpu_out, cpu_chunk = cpu_baseline(w, args.cpu_seconds, os.cpu_count() or 1)
        result["cpu_baseline"] = cb
        dch, ptrs = ctx.device_chunk(cpu_chunk)
        got = ops.run_scan_agg(ctx, w["plan"](), dch)
        for p in ptrs:
            ctx.device_free(p)
        result["parity_on_cpu_sample"] = K.chunk_rows(got) == K.chunk_rows(cpu_out)
    if rank == 0:
        print(json.dumps(result))
    if comm is not None:
        part.close()
        comm.close()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def ops_shape_name(stats):
    shapes = {p["shape"] for _, p in stats}
    static = sum(p["static_launches"] for _, p in stats)
    return f"static shape {sorted(shapes)} (TMA-staged), {static} static launches" if static else "generic interpreter"


def run_e2e(args, ctx, w, chunk, rows, dev, rank, world):
    """Host buffers -> C ABI -> result on the host. Pinned host copies of a slice of the shard are pushed as SRB_MEM_PINNED chunks;
    the library stages them into HBM in slices overlapping copies with the kernel; the 4 result rows are read back."""
    import ctypes as C
    import torch
    import torch.distributed as dist
    from starrocks_b200 import capi as K, ops
    n = int(min(args.e2e_rows, rows))
    host = K.Chunk(n)
    pins = []
    h2d = 0
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
        out = op.pull_all() if args.query != "highcard" else op.pull_device()
        op.close()
        return out

    for _ in range(max(1, args.warmup - 1)):
        out = step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    steps = max(2, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(steps):
        out = step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    d2h = sum(len(K.column_pyvalues(c)) * max(K.type_width(c.type), 24) for c in out.columns.values()) if args.query != "highcard" else 8
    for p in pins:
        ctx.L.srb_pinned_free(ctx.h, p)
    return {"value": n * world * steps / t.item(), "unit": "rows/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "rows_per_step": n,
            "steps": steps, "note": "pinned host columns pushed through srb_op_push_chunk; staged H2D in 4 Mi-row slices overlapped with the kernel"}


if __name__ == "__main__":
    main()
