"""Host-side breakdown of one bench step (operator create / push / set_finishing / pull / close), in milliseconds.

    python tools/step_profile.py --query q1 [--rows N] [--steps K]
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--query", default="q1")
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--two-phase", action="store_true", help="q1 only: phase 1 -> pull -> phase-2 merge operator (the N > 1 step without the gather)")
    args = ap.parse_args()
    import torch
    import bench
    from starrocks_b200 import ops, tpch_synth as T

    w = bench.workload(args.query)
    rows = args.rows or w["rows"]
    dev = torch.device("cuda", 0)
    ctx = ops.Context(device=0)
    chunk, keep = T.device_chunk(w["table"], 0, rows, w["cols"], dev, **w["kw"])
    torch.cuda.synchronize()
    if args.two_phase:
        two_phase(args, ctx, w, chunk)
        return
    names = ["plan", "create", "push", "finish", "pull", "close"]
    acc = {n: [] for n in names}
    kern = []
    for it in range(args.steps + 2):
        t = [time.perf_counter()]
        plan = w["plan"]()
        t.append(time.perf_counter())
        op = ops.ScanAggOperator(ctx, plan)
        t.append(time.perf_counter())
        op.push_chunk(chunk)
        t.append(time.perf_counter())
        op.set_finishing()
        t.append(time.perf_counter())
        out = op.pull_all() if args.query != "highcard" else op.pull_device()
        t.append(time.perf_counter())
        p = op.profile()
        op.close()
        t.append(time.perf_counter())
        if it >= 2:
            for i, n in enumerate(names):
                acc[n].append((t[i + 1] - t[i]) * 1e3)
            kern.append(p["kernel_ms"])
    for n in names:
        v = sorted(acc[n])
        print(f"{n:8s} median {v[len(v) // 2]:8.3f} ms   min {v[0]:8.3f}   max {v[-1]:8.3f}")
    v = sorted(kern)
    print(f"kernel_ms (device events, inside push) median {v[len(v) // 2]:8.3f}")


def two_phase(args, ctx, w, chunk):
    from starrocks_b200 import capi as K, ops, queries as Q
    names = ["phase1", "pull_all", "merge_plan", "merge_create", "merge_push", "merge_finish", "merge_pull", "close"]
    acc = {n: [] for n in names}
    for it in range(args.steps + 2):
        t = [time.perf_counter()]
        op = ops.ScanAggOperator(ctx, w["plan"](phase=K.AGG_UPDATE_SERIALIZE))
        op.push_chunk(chunk)
        op.set_finishing()
        t.append(time.perf_counter())
        inter = op.pull_all()
        t.append(time.perf_counter())
        plan = Q.q1_merge_plan(inter)
        t.append(time.perf_counter())
        op2 = ops.ScanAggOperator(ctx, plan)
        t.append(time.perf_counter())
        op2.push_chunk(inter)
        t.append(time.perf_counter())
        op2.set_finishing()
        t.append(time.perf_counter())
        op2.pull_all()
        t.append(time.perf_counter())
        op.close()
        op2.close()
        t.append(time.perf_counter())
        if it >= 2:
            for i, n in enumerate(names):
                acc[n].append((t[i + 1] - t[i]) * 1e3)
    for n in names:
        v = sorted(acc[n])
        print(f"{n:13s} median {v[len(v) // 2]:8.3f} ms   min {v[0]:8.3f}   max {v[-1]:8.3f}")


if __name__ == "__main__":
    main()
