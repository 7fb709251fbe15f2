"""Phase 2 of the two-phase high-cardinality GROUP BY on one GPU: the intermediate tuples of a phase-1 operator are merged by a
phase-2 operator (what a rank does with the tuples it receives from the shuffle). Prints the host times of the merge step.

    python tools/merge_profile.py [--rows N] [--keys K] [--steps S]
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=600_000_000)
    ap.add_argument("--keys", type=int, default=100_000_000)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    import torch
    from starrocks_b200 import capi as K, ops, queries as Q, tpch_synth as T

    dev = torch.device("cuda", 0)
    ctx = ops.Context(device=0)
    chunk, keep = T.device_chunk("highcard", 0, args.rows, ["k", "v"], dev, n_keys=args.keys)
    op1 = ops.ScanAggOperator(ctx, Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE, group_hint=args.keys))
    op1.push_chunk(chunk)
    op1.set_finishing()
    inter = op1.pull_device()
    print("intermediate rows", inter.n_rows)
    for it in range(args.steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        op2 = ops.ScanAggOperator(ctx, Q.highcard_merge_plan(group_hint=args.keys))
        t1 = time.perf_counter()
        op2.push_chunk(inter)
        t2 = time.perf_counter()
        op2.set_finishing()
        t3 = time.perf_counter()
        out = op2.pull_device()
        t4 = time.perf_counter()
        p, c = op2.profile(), op2.counters()
        op2.close()
        print(f"create {(t1 - t0) * 1e3:.2f} push {(t2 - t1) * 1e3:.2f} finish {(t3 - t2) * 1e3:.2f} pull {(t4 - t3) * 1e3:.2f} ms; kernel_ms {p['kernel_ms']:.2f} "
              f"shape {p['shape']} static launches {p['static_launches']} launches {c['kernel_launches']} groups {out.n_rows}")
    op1.close()
    ctx.close()


if __name__ == "__main__":
    main()
