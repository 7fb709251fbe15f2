"""TPC-H Q3 operator pipeline on one B200: filters -> hash join build/probe x2 -> group by, every
operator output handed to the next as a device chunk.

    python tools/bench_q3.py [--rows LINEITEM_ROWS] [--steps K] [--warmup W]

customer (rows/40) --filter mktsegment--> build J1
orders   (rows/4)  --filter orderdate---> probe J1 --> build J2
lineitem (rows)    --filter shipdate----> probe J2 --> group by (l_orderkey, o_orderdate, o_shippriority): sum(ext * (1 - disc))

Prints one JSON line: rows/s over lineitem rows, per-stage wall times (each stage synchronised) and the algorithmic bytes
(customer 5 B, orders 16 B, lineitem 24 B per row).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=600_037_902)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    args = ap.parse_args()
    import torch
    from starrocks_b200 import ops, queries as Q, tpch_synth as T

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    ctx = ops.Context(device=0)
    n_l, n_o, n_c = args.rows, args.rows // 4, max(30, args.rows // 40)
    cust, k1 = T.device_chunk("customer", 0, n_c, ["c_custkey", "c_mktsegment_code"], dev)
    orders, k2 = T.device_chunk("orders", 0, n_o, ["o_orderkey", "o_custkey", "o_orderdate", "o_shippriority"], dev, n_customers=n_c)
    line, k3 = T.device_chunk("lineitem", 0, n_l, ["l_orderkey", "l_extendedprice", "l_discount", "l_shipdate"], dev)
    torch.cuda.synchronize()

    def filt(plan, ch):
        op = ops.ScanAggOperator(ctx, plan)
        op.push_chunk(ch)
        op.set_finishing()
        return op, op.pull_device()

    def step(times=None):
        marks = [time.perf_counter()]

        def mark():
            if times is not None:
                torch.cuda.synchronize()
                marks.append(time.perf_counter())

        fc, dc = filt(Q.q3_customer_filter(), cust)
        j1 = ops.HashJoinOperator(ctx, Q.q3_join_customer_orders())
        j1.push_chunk(dc)
        j1.set_finishing()
        mark()
        fo, do = filt(Q.q3_orders_filter(), orders)
        mark()
        j1.probe.push_chunk(do)
        d1 = j1.probe.pull_device()
        j2 = ops.HashJoinOperator(ctx, Q.q3_join_orders_lineitem())
        j2.push_chunk(d1)
        j2.set_finishing()
        mark()
        fl, dl = filt(Q.q3_lineitem_filter(), line)
        mark()
        j2.probe.push_chunk(dl)
        d2 = j2.probe.pull_device()
        mark()
        agg = ops.ScanAggOperator(ctx, Q.q3_agg_plan(group_hint=max(1024, d2.n_rows // 2)))
        agg.push_chunk(d2)
        agg.set_finishing()
        out = agg.pull_device()
        mark()
        rows = dict(customer_kept=dc.n_rows, orders_kept=do.n_rows, j1_out=d1.n_rows, lineitem_kept=dl.n_rows, j2_out=d2.n_rows, groups=out.n_rows)
        launches = sum(o.counters()["kernel_launches"] for o in (fc, fo, fl, j1, j2, agg))
        for o in (fc, fo, fl, j1, j2, agg):
            o.close()
        if times is not None:
            times.append([(marks[i + 1] - marks[i]) * 1e3 for i in range(len(marks) - 1)])
        return rows, launches

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        rows, launches = step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.steps
    stage = []
    step(stage)  # one extra, synchronised pass for the per-stage breakdown (not part of the timed region)
    names = ["customer filter + build J1", "orders filter", "probe J1 + build J2", "lineitem filter", "probe J2 + gather", "group by"]
    alg_bytes = n_c * 5 + n_o * 16 + n_l * 24
    peak = 6450.3
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    print(json.dumps({
        "metric": "TPC-H Q3 operator pipeline rows/s (lineitem rows)", "value": n_l / dt, "unit": "rows/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "dtype": "int128", "data": "synthetic",
        "config": {"workload": "TPC-H Q3: 3 filters, 2 hash joins (build/probe/gather), GROUP BY 3 keys", "lineitem_rows": n_l, "orders_rows": n_o,
                   "customer_rows": n_c},
        "rows": rows, "gpu_launches": launches,
        "stages_ms": dict(zip(names, stage[0])),
        "roofline": {"bound": "hbm", "achieved": alg_bytes / dt / 1e9, "peak": peak, "unit": "GB/s", "frac": alg_bytes / dt / 1e9 / peak,
                     "note": "whole pipeline: algorithmic input bytes (5 / 16 / 24 B per customer / orders / lineitem row) over the step's wall time"},
    }))
    ctx.close()


if __name__ == "__main__":
    main()
