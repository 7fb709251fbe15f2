"""SSB Q4.1 operator pipeline on one B200: broadcast star join + GROUP BY, device chunks operator to operator.

    python tools/bench_ssb.py [--rows LINEORDER_ROWS] [--steps K] [--warmup W]

A step = filter + build of the four dimensions (customer rows/200, supplier rows/3000, part rows/430, dates 2556: the SF100
proportions) and one pass of lineorder (24 B/row: six INT columns) through the four probes and the aggregation.
Prints one JSON line: lineorder rows/s, per-stage times, rows after every probe.
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
    from starrocks_b200 import ops, pipelines as P, tpch_synth as T

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    ctx = ops.Context(device=0)
    n_lo = args.rows
    n_c, n_s, n_p = max(100, n_lo // 200), max(50, n_lo // 3000), max(100, n_lo // 430)
    kw = dict(n_customers=n_c, n_suppliers=n_s, n_parts=n_p)
    keep, dims = [], {}
    for name, n in (("ssb_customer", n_c), ("ssb_supplier", n_s), ("ssb_part", n_p), ("ssb_dates", T.SSB_DAYS)):
        dims[name], k = T.device_chunk(name, 0, n, P.SSB_DIM_COLS[name], dev)
        keep.append(k)
    lo, k = T.device_chunk("lineorder", 0, n_lo, T.LINEORDER_Q41, dev, **kw)
    keep.append(k)
    torch.cuda.synchronize()

    def step(times=None):
        t0 = time.perf_counter()
        q = P.SsbQ41(ctx).build(dims)
        if times is not None:
            torch.cuda.synchronize()
            t1 = time.perf_counter()
        stage_rows = []
        agg = q.probe(lo, stage_rows=stage_rows)
        out = agg.pull_all()
        if times is not None:
            torch.cuda.synchronize()
            times.append({"filter + build 4 dimensions": (t1 - t0) * 1e3, "4 probes + group by": (time.perf_counter() - t1) * 1e3})
        launches = sum(o.counters()["kernel_launches"] for o in q.joins + q.keep + [agg])
        agg.close()
        q.close()
        return out, stage_rows, launches

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out, stage_rows, launches = step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.steps
    stage = []
    step(stage)
    alg_bytes = n_lo * 24 + n_c * 12 + n_s * 8 + n_p * 8 + T.SSB_DAYS * 8
    peak = 6450.3
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    print(json.dumps({
        "metric": "SSB Q4.1 operator pipeline rows/s (lineorder rows)", "value": n_lo / dt, "unit": "rows/s", "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "dtype": "int64", "data": "synthetic",
        "config": {"workload": "SSB Q4.1: 3 dimension filters, 4 hash joins (broadcast build sides), GROUP BY d_year, c_nation", "lineorder_rows": n_lo,
                   "customer_rows": n_c, "supplier_rows": n_s, "part_rows": n_p, "dates_rows": T.SSB_DAYS},
        "rows_after_probe": dict(zip(["supplier", "customer", "part", "dates"], stage_rows)), "groups": out.n_rows, "gpu_launches": launches,
        "stages_ms": stage[0],
        "roofline": {"bound": "hbm", "achieved": alg_bytes / dt / 1e9, "peak": peak, "unit": "GB/s", "frac": alg_bytes / dt / 1e9 / peak,
                     "note": "whole pipeline: algorithmic input bytes (24 B per lineorder row + the dimensions) over the step's wall time"},
    }))
    ctx.close()


if __name__ == "__main__":
    main()
