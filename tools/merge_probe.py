"""phase-2 merge of the high-cardinality GROUP BY on one GPU: phase-1 (serialize) output of 600 M rows fed to the merge operator; prints times"""
import sys, time
import torch
from starrocks_b200 import capi as K, ops, queries as Q, tpch_synth as T

rows, keys = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000_000, int(sys.argv[2]) if len(sys.argv) > 2 else 50_000_000
ctx = ops.Context(device=0)
dev = torch.device("cuda", 0)
chunk, keep = T.device_chunk("highcard", 0, rows, ["k", "v"], dev, n_keys=keys)
torch.cuda.synchronize()
op = ops.ScanAggOperator(ctx, Q.highcard_plan(phase=K.AGG_UPDATE_SERIALIZE, group_hint=keys))
op.push_chunk(chunk)
op.set_finishing()
inter = op.pull_device()
print("phase 1:", op.profile(), "tuples", inter.n_rows, {s: (c.type, bool(c.nulls)) for s, c in inter.columns.items()})
for it in range(3):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    op2 = ops.ScanAggOperator(ctx, Q.highcard_merge_plan(group_hint=keys))
    op2.push_chunk(inter)
    t1 = time.perf_counter()
    op2.set_finishing()
    t2 = time.perf_counter()
    out = op2.pull_device()
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    print(f"merge: push {1e3*(t1-t0):.2f} ms, finish {1e3*(t2-t1):.2f} ms, pull {1e3*(t3-t2):.2f} ms", op2.profile(), op2.counters(), out.n_rows)
    op2.close()
op.close()
ctx.close()
