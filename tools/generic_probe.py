"""TPC-H Q1 through the general kernel path (nullable measures: no static shape applies) on one GPU; prints kernel times.
    PYTHONPATH=. python tools/generic_probe.py [rows] [variant: nullable | minmax | binary]
Used under `ncu --set full -k regex:k_scan_agg -c 1` to read the interpreter's bound."""
import sys
import torch
from starrocks_b200 import capi as K, ops, queries as Q, tpch_synth as T

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 64 << 20
variant = sys.argv[2] if len(sys.argv) > 2 else "nullable"
ctx = ops.Context(device=0)
dev = torch.device("cuda", 0)
chunk, keep = T.device_chunk("lineitem", 0, rows, T.LINEITEM_Q1, dev)
plan_fn = Q.q1_plan
if variant == "nullable":
    for slot in (Q.L_QUANTITY, Q.L_DISCOUNT):
        nm = (torch.rand(rows, device=dev) < 0.01).to(torch.uint8)
        keep[f"null{slot}"] = nm
        c = chunk.columns[slot]
        chunk.columns[slot] = K.Column(c.type, c.data, nm.data_ptr(), None, c.precision, c.scale, K.MEM_DEVICE)
elif variant == "extra":  # two more aggregates than the built-in Q1 shape
    from tests.test_jit import lineitem_plan
    plan_fn = lambda: lineitem_plan(charge=True)
elif variant == "minmax":
    plan_fn = lambda: Q.q1_plan(with_min_max=True)
elif variant == "binary":
    for slot, code_name, names in ((Q.L_RETURNFLAG, "l_returnflag_code", T.RETURNFLAGS), (Q.L_LINESTATUS, "l_linestatus_code", T.LINESTATUS)):
        lut = torch.tensor(list(b"".join(names)), dtype=torch.uint8, device=dev)
        data = torch.cat([lut[keep[code_name][:rows].long()], torch.zeros(16, dtype=torch.uint8, device=dev)])
        offs = torch.arange(rows + 1, dtype=torch.int32, device=dev)
        keep[f"bin{slot}"] = (data, offs)
        chunk.columns[slot] = K.Column(K.TYPE_VARCHAR, data.data_ptr(), None, offs.data_ptr(), 0, 0, K.MEM_DEVICE, rows)
    for s_ in (Q.L_RETURNFLAG_CODE, Q.L_LINESTATUS_CODE):
        chunk.columns.pop(s_, None)
    plan_fn = lambda: Q.q1_plan(dict_keys=False)
torch.cuda.synchronize()
for it in range(3):
    op = ops.ScanAggOperator(ctx, plan_fn())
    op.push_chunk(chunk)
    op.set_finishing()
    out = op.pull_all()
    p = op.profile()
    print(f"{variant}: {rows} rows, kernel {p['kernel_ms']:.3f} ms, {rows / p['kernel_ms'] / 1e6:.2f} G rows/s, shape {p['shape']}, static launches {p['static_launches']}, result rows {out.n_rows}")
    op.close()
ctx.close()
