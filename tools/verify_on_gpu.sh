#!/usr/bin/env bash
# One gpurun call that re-checks the round's claims on a B200:
#   /usr/local/graft/bin/gpurun --timeout 900 -- 'bash tools/verify_on_gpu.sh'
# Everything lands in gpurun_out/v_*; nothing here is a bench value taken under a profiler.
set -u
mkdir -p gpurun_out
# 1. the whole GPU suite (about 5.5 minutes, most of it the high-cardinality cases) and the smoke entry
timeout 500 python -m pytest tests -m gpu -q > gpurun_out/v_gpu_suite.log 2>&1
tail -3 gpurun_out/v_gpu_suite.log
python __graft_entry__.py smoke 2>&1 | tail -2
# 2. the default bench line (Q1 headline + queries.*, e2e from encoded pages, parity) and the reference arm
timeout 300 python bench.py --steps 10 --warmup 3 2> gpurun_out/v_bench.err | tail -1 > gpurun_out/v_bench.json
timeout 200 python bench.py --impl reference --steps 10 --warmup 3 2>/dev/null | tail -1 > gpurun_out/v_bench_reference.json
python - <<'EOF'
import json
d = json.loads(open("gpurun_out/v_bench.json").read()); r = json.loads(open("gpurun_out/v_bench_reference.json").read())
print("Q1", round(d["ms_per_step"], 2), "ms/step, roofline", round(d["roofline"]["frac"], 3), "e2e", round(d["e2e"]["value"] / 1e9, 2), "G rows/s =",
      round(d["e2e"]["value"] / r["value"], 1), "x the reference arm; parity_all", d.get("parity_all"))
for k, x in d["queries"].items():
    if k == "q1_interpreter_variants":
        for kk, v in x.items():
            if isinstance(v, dict) and "kernel_ms" in v:
                print("  ", kk, round(v["kernel_ms"], 2), "ms, roofline", round(v["roofline"]["frac"], 2), "interpreter", round(v["interpreter_kernel_ms"], 1), "ms, same rows", v["same_rows_as_interpreter"])
    else:
        print(k, round(x["ms_per_step"], 2) if x.get("ms_per_step") else x.get("error"), "parity", x.get("parity"))
EOF
# 3. the compiled kernels on their own (134 M rows per launch), and the built-in Q1 shape against its run-time compiled twin
export PYTHONPATH=.
for v in binary nullable minmax extra; do SRB_JIT=sync python tools/generic_probe.py 134217728 $v | tail -1; done
SRB_JIT=sync SRB_NO_BUILTIN_SHAPES=1 python tools/generic_probe.py 134217728 q1 | tail -1
python tools/generic_probe.py 134217728 q1 | tail -1
# 4. launch lists of the two join pipelines (library kernels only)
for q in q3 ssb; do
    timeout 90 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv --log-file gpurun_out/v_${q}_launches.csv \
        python bench.py --query $q --steps 1 --warmup 1 --no-cpu --no-e2e --no-parity > gpurun_out/v_ncu_$q.log 2>&1
done
