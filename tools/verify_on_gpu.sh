#!/usr/bin/env bash
# One gpurun call that settles what round 1 left unverified on hardware:
#   /usr/local/graft/bin/gpurun --timeout 600 -- 'bash tools/verify_on_gpu.sh'
# Everything lands in gpurun_out/v_*; nothing here is a bench value taken under a profiler.
set -u
mkdir -p gpurun_out
# 1. the tests marked not-yet-on-hardware, run as ordinary tests (a failure here is a real failure)
timeout 180 python -m pytest tests/test_gpu_topn.py tests/test_gpu_unverified.py -m gpu -q --runxfail > gpurun_out/v_unverified_tests.log 2>&1
tail -15 gpurun_out/v_unverified_tests.log
# 2. the whole GPU suite (about 5 minutes, most of it the high-cardinality cases)
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/v_gpu_suite.log 2>&1
tail -3 gpurun_out/v_gpu_suite.log
# 3. bench lines: default (Q1), then the join pipelines and the high-cardinality GROUP BY
for q in q1 q6 q3 ssb highcard; do
    timeout 120 python bench.py --query $q --steps 5 --warmup 3 --cpu-seconds 8 2> gpurun_out/v_bench_$q.err | tail -1 > gpurun_out/v_bench_$q.json
    cut -c1-220 gpurun_out/v_bench_$q.json
done
# 3b. A/B of the fused runtime-filter kernel on SSB
SRB_RF_FUSED=1 timeout 60 python bench.py --query ssb --steps 5 --warmup 3 --no-cpu --no-e2e 2>/dev/null | tail -1 > gpurun_out/v_bench_ssb_fused.json
cut -c1-220 gpurun_out/v_bench_ssb_fused.json
# 4. launch lists of the two join pipelines (library kernels only)
for q in q3 ssb; do
    timeout 90 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 400 --csv --log-file gpurun_out/v_${q}_launches.csv \
        python bench.py --query $q --steps 1 --warmup 1 --no-cpu --no-e2e > gpurun_out/v_ncu_$q.log 2>&1
done
