#!/bin/bash
# Job-pool kernel, threshold policy: parity, then a parameter sweep against variant 2 on the same box.
mkdir -p gpurun_out
OUT=gpurun_out/pool_ab2.log; : > $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "variants_agree" > gpurun_out/pool_parity.log 2>&1; echo "parity rc=$?" | tee -a $OUT; tail -2 gpurun_out/pool_parity.log | tee -a $OUT
export WAVES=${WAVES:-64} CASES=${CASES:-dense512,wdas_cloud}
VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
VARIANTS=3 POOL="${POOL:-2,8,28,28;3,8,28,28;4,8,28,28;3,8,24,24;3,8,32,32;3,16,28,28;3,8,32,24;2,8,32,32;3,4,30,30}" timeout 1200 python tools/perf_probe.py 2>&1 | tee -a $OUT
VPT_KERNEL_VARIANT=3 timeout 600 python tools/sched_stats.py 2>&1 | grep -v "^  per" | tee -a $OUT
