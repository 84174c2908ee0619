#!/bin/bash
# Variant 4 (job pool, cold state in global memory, 30 warps/SM): parity of all variants, then same-box timing against variants 2 and 3.
mkdir -p gpurun_out
OUT=gpurun_out/pool4.log; : > $OUT
timeout 300 python -m pytest tests/test_gpu_parity.py -q -x -k "variants_agree or tile_sizes" > gpurun_out/pool4_parity.log 2>&1; echo "parity rc=$?" | tee -a $OUT; tail -3 gpurun_out/pool4_parity.log | tee -a $OUT
export WAVES=${WAVES:-64} CASES=${CASES:-dense512,wdas_cloud}
VARIANTS=2 timeout 300 python tools/perf_probe.py 2>&1 | tee -a $OUT
VARIANTS=3 POOL="3,8,24,24" timeout 300 python tools/perf_probe.py 2>&1 | tee -a $OUT
VARIANTS=4 POOL="${POOL:-3,8,24,24;3,8,28,28;2,8,24,24;4,8,24,24}" timeout 300 python tools/perf_probe.py 2>&1 | tee -a $OUT
