#!/bin/bash
# One GPU-box visit for the job-pool kernel (variant 3): parity of all variants, then same-box A/B timing against
# variant 2 for the side-by-side builds in build/libs (pool slots per warp x warps per block) and a parameter sweep.
mkdir -p gpurun_out
OUT=gpurun_out/pool_ab.log; : > $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "variants_agree or per_job_radiance or film_and_counters or edge_cases" > gpurun_out/pool_parity.log 2>&1; echo "parity(default lib) rc=$?" | tee -a $OUT; tail -3 gpurun_out/pool_parity.log | tee -a $OUT
for L in ${LIBS:-p48 p56}; do
  VPT_LIBVPTCUDA=build/libs/libvptcuda_$L.so timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "variants_agree" > gpurun_out/pool_parity_$L.log 2>&1; echo "parity($L) rc=$?" | tee -a $OUT
done
export WAVES=${WAVES:-32} CASES=${CASES:-dense512,wdas_cloud,fire}
echo "== variant 2 (default lib)" | tee -a $OUT
VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
echo "== variant 3, default lib (P=64 x 10 warps)" | tee -a $OUT
VARIANTS=3 POOL="${POOL:-1,1;2,1;3,1;2,16;3,20;4,24;6,24}" timeout 900 python tools/perf_probe.py 2>&1 | tee -a $OUT
for L in ${LIBS:-p48 p56}; do
  echo "== variant 3, $L" | tee -a $OUT
  VPT_LIBVPTCUDA=build/libs/libvptcuda_$L.so VARIANTS=3 POOL="${POOL2:-2,1;3,20;4,24}" timeout 900 python tools/perf_probe.py 2>&1 | tee -a $OUT
done
echo "== stage statistics, variant 3 default lib" | tee -a $OUT
VPT_KERNEL_VARIANT=3 timeout 600 python tools/sched_stats.py 2>&1 | tee -a $OUT
