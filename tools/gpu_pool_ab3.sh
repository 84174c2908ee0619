#!/bin/bash
# Is the job-pool kernel latency bound? Same kernel at 10 and 20 warps per SM, and smaller pools at 28 / 32 warps per SM.
mkdir -p gpurun_out
OUT=gpurun_out/pool_ab3.log; : > $OUT
export WAVES=${WAVES:-64} CASES=${CASES:-dense512}
echo "== P=64, 20 warps/SM" | tee -a $OUT
VARIANTS=3 POOL="3,8,24,24" timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
echo "== P=64, 10 warps/SM" | tee -a $OUT
VPT_MAX_BLOCKS_PER_SM=1 VARIANTS=3 POOL="3,8,24,24" timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
echo "== variant 2 at 7 / 5 / 3 blocks (28 / 20 / 12 warps) per SM" | tee -a $OUT
for b in 7 5 3; do VPT_MAX_BLOCKS_PER_SM=$b VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT; done
echo "== P=48, 28 warps/SM" | tee -a $OUT
VPT_LIBVPTCUDA=build/libs/libvptcuda_p48w14.so VARIANTS=3 POOL="3,8,20,20;3,8,16,16;3,8,24,24;2,8,20,20" timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
echo "== P=40, 32 warps/SM" | tee -a $OUT
VPT_LIBVPTCUDA=build/libs/libvptcuda_p40w16.so VARIANTS=3 POOL="3,8,16,16;3,8,12,12;3,8,20,20;2,8,16,16" timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
