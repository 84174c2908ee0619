#!/bin/bash
# r02 kernel A/B, second series (profiles/r02_kernel_ab.txt): HDDA look-ahead, split / blocked majorant table, scheduler thresholds.
# Each build: bit-exactness subset, device-time throughput; then one ncu --set full capture per build named in CAPTURE.
mkdir -p gpurun_out
LIBS="${LIBS:-default la maj1 maj2 lamaj2}" ABLOG=kernel_ab2.log bash tools/gpu_ab.sh
OUT=gpurun_out/kernel_ab2.log
if [ -n "$THRESH" ]; then
  unset VPT_LIBVPTCUDA
  echo "== default, scheduler thresholds (setup,hot,-,eval)" | tee -a $OUT
  WAVES=64 CASES=dense512,wdas_cloud VARIANTS=2 THRESH="$THRESH" timeout 900 python tools/perf_probe.py 2>&1 | tee -a $OUT
fi
for L in $CAPTURE; do
  if [ $L = default ]; then unset VPT_LIBVPTCUDA; else export VPT_LIBVPTCUDA=build/libs/libvptcuda_$L.so; fi
  VARIANTS=2 timeout 900 bash tools/gpu_ncu.sh dense512 16 1.0 r02_kernel_$L 2>&1 | tee -a $OUT
done
