#!/bin/bash
# Same-box A/B of micro-variants of the default (variant 2) kernel: side-by-side libraries in build/libs, parity first.
mkdir -p gpurun_out
OUT=gpurun_out/micro_ab.log; : > $OUT
for L in ${LIBS:-default nosplit pmul}; do
  if [ $L = default ]; then unset VPT_LIBVPTCUDA; else export VPT_LIBVPTCUDA=build/libs/libvptcuda_$L.so; fi
  echo "== $L" | tee -a $OUT
  timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "per_job_radiance or variants_agree or edge_cases" > gpurun_out/micro_parity_$L.log 2>&1; echo "parity($L) rc=$?" | tee -a $OUT
  WAVES=64 CASES=dense512,wdas_cloud VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
  WAVES=16 CASES=fire VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
  [ -n "$FULL" ] && ONLY_FULL_CLOUD=1 WAVES=32 VARIANTS=2 timeout 900 python tools/perf_probe.py 2>&1 | tee -a $OUT
done
