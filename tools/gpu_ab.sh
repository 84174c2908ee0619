#!/bin/bash
# Same-box A/B of side-by-side builds of the default kernel (build/libs/libvptcuda_<name>.so, chosen with VPT_LIBVPTCUDA): bit-exactness first,
# then device-time throughput. LIBS="default e1 e2" bash tools/gpu_ab.sh  -> gpurun_out/ab.log
mkdir -p gpurun_out
OUT=gpurun_out/${ABLOG:-ab.log}; : > $OUT
for L in ${LIBS:-default}; do
  if [ $L = default ]; then unset VPT_LIBVPTCUDA; else export VPT_LIBVPTCUDA=build/libs/libvptcuda_$L.so; fi
  echo "== $L" | tee -a $OUT
  timeout 600 python -m pytest tests/test_gpu_parity.py -q -k "per_job_radiance or edge_cases or film_and_counters or event_log or async or stripes" > gpurun_out/ab_parity_$L.log 2>&1; echo "parity($L) rc=$? $(tail -1 gpurun_out/ab_parity_$L.log)" | tee -a $OUT
  WAVES=${WAVES:-64} CASES=${CASES:-dense512,wdas_cloud} VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
  [ -n "$FIRE" ] && WAVES=32 CASES=fire VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
done
