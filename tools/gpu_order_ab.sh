#!/bin/bash
# Tile order (longest tiles first) A/B: parity, then device-time throughput at 64 and 16 waves per call, ordered / order disabled / kernel built without it.
mkdir -p gpurun_out; OUT=gpurun_out/order_ab.log; : > $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py -q -x > gpurun_out/order_parity.log 2>&1; echo "parity rc=$? $(tail -1 gpurun_out/order_parity.log)" | tee -a $OUT
for MODE in on off nolib on off; do
  unset VPT_LIBVPTCUDA VPT_TILE_ORDER
  [ $MODE = off ] && export VPT_TILE_ORDER=0
  [ $MODE = nolib ] && export VPT_LIBVPTCUDA=build/libs/libvptcuda_noorder.so
  for W in 64 16; do
    echo "== order $MODE, $W waves per call" | tee -a $OUT
    WAVES=$W CASES=dense512,wdas_cloud VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
  done
  [ $MODE != nolib ] && { echo "== order $MODE, full cloud, 16 waves per call" | tee -a $OUT; ONLY_FULL_CLOUD=1 WAVES=16 VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT; }
done
