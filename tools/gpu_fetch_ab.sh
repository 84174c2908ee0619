#!/bin/bash
# Fetch-layout A/B on one box (profiles/r02_fetch_layout.txt): (i) default 9^3 bricks in leaf-array order, (ii) the same bricks in Morton
# order of the leaf origins (VPT_BAKE_BRICK_ORDER=morton), (iii) z-pair bricks, 4 x LDG.64 per request (build/libs/libvptcuda_pairs.so),
# (iv) one 32-byte stencil per voxel, 2 x LDG.128 per request (build/libs/libvptcuda_stencil.so).
# For each: bit-exactness subset, device-time throughput, one ncu --set full capture of the default kernel (dense512, 16 waves).
mkdir -p gpurun_out
OUT=gpurun_out/fetch_ab.log; : > $OUT
for V in ${LAYOUTS:-default morton pairs}; do
  unset VPT_LIBVPTCUDA VPT_BAKE_BRICK_ORDER
  [ $V = pairs ] && export VPT_LIBVPTCUDA=build/libs/libvptcuda_pairs.so
  [ $V = stencil ] && export VPT_LIBVPTCUDA=build/libs/libvptcuda_stencil.so   # one 32-byte 2x2x2 stencil per voxel (-DVPT_BRICK_PAIRS=2)
  [ $V = morton ] && export VPT_BAKE_BRICK_ORDER=morton
  echo "== $V" | tee -a $OUT
  timeout 600 python -m pytest tests/test_gpu_parity.py -q -k "per_job_radiance or edge_cases or film_and_counters or sample_grid" > gpurun_out/fetch_parity_$V.log 2>&1
  echo "parity($V) rc=$? $(tail -1 gpurun_out/fetch_parity_$V.log)" | tee -a $OUT
  WAVES=64 CASES=dense512,wdas_cloud VARIANTS=2 timeout 600 python tools/perf_probe.py 2>&1 | tee -a $OUT
  echo "host MemAvailable: $(awk '/MemAvailable/ {print int($2/1048576)}' /proc/meminfo) GiB" | tee -a $OUT
  [ -n "$FULL" ] && [ $(awk '/MemAvailable/ {print int($2/1048576)}' /proc/meminfo) -gt 120 ] && ONLY_FULL_CLOUD=1 WAVES=32 VARIANTS=2 timeout 900 python tools/perf_probe.py 2>&1 | tee -a $OUT
  VARIANTS=2 timeout 900 bash tools/gpu_ncu.sh dense512 16 1.0 r02_fetch_$V 2>&1 | tee -a $OUT
done
