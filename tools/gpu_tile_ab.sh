#!/bin/bash
# Launch-size sensitivity against the tile size (job = tile x wave; pixels of a job are serial): 8x8 (shipped scenes), 4x4, 2x2, 1x1.
mkdir -p gpurun_out
OUT=gpurun_out/tile_ab.log; : > $OUT
for T in 8,8 4,4 2,2 1,1; do
  for W in ${WAVES_LIST:-8 32 128}; do
    echo "== tile $T waves $W" | tee -a $OUT
    TILE=$T WAVES=$W CASES=${CASES:-dense512,wdas_cloud} VARIANTS=2 timeout 900 python tools/perf_probe.py 2>&1 | tee -a $OUT
  done
done
