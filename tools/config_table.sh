#!/bin/bash
# One bench line per BASELINE.json config at full size on one B200 (CPU-port leg included) -> gpurun_out/r02_configs.jsonl
mkdir -p gpurun_out; : > gpurun_out/r02_configs.jsonl
for c in fire fire_lowscattering wdas_cloud dense512 wdas_cloud_4k; do
  echo "== $c" >&2
  timeout 900 python bench.py --config $c --steps 3 --warmup 3 --no-strong >> gpurun_out/r02_configs.jsonl 2>> gpurun_out/r02_configs.err || echo "{\"config\": \"$c\", \"failed\": $?}" >> gpurun_out/r02_configs.jsonl
done
