#!/bin/bash
# Device brick stage of the baker vs the host baker: parity test, then set_volume wall time for dense512 and the full cloud (both paths).
mkdir -p gpurun_out; OUT=gpurun_out/bake_ab.log; : > $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "device_bake or per_job_radiance or sample_grid or segments" > gpurun_out/bake_parity.log 2>&1; echo "parity rc=$? $(tail -1 gpurun_out/bake_parity.log)" | tee -a $OUT
python - <<'PY' 2>&1 | tee -a $OUT
import os, time
from volume_path_tracer_b200 import api, scenes
for cfg in ("dense512", "wdas_cloud"):
    scene, spec = scenes.load_config(cfg)
    g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
    for mode in ("device", "host", "device", "host"):
        os.environ.pop("VPT_HOST_BAKE", None)
        if mode == "host": os.environ["VPT_HOST_BAKE"] = "1"
        r = api.Renderer(0)
        t = time.time(); r.set_volume(g); dt = time.time() - t
        print(f"{cfg:12s} {g.density.nbytes / 1e6:8.0f} MB grid, brick stage on the {mode:6s}: vptc_set_volume {dt * 1e3:8.1f} ms, {r.device_volume_bytes() / 1e6:.0f} MB baked", flush=True)
        r.close()
PY
