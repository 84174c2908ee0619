"""Development aid: time the render kernel variants on the BASELINE configs (device time, film resident)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from volume_path_tracer_b200 import api, scenes

variants = [int(v) for v in os.environ.get("VARIANTS", "1,2").split(",")]
threshes = os.environ.get("THRESH", "").split(";") if os.environ.get("THRESH") else [None]  # "12,8,8,12;16,8,8,16"
cases = [("fire", 1.0, 16), ("fire_lowscattering", 1.0, 16), ("wdas_cloud", 0.25, 8), ("dense512", 1.0, 8)]
if os.environ.get("FULL_CLOUD"):  # the full-resolution WDAS-shaped cloud (3.6 GB grid) as an extra case
    cases.append(("wdas_cloud", 1.0, 8))
if os.environ.get("WAVES"):  # override the waves per launch of every case
    cases = [(n, f, int(os.environ["WAVES"])) for n, f, _ in cases]
if os.environ.get("CASES"):
    cases = [c for c in cases if c[0] in os.environ["CASES"].split(",")]
if os.environ.get("ONLY_FULL_CLOUD"):
    cases = [("wdas_cloud", 1.0, int(os.environ.get("WAVES", 32)))]
pools = os.environ.get("POOL", "").split(";") if os.environ.get("POOL") else [None]  # variant 3: "3,20;2,16" = HOT passes, min HOT lanes
if pools != [None]:
    threshes = pools
for name, factor, waves in cases:
    scene, spec = scenes.load_config(name)
    if factor != 1.0:
        spec = spec.scaled(factor)
    if os.environ.get("TILE"):  # e.g. TILE=1,1: per-pixel jobs (a configuration the reference accepts: tile_size [1,1])
        scene.tile_size[0], scene.tile_size[1] = [int(v) for v in os.environ["TILE"].split(",")]
    g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
    d = torch.zeros((scene.height, scene.width, 4), dtype=torch.float32, device="cuda:0")
    ps = scene.width * scene.height * waves
    for th in threshes:
        if th:
            os.environ["VPT_POOL_PARAMS" if pools != [None] else "VPT_SCHED_THRESH"] = th
        r = api.Renderer(0)
        r.set_volume(g)
        for v in variants:
            r.set_kernel_variant(v)
            r.run_device(scene, d.data_ptr(), 1, 2, 1)
            ms = []
            for i in range(3):
                r.run_device(scene, d.data_ptr(), 1 + i * waves, waves, 1); ms.append(r.last_render_ms())
            print(f"{name:20s} dims {spec.dims} variant {v} thresh {th}: {min(ms):9.2f} ms  {ps / min(ms) / 1e3:9.2f} Mpx-samples/s", flush=True)
        r.close()
