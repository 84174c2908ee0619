import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["VPT_SCHED_STATS"] = "1"
import torch
from volume_path_tracer_b200 import api, scenes
for name, factor, waves in [("dense512", 1.0, int(os.environ.get("SWAVES", 16))), ("wdas_cloud", 0.25, int(os.environ.get("SWAVES", 16))), ("fire", 1.0, 32)]:
    scene, spec = scenes.load_config(name)
    if factor != 1.0: spec = spec.scaled(factor)
    g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
    for th in os.environ.get("THRESH", "12,8,8,12").split(";"):
        os.environ["VPT_SCHED_THRESH"] = th
        r = api.Renderer(0); r.set_volume(g); r.enable_counters(True)
        d = torch.zeros((scene.height, scene.width, 4), dtype=torch.float32, device="cuda:0")
        print(name, "thresh", th, flush=True)
        r.run_device(scene, d.data_ptr(), 1, waves, 1)
        c = r.counters().as_dict(); print("  per sample:", {k: round(v / c["pixel_samples"], 2) for k, v in c.items()}, "ms", r.last_render_ms(), flush=True)
        r.close()
