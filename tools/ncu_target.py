"""ncu target: warm both render-kernel variants once, then launch each once more (profile with -s 2 -c 2).
Usage: python tools/ncu_target.py [config] [waves] [volume_scale]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from volume_path_tracer_b200 import api, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "dense512"
waves = int(sys.argv[2]) if len(sys.argv) > 2 else 4
factor = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
variants = [int(v) for v in os.environ.get("VARIANTS", "0,1").split(",")]
scene, spec = scenes.load_config(name)
if factor != 1.0:
    spec = spec.scaled(factor)
g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
r = api.Renderer(0)
r.set_volume(g)
d = torch.zeros((scene.height, scene.width, 4), dtype=torch.float32, device="cuda:0")
for phase in range(2):
    for v in variants:
        r.set_kernel_variant(v)
        r.run_device(scene, d.data_ptr(), 1 + phase * waves, waves, 1)
        print(f"{name} variant {v} phase {phase}: {r.last_render_ms():.2f} ms", flush=True)
