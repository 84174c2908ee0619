import os, sys, time
sys.path.insert(0, '.')
import torch
from volume_path_tracer_b200 import api, scenes
scene, spec = scenes.load_config("wdas_cloud"); spec = spec.scaled(0.25)
g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
ntx = -(-scene.width // int(scene.tile_size[0])); nty = -(-scene.height // int(scene.tile_size[1]))
r = api.Renderer(0); r.set_volume(g)
d = torch.zeros((scene.height, scene.width, 4), dtype=torch.float32, device="cuda:0")
r.run_device(scene, d.data_ptr(), 1, 2, 1)
for i in range(3):
    r.run_device_range(scene, d.data_ptr(), 1 + 64 * i, 64, tile_begin=0, tile_count=ntx * len(range(0, nty, 2)), tile_row_stride=2)
    print("order", os.environ.get("VPT_TILE_ORDER", "1"), "interleaved half frame, 64 waves:", round(r.last_render_ms(), 2), "ms", flush=True)
