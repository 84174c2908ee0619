"""Launch drain A/B (development aid + profiles/r02_launch_drain.txt): the same 128 waves rendered as one call, and as many small calls —
blocking (every call ends with a drain one job long) and asynchronous (device-side range ring: the running kernel carries on into the
next call). Device film and host film. Usage: python tools/drain_ab.py [config] [total_waves] [volume_scale]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from volume_path_tracer_b200 import api, scenes

name = sys.argv[1] if len(sys.argv) > 1 else "dense512"
total = int(sys.argv[2]) if len(sys.argv) > 2 else 128
factor = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
scene, spec = scenes.load_config(name)
if factor != 1.0:
    spec = spec.scaled(factor)
g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
r = api.Renderer(0)
r.set_volume(g)
d = torch.zeros((scene.height, scene.width, 4), dtype=torch.float32, device="cuda:0")
h = np.zeros((scene.height, scene.width, 4), np.float32)
ps = scene.width * scene.height * total
r.run_device(scene, d.data_ptr(), 1, 8, 1)  # warm-up
torch.cuda.synchronize()


def timed(fn):
    torch.cuda.synchronize(); t = time.perf_counter(); fn(); torch.cuda.synchronize(); return time.perf_counter() - t


print(f"{name} {spec.dims} {scene.width}x{scene.height}, {total} waves, tile {scene.tile_size[0]}x{scene.tile_size[1]}; M pixel-samples/s (wall clock incl. all copies)")
for per in (total, 32, 16, 8, 4):
    calls = [(1 + i, min(per, total - i)) for i in range(0, total, per)]
    def dev_block():
        for f, n in calls: r.run_device(scene, d.data_ptr(), f, n, 1)
    def dev_async():
        for f, n in calls: r.run_device_range(scene, d.data_ptr(), f, n, 1, wait=False)
        r.wait()
    def host_block():
        for f, n in calls: r.run(scene, h, first_wave=f, n_waves=n)
    def host_async():
        for f, n in calls: r.run_async(scene, h, f, n)
        r.wait()
    row = []
    for fn in (dev_block, dev_async, host_block, host_async):
        d.zero_(); h[:] = 0
        s = timed(fn)
        row.append(ps / s / 1e6)
    ok = bool((h[..., 3] == total).all())
    print(f"  {per:4d} waves/call x {len(calls):3d}: device film blocking {row[0]:7.1f}  async {row[1]:7.1f} | host film blocking {row[2]:7.1f}  async {row[3]:7.1f}   counts ok {ok}", flush=True)
