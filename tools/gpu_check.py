"""Quick GPU-vs-oracle sanity script (development aid; the real parity tests live in tests/)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from volume_path_tracer_b200 import api, scenes
from oracle import oracle_lib as ol

orc = ol.Oracle("own")
r = api.Renderer(0)
for name, factor, size in [("fire", 0.25, (70, 100)), ("wdas_cloud", 0.05, (96, 54)), ("dense512", 0.125, (96, 54))]:
    scene, spec = scenes.small_variant(name, size, 2, factor)
    g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
    vo = orc.volume(g)
    r.set_volume(g)
    # sampler parity
    rs = np.random.default_rng(0)
    lo = np.array([-8, -8, -8], np.float32); hi = np.array(spec.dims, np.float32) + 8
    # index bbox is not at origin necessarily: sample around segments instead
    pts = (rs.random((20000, 3)).astype(np.float32) * 2 - 1) * np.array(spec.dims, np.float32)
    v0, d0, m0 = vo.sample_grid(0, pts); v1, d1, m1 = r.sample_grid(0, pts)
    print(name, "sample eq", np.array_equal(v0, v1), np.array_equal(d0, d1), np.array_equal(m0, m1), "nonzero", (v0 != 0).mean())
    jids = np.arange(0, scene.ntiles * 2, 1, dtype=np.uint64)
    t = time.time(); x0, n0, c0 = vo.trace_jobs(scene, jids); t0 = time.time() - t
    t = time.time(); x1, n1 = r.trace_jobs(scene, jids); t1 = time.time() - t
    agree = (n0 == n1).all(axis=1)
    rel = np.abs(x0 - x1) / np.maximum(np.abs(x0), 1e-20)
    print(name, "jobs", len(jids), "draw-count agreement (jobs)", agree.mean(), "max rel err on agreeing", rel[agree].max() if agree.any() else None,
          "bit-equal frac", (x0[agree] == x1[agree]).mean(), "cpu s", t0, "gpu s", t1)
    fo = np.zeros((scene.height, scene.width, 4), np.float32); fg = fo.copy()
    vo.render(scene, fo); r.enable_counters(True); r.run(scene, fg); c = r.counters(); r.enable_counters(False)
    print("  film close", np.abs(fo - fg).max(), "counters", c.as_dict(), "ms", r.last_render_ms())

# throughput on bigger configs
for name, factor, size, waves in [("fire", 1.0, (700, 1000), 4), ("wdas_cloud", 0.25, (1920, 1080), 2), ("dense512", 1.0, (1920, 1080), 2)]:
    scene, spec = scenes.small_variant(name, size, waves, factor)
    t = time.time(); g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor); ts = time.time() - t
    t = time.time(); r.set_volume(g); tb = time.time() - t
    film = r.new_film(scene)
    r.run(scene, film)
    r.enable_counters(True); film[:] = 0; r.run(scene, film); c = r.counters(); msc = r.last_render_ms(); r.enable_counters(False)
    film[:] = 0; r.run(scene, film); ms = r.last_render_ms()
    ps = scene.width * scene.height * waves
    print(name, spec.dims, "synth s", ts, "bake+upload s", tb, "grid MB", g.density.nbytes / 1e6, "ms", ms, "ms(count)", msc, "Mpx-samples/s", ps / ms / 1e3,
          "alg GB/s", c.algorithmic_bytes() / ms / 1e6, c.as_dict(), "mean", film[..., :3].mean())
