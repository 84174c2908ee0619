"""Mint the golden vectors under tests/golden/ from oracle/_ref — the oracle build that runs the
REFERENCE'S OWN vendored NanoVDB (HDDA, Ray, ReadAccessor, SampleFromVoxels, Map) and pcg-cpp headers from
/root/reference/external, included where they lie. The reference has no tests or stored outputs
(SURVEY.md §4), so these are produced by executing its arithmetic here; the fixtures travel to the GPU box,
the reference tree does not.

    python tests/golden/make_golden.py        # needs /root/reference (oracle/_ref is built on demand)

Inputs are the seeded procedural volumes of volume_path_tracer_b200 (bit-reproducible: integer hashes and
+ - * / sqrt only); their SHA-256 is stored so generator drift is noticed.
"""
import hashlib
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))
sys.path.insert(0, str(HERE.parent))

from oracle import oracle_lib  # noqa: E402
from volume_path_tracer_b200 import api, scenes  # noqa: E402
from conftest import SMALL  # noqa: E402

RNG_CASES = [(500, 0), (500, 1), (500, 10999), (500, 11000), (10, 0), (10, 1), (10, 10999), (10, 11000), (0, 0), (4294967295, 2**40 + 7)]


def ray_set(scene, orc, n=64, seed=3):
    rs = np.random.default_rng(seed)
    px = rs.integers(0, scene.width, n); py = rs.integers(0, scene.height, n)
    jx = (rs.random(n) * 0.5).astype(np.float32); jy = (rs.random(n) * 0.5).astype(np.float32)
    return orc.camera_rays(scene, px, py, jx, jy)


def main():
    oracle_lib.build(ref=True)
    ref = oracle_lib.Oracle("ref")
    assert ref.substrate() == "nanovdb"
    out = {}
    # ---- RNG KATs (SURVEY.md Appendix C + extremes) ----
    out["rng_cases"] = np.array(RNG_CASES, dtype=np.uint64)
    st, u32, f32 = [], [], []
    for seed, jid in RNG_CASES:
        s0, u, f = ref.rng(seed, jid, 16)
        st.append(s0); u32.append(u); f32.append(f)
    out["rng_state0"] = np.array(st, dtype=np.uint64); out["rng_u32"] = np.stack(u32); out["rng_f32"] = np.stack(f32)
    # ---- blackbody table + lookups, HG ----
    out["bb_table"] = ref.blackbody_table()
    temps = np.array([-5, 0, 1, 99.9, 100, 150, 300, 343, 1000, 1234.5, 2450, 6504, 12000, 49800, 49850, 49899.9], np.float32)
    out["bb_temps"] = temps; out["bb_xyz"] = ref.blackbody(temps)
    rs = np.random.default_rng(11)
    w = rs.normal(size=(64, 3)).astype(np.float32); w /= np.linalg.norm(w, axis=1, keepdims=True).astype(np.float32)
    u = rs.random((64, 2)).astype(np.float32)
    out["hg_w"], out["hg_u"] = w, u
    for g in (0.7, 0.4, 0.0, -0.3):
        out[f"hg_dir_g{g}"] = ref.hg_sample(w, u, g)
        out[f"hg_eval_g{g}"] = ref.hg_eval(np.linspace(-1, 1, 33, dtype=np.float32), g)
    # ---- per-config: camera rays, segments, samples, per-job traces, low-spp film ----
    for name in SMALL:
        cfg, size, waves, factor = SMALL[name]
        scene, spec = scenes.small_variant(cfg, size, waves, factor)
        grids = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
        out[f"{name}_density_sha256"] = np.frombuffer(hashlib.sha256(grids.density.view().tobytes()).digest(), np.uint8)
        vol = ref.volume(grids)
        o, d = ray_set(scene, ref)
        out[f"{name}_ray_o"], out[f"{name}_ray_d"] = o, d
        segs, nseg = vol.trace_segments(o, d, 96)
        out[f"{name}_segs"], out[f"{name}_nseg"] = segs, nseg
        lo = np.array([segs[..., 0].min()], np.float32)  # sample points along the rays (index space is internal: use world->index via segments' t is awkward) -> random index positions instead
        pts = (np.random.default_rng(5).random((512, 3)).astype(np.float32) * 2 - 1) * np.array(spec.dims, np.float32)
        v, dm, mj = vol.sample_grid(0, pts)
        out[f"{name}_pts"], out[f"{name}_val"], out[f"{name}_dim"], out[f"{name}_maj"] = pts, v, dm, mj
        jids = np.linspace(0, scene.ntiles * waves - 1, 48).astype(np.uint64)
        xyz, nd, cnt = vol.trace_jobs(scene, jids)
        out[f"{name}_jids"], out[f"{name}_job_xyz"], out[f"{name}_job_ndraws"] = jids, xyz, nd
        out[f"{name}_job_counters"] = np.array(list(cnt.as_dict().values()), np.uint64)
        film = np.zeros((scene.height, scene.width, 4), np.float32)
        _, c = vol.render(scene, film, threads=4)
        out[f"{name}_film"] = film
        out[f"{name}_film_counters"] = np.array(list(c.as_dict().values()), np.uint64)
        out[f"{name}_leafmax_sum"] = np.array([vol.leaf_maxima().astype(np.float64).sum()])
        # ---- Logger<true> + single_pixel path event logs (worker.cpp:15-49,113-119) of a few (job, pixel) pairs ----
        rs = np.random.default_rng(13)
        ntx = -(-scene.width // int(scene.tile_size[0]))
        args, logs = [], []
        for _ in range(8):
            px, py, wave = int(rs.integers(0, scene.width)), int(rs.integers(0, scene.height)), int(rs.integers(1, waves + 1))
            jid = (wave - 1) * scene.ntiles + (py // int(scene.tile_size[1])) * ntx + px // int(scene.tile_size[0])
            ev = vol.path_log(scene, jid, px, py)
            args.append((jid, px, py, len(ev))); logs.append(ev)
        out[f"{name}_pathlog_args"] = np.array(args, np.int64)
        out[f"{name}_pathlog"] = np.concatenate(logs)
        print(name, spec, "segments/ray", nseg.mean(), "film mean", film[..., :3].mean(), c.as_dict())
    np.savez_compressed(HERE / "vectors.npz", **out)
    print("wrote", HERE / "vectors.npz", (HERE / "vectors.npz").stat().st_size, "bytes")


if __name__ == "__main__":
    main()
