"""BASELINE.json's configurations at their FULL sizes, through size-independent properties (the oracle cannot
render them whole in test time) plus an oracle comparison on a sample of (tile, wave) jobs:
  * every pixel receives exactly n_waves samples (sample-count plane == n_waves);
  * waves are additive and individually addressable: [1..2] == [1] + [2] == stride-2 shards;
  * the work counters of a render equal the sum over its waves;
  * a random sample of jobs is bit-equal to the CPU oracle (RNG draw counts and per-pixel radiance).
All five BASELINE.json configurations at their named sizes: the WDAS-shaped cloud at its full 2000x1350x2450 extent (3.6 GB NanoVDB grid,
synthesised once and shared by wdas_cloud and wdas_cloud_4k)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

_CLOUD_FACTOR = float(os.environ.get("VPT_CLOUD_FACTOR", "1.0"))  # development aid: < 1 shrinks the cloud's resolution
CASES = [("fire", 1.0), ("fire_lowscattering", 1.0), ("dense512", 1.0), ("wdas_cloud", _CLOUD_FACTOR), ("wdas_cloud_4k", _CLOUD_FACTOR)]
_volumes = {}  # (kind, dims, voxel_size, seed, anchor) -> grids: the two cloud configurations share one volume


@pytest.mark.parametrize("name,factor", CASES)
def test_fullsize_properties(renderer, oracle, api, scenes, name, factor):
    import torch
    scene, spec = scenes.load_config(name)
    if factor != 1.0:
        spec = spec.scaled(factor)
    key = (spec.kind, tuple(spec.dims), spec.voxel_size, spec.seed, tuple(spec.anchor))
    if key not in _volumes:
        _volumes.clear()
        _volumes[key] = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
        renderer.set_volume(_volumes[key])
    grids = _volumes[key]
    H, W = scene.height, scene.width
    dev = "cuda:0"
    f12 = torch.zeros((H, W, 4), dtype=torch.float32, device=dev)
    f1, f2, fs = torch.zeros_like(f12), torch.zeros_like(f12), torch.zeros_like(f12)
    renderer.enable_counters(True)
    try:
        renderer.run_device(scene, f12.data_ptr(), 1, 2, 1); c12 = renderer.counters().as_dict()
        renderer.run_device(scene, f1.data_ptr(), 1, 1, 1); c1 = renderer.counters().as_dict()
        renderer.run_device(scene, f2.data_ptr(), 2, 1, 1); c2 = renderer.counters().as_dict()
    finally:
        renderer.enable_counters(False)
    renderer.run_device(scene, fs.data_ptr(), 1, 1, 2)  # stride-2 shard: wave 1
    renderer.run_device(scene, fs.data_ptr(), 2, 1, 2)  # stride-2 shard: wave 2
    torch.cuda.synchronize()
    assert bool((f12[..., 3] == 2).all()) and bool((f1[..., 3] == 1).all())
    assert torch.isfinite(f12).all()
    assert torch.allclose(f12, f1 + f2, rtol=1e-5, atol=1e-7) and torch.allclose(fs, f12, rtol=1e-5, atol=1e-7)
    assert {k: c1[k] + c2[k] for k in c1} == c12
    assert c12["pixel_samples"] == 2 * W * H and float(f12[..., :3].sum()) > 0
    # oracle on a sample of jobs from both waves
    rs = np.random.default_rng(5)
    jids = np.sort(rs.choice(scene.ntiles * 2, size=48, replace=False)).astype(np.uint64)
    xo, no, co = oracle.volume(grids).trace_jobs(scene, jids)
    xg, ng = renderer.trace_jobs(scene, jids)
    assert np.array_equal(no, ng), "RNG draw counts differ from the CPU oracle"
    assert np.array_equal(xo, xg), "per-pixel radiance differs from the CPU oracle"
    print(f"\n[{name} x{factor}] {W}x{H}: 48 sampled jobs bit-equal to the oracle; per pixel-sample: "
          f"{c12['density_evals'] / c12['pixel_samples']:.1f} density evals, {c12['majorant_fetches'] / c12['pixel_samples']:.1f} majorant fetches, "
          f"{c12['rng_draws'] / c12['pixel_samples']:.1f} draws")


# (config, GPU waves with the second-moment plane, independent CPU waves): the CPU renders the full frame, so its share is kept to ~20 s
STATS_CASES = [("wdas_cloud", 64, 8), ("fire", 64, 16), ("fire_lowscattering", 64, 16)]


@pytest.mark.parametrize("name,gpu_waves,cpu_waves", STATS_CASES)
def test_fullsize_image_statistics(renderer, oracle, api, scenes, name, gpu_waves, cpu_waves):
    """BASELINE configs 1-3 at FULL resolution: the GPU image against INDEPENDENT waves of the CPU oracle, per pixel inside 3 sigma, relMSE
    stated. sigma comes from the GPU's own second-moment plane (vptc_render_adaptive with rel_error = 0 is the plain render + sum of y^2):
    var(y_gpu_mean - y_cpu_mean) = s^2 (1/Ng + 1/Nc). Radiance is heavy-tailed, so the score sits below the Gaussian 0.9973; the CONTROL is a
    second GPU render of yet other waves scored the same way (per wave the GPU film IS the CPU film, so it is the same estimator)."""
    import json
    scene, spec = scenes.load_config(name)
    if name.startswith("wdas") and _CLOUD_FACTOR != 1.0:
        spec = spec.scaled(_CLOUD_FACTOR)
    key = (spec.kind, tuple(spec.dims), spec.voxel_size, spec.seed, tuple(spec.anchor))
    if key not in _volumes:
        _volumes.clear()
        _volumes[key] = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
        renderer.set_volume(_volumes[key])
    grids = _volumes[key]
    H, W = scene.height, scene.width
    film, sq, tw = renderer.run_adaptive(scene, np.zeros((H, W, 4), np.float32), gpu_waves, gpu_waves, 0.0, first_wave=1000)
    assert (film[..., 3] == gpu_waves).all() and (tw == gpu_waves).all()
    ctrl = renderer.run(scene, np.zeros((H, W, 4), np.float32), first_wave=5000, n_waves=cpu_waves)
    cpu = np.zeros((H, W, 4), np.float32)
    oracle.volume(grids).render(scene, cpu, first_wave=1, n_waves=cpu_waves)
    assert (cpu[..., 3] == cpu_waves).all()
    yg = film[..., 1].astype(np.float64) / gpu_waves
    s2 = np.maximum(sq.astype(np.float64) - gpu_waves * yg * yg, 0.0) / (gpu_waves - 1)
    sigma = np.sqrt(s2 * (1.0 / gpu_waves + 1.0 / cpu_waves))
    floor = 1e-3 * max(yg.mean(), 1e-12)

    def score(other):
        yo = other[..., 1].astype(np.float64) / cpu_waves
        z = np.abs(yo - yg) / np.maximum(sigma, floor)
        e = np.sort(((yo - yg) ** 2 / (yg ** 2 + 1e-4 * yg.mean() ** 2)).ravel())
        # relMSE is carried by a few fireflies (rare bright emission in dark pixels): also state it with the top 0.1 % of the pixels left out
        return float((z <= 3).mean()), float(e.mean()), float(yo.mean()), float(e[: int(len(e) * 0.999)].mean())
    (f_cpu, r_cpu, m_cpu, t_cpu), (f_ctl, r_ctl, m_ctl, t_ctl) = score(cpu), score(ctrl)
    line = {"config": name, "frame": [W, H], "gpu_waves": gpu_waves, "independent_cpu_waves": cpu_waves, "within_3_sigma_cpu": round(f_cpu, 4),
            "within_3_sigma_control": round(f_ctl, 4), "relMSE_cpu": r_cpu, "relMSE_control": r_ctl, "relMSE_trimmed_cpu": t_cpu, "relMSE_trimmed_control": t_ctl, "mean_y_gpu": float(yg.mean()), "mean_y_cpu": m_cpu,
            "mean_y_control": m_ctl}
    print("\n[fullres]", json.dumps(line))
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/fullres_stats.jsonl", "a") as f:
            f.write(json.dumps(line) + "\n")
    assert f_cpu >= 0.95 and abs(f_cpu - f_ctl) <= 0.005
    assert abs(t_cpu - t_ctl) <= 0.1 * max(t_ctl, 1e-6) + 1e-6 and r_ctl / 4 <= r_cpu <= r_ctl * 4
    assert abs(m_cpu - yg.mean()) <= 0.02 * yg.mean()
