"""Parity tests proper: the CUDA path, called through the C ABI (include/vptc.h), against the CPU oracle on the
same seeded inputs and against the committed golden vectors.

Bars (BASELINE.json north_star):
  * integer / index / lookup work (RNG, getValue/getDim/majorant, trilinear, HDDA segments): BIT-EXACT;
  * per-path radiance with identical RNG streams: relative error <= 1e-4 wherever branch decisions agree
    (agreement = the job's per-pixel RNG draw counts all match), agreeing fraction reported and bounded below;
  * images: sample counts exact; per-pixel difference inside 3 sigma of the Monte-Carlo noise; relMSE stated.
"""
import numpy as np
import pytest

from conftest import GOLDEN, SMALL

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4  # north_star: per-path radiance within 1e-4 relative where branch decisions agree


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLDEN / "vectors.npz")


def rel_err(a, b):
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), 1e-12)


@pytest.fixture(scope="module")
def loaded(renderer, small_case):
    """Set the renderer's volume lazily per case (one context, re-baked when the case changes)."""
    state = {"name": None}

    def use(name):
        scene, spec, grids = small_case(name)
        if state["name"] != name:
            renderer.set_volume(grids)
            state["name"] = name
        return scene, spec, grids
    return use


@pytest.mark.parametrize("name", list(SMALL))
def test_grid_lookups_bit_exact(renderer, oracle, loaded, gold, name):
    """getValue-based trilinear sampler, getDim and the patched majorant: brick layout == NanoVDB tree."""
    scene, spec, grids = loaded(name)
    v, dm, mj = renderer.sample_grid(0, gold[f"{name}_pts"])
    assert np.array_equal(v, gold[f"{name}_val"]) and np.array_equal(dm, gold[f"{name}_dim"]) and np.array_equal(mj, gold[f"{name}_maj"])
    rs = np.random.default_rng(42)
    pts = (rs.random((200000, 3)).astype(np.float32) * 2.2 - 1.1) * np.array(spec.dims, np.float32)
    vol = oracle.volume(grids)
    for got, want in zip(renderer.sample_grid(0, pts), vol.sample_grid(0, pts)):
        assert np.array_equal(got, want)
    if grids.has_temperature():
        for got, want in zip(renderer.sample_grid(1, pts)[:1], vol.sample_grid(1, pts)[:1]):
            assert np.array_equal(got, want)


@pytest.mark.parametrize("name", list(SMALL))
def test_majorant_segments_bit_exact(renderer, oracle, loaded, gold, name):
    """Volume::intersect + RayMajorantIterator::next (HDDA at dims 8/128/4096, merging): t0/t1/d_maj bit-equal."""
    scene, spec, grids = loaded(name)
    segs, n = renderer.trace_segments(gold[f"{name}_ray_o"], gold[f"{name}_ray_d"], 96)
    assert np.array_equal(n, gold[f"{name}_nseg"]) and np.array_equal(segs, gold[f"{name}_segs"])
    rs = np.random.default_rng(7)
    px, py = rs.integers(0, scene.width, 4000), rs.integers(0, scene.height, 4000)
    o, d = oracle.camera_rays(scene, px, py, (rs.random(4000) * 0.5).astype(np.float32), (rs.random(4000) * 0.5).astype(np.float32))
    # plus rays starting inside the volume in random directions (shadow/bounce rays)
    c = np.array(spec.anchor, np.float32) + (np.array([0, spec.dims[1] * spec.voxel_size / 2, 0], np.float32) if spec.kind == "fire" else 0)
    o2 = c + (rs.random((4000, 3)).astype(np.float32) - 0.5) * np.array(spec.dims, np.float32) * np.float32(spec.voxel_size)
    d2 = rs.normal(size=(4000, 3)).astype(np.float32); d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    o, d = np.concatenate([o, o2.astype(np.float32)]), np.concatenate([d, d2.astype(np.float32)])
    want_s, want_n = oracle.volume(grids).trace_segments(o, d, 128)
    got_s, got_n = renderer.trace_segments(o, d, 128)
    assert np.array_equal(got_n, want_n) and np.array_equal(got_s, want_s)
    assert want_n.max() >= 2


def test_tiles_volume_bit_exact(renderer, oracle, api):
    """Active/inactive tiles at every tree level: getDim in {8,128,4096}, probeValue majorants, tile values."""
    grids = api.synth_volume("tiles", (600, 300, 300), 1.0, 3, None)
    renderer.set_volume(grids)
    vol = oracle.volume(grids)
    rs = np.random.default_rng(1)
    pts = (rs.random((100000, 3)).astype(np.float32) * 2 - 0.5) * np.array([600, 300, 300], np.float32)
    got, want = renderer.sample_grid(0, pts), vol.sample_grid(0, pts)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    assert set(np.unique(want[1]).tolist()) >= {8, 128}
    o = (rs.random((5000, 3)).astype(np.float32) * 3 - 1) * np.array([600, 300, 300], np.float32)
    d = rs.normal(size=(5000, 3)).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
    gs, gn = renderer.trace_segments(o, d, 128); ws, wn = vol.trace_segments(o, d, 128)
    assert np.array_equal(gn, wn) and np.array_equal(gs, ws)


# minimum fraction of JOBS (64 pixels sharing one RNG stream) whose draw counts all agree with the oracle
# (the device's logf/sinf/cosf are glibc-exact — csrc/device/libm_exact.cuh — so every job agrees in practice)
MIN_AGREE = {"fire": 0.999, "fire_lowscattering": 0.999, "wdas_cloud": 0.999, "dense512": 0.999}


@pytest.mark.parametrize("name", list(SMALL))
def test_per_job_radiance_parity(renderer, oracle, loaded, gold, name):
    """Identical pcg32_fast streams => per-path radiance within 1e-4 wherever branch decisions agree."""
    scene, spec, grids = loaded(name)
    jids = np.arange(scene.ntiles * scene.num_waves, dtype=np.uint64)
    xo, no, co = oracle.volume(grids).trace_jobs(scene, jids)
    xg, ng = renderer.trace_jobs(scene, jids)
    agree = (no == ng).all(axis=1)
    frac = agree.mean()
    # pixel-level agreement: pixels before the first diverging pixel of a job still share the stream
    first_bad = np.where(no != ng, np.arange(no.shape[1])[None, :], no.shape[1]).min(axis=1)
    pix_ok = np.arange(no.shape[1])[None, :] < first_bad[:, None]
    used = no > 0
    err = rel_err(xo, xg).max(axis=2)
    worst = err[pix_ok & used].max()
    print(f"\n[{name}] jobs {len(jids)}: draw-count agreement {frac:.4f} of jobs, {pix_ok[used].mean():.4f} of pixels; "
          f"max rel radiance err on agreeing pixels {worst:.3g}; bit-equal {(xo[pix_ok & used] == xg[pix_ok & used]).all(axis=-1).mean():.4f}")
    assert frac >= MIN_AGREE[name]
    assert worst <= REL_TOL
    # golden subset (minted from the vendored-NanoVDB oracle build)
    gj = gold[f"{name}_jids"]
    sel = np.searchsorted(jids, gj)
    ok = (ng[sel] == gold[f"{name}_job_ndraws"]).all(axis=1)
    assert rel_err(xg[sel][ok], gold[f"{name}_job_xyz"][ok]).max() <= REL_TOL


@pytest.mark.parametrize("name", list(SMALL))
def test_film_and_counters(renderer, oracle, loaded, gold, name):
    """vptc_render adds (imaging_ratio*L, 1) per sample: counts exact, sums close, work counters equal the oracle's."""
    scene, spec, grids = loaded(name)
    renderer.enable_counters(True)
    film = renderer.run(scene, renderer.new_film(scene))
    c = renderer.counters().as_dict()
    renderer.enable_counters(False)
    want = gold[f"{name}_film"]
    assert np.array_equal(film[..., 3], want[..., 3]) and (film[..., 3] == scene.num_waves).all()
    assert np.isfinite(film).all()
    wc = dict(zip(c.keys(), gold[f"{name}_film_counters"].tolist()))
    assert c["pixel_samples"] == wc["pixel_samples"]
    for k in ("density_evals", "majorant_fetches", "rng_draws", "temperature_evals"):
        assert c[k] == wc[k], (k, c[k], wc[k])  # same branch decisions => same work, exactly
    # most pixels are bit-close; the rest differ by whole paths where a branch flipped
    close = np.isclose(film[..., :3], want[..., :3], rtol=1e-4, atol=1e-7).all(axis=-1)
    assert close.mean() >= MIN_AGREE[name] * 0.9
    assert abs(film[..., :3].mean() - want[..., :3].mean()) <= 0.05 * want[..., :3].mean()
    # without counters the film is the same up to atomic add order
    film2 = renderer.run(scene, renderer.new_film(scene))
    assert np.allclose(film2, film, rtol=1e-5, atol=1e-7)


def test_render_is_additive_and_wave_addressed(renderer, loaded):
    """ADDS into the caller's film; waves are 1-based and independent: [1..4] == [1..2] + [3..4] == strided shards."""
    scene, spec, grids = loaded("fire")
    s = scene.copy(); s.num_waves = 4
    full = renderer.run(s, renderer.new_film(s))
    parts = renderer.new_film(s)
    renderer.run(s, parts, first_wave=1, n_waves=2)
    renderer.run(s, parts, first_wave=3, n_waves=2)
    assert np.array_equal(parts[..., 3], full[..., 3])
    assert np.allclose(parts, full, rtol=1e-5, atol=1e-7)
    empty = renderer.run(s, renderer.new_film(s), first_wave=1, n_waves=0)
    assert (empty == 0).all()
    # device-film path with wave_stride (multi-GPU sharding of a progressive render)
    import torch
    d = torch.zeros((s.height, s.width, 4), dtype=torch.float32, device="cuda:0")
    renderer.run_device(s, d.data_ptr(), 1, 2, 2)  # waves 1, 3
    renderer.run_device(s, d.data_ptr(), 2, 2, 2)  # waves 2, 4
    torch.cuda.synchronize()
    assert np.allclose(d.cpu().numpy(), full, rtol=1e-5, atol=1e-7)


def test_edge_cases(renderer, oracle, api, scenes, loaded):
    scene, spec, grids = loaded("fire")
    vol = oracle.volume(grids)
    # ragged tiles (image not a multiple of the tile size; non-square tiles)
    s = scene.copy()
    s.output_size[0], s.output_size[1] = 37, 29
    s.tile_size[0], s.tile_size[1] = 8, 5
    film = renderer.run(s, renderer.new_film(s))
    ref = np.zeros_like(film); vol.render(s, ref)
    assert np.array_equal(film[..., 3], ref[..., 3]) and (film[..., 3] == s.num_waves).all()
    jids = np.arange(s.ntiles, dtype=np.uint64)
    xo, no, _ = vol.trace_jobs(s, jids); xg, ng = renderer.trace_jobs(s, jids)
    ok = (no == ng).all(axis=1)
    assert ok.mean() > 0.9 and rel_err(xo[ok], xg[ok]).max() <= REL_TOL
    # single_pixel: only that pixel is touched, its job's stream starts at it (worker.cpp:113-119)
    s = scene.copy()
    s.worker_parameters.single_pixel_enabled = 1
    s.worker_parameters.single_pixel_coord[0], s.worker_parameters.single_pixel_coord[1] = 35, 50
    film = renderer.run(s, renderer.new_film(s))
    ref = np.zeros_like(film); vol.render(s, ref)
    assert film[..., 3].sum() == s.num_waves and film[50, 35, 3] == s.num_waves
    assert np.allclose(film, ref, rtol=1e-4, atol=1e-7)
    # 1x1 tiles: literally per-pixel, per-sample seeding
    s = scene.copy()
    s.output_size[0], s.output_size[1] = 24, 16
    s.tile_size[0], s.tile_size[1] = 1, 1
    jids = np.arange(s.ntiles * 2, dtype=np.uint64)
    xo, no, _ = vol.trace_jobs(s, jids); xg, ng = renderer.trace_jobs(s, jids)
    ok = (no == ng).all(axis=1)
    assert ok.mean() > 0.97 and rel_err(xo[ok], xg[ok]).max() <= REL_TOL
    # no distant light (Li == 0: sample_Ld returns before any draw), no jitter, tiny max_depth
    s = scene.copy()
    s.worker_parameters.distant_light.multiplier = 0.0
    s.worker_parameters.use_jitter = 0
    s.worker_parameters.max_depth = 3
    jids = np.arange(0, s.ntiles * 2, 3, dtype=np.uint64)
    xo, no, _ = vol.trace_jobs(s, jids); xg, ng = renderer.trace_jobs(s, jids)
    ok = (no == ng).all(axis=1)
    assert ok.mean() > 0.95 and rel_err(xo[ok], xg[ok]).max() <= REL_TOL
    # camera that misses the volume entirely: every path returns the infinite light
    s = scene.copy()
    s.camera_parameters.look[0], s.camera_parameters.look[1], s.camera_parameters.look[2] = 240.0, 30.0, 0.0
    film = renderer.run(s, renderer.new_film(s))
    inf = np.array(s.worker_parameters.infinite_light.xyz, np.float32) * np.float32(s.worker_parameters.infinite_light.multiplier)
    want = np.float32(s.camera_parameters.imaging_ratio) * inf * np.float32(s.num_waves)
    assert np.allclose(film[..., :3], want[None, None, :], rtol=1e-6)


def test_more_edge_cases(renderer, oracle, api, scenes, loaded):
    """Corners of the parameter space the reference code branches on; the GPU must equal the oracle draw for draw."""
    scene, spec, grids = loaded("fire")
    vol = oracle.volume(grids)

    def same(s, jids):
        xo, no, _ = vol.trace_jobs(s, jids); xg, ng = renderer.trace_jobs(s, jids)
        assert np.array_equal(no, ng)
        assert rel_err(xo, xg).max() <= REL_TOL

    some = np.arange(0, scene.ntiles * 2, 5, dtype=np.uint64)
    # n_waves == 0 renders nothing; so does a wave range past num_waves' end — waves are only addresses (tile_provider.cpp:30)
    film = renderer.new_film(scene)
    renderer.run(scene, film, first_wave=1, n_waves=0)
    assert not film.any()
    renderer.run(scene, film, first_wave=1000, n_waves=1)
    ref = np.zeros_like(film); vol.render(scene, ref, first_wave=1000, n_waves=1)
    assert np.array_equal(film[..., 3], ref[..., 3]) and np.allclose(film, ref, rtol=1e-5, atol=1e-7)
    # isotropic limit of the phase function: |g| < 1e-3 takes the `1 - 2u` branch (random.hpp:60-62)
    s = scene.copy(); s.volume_parameters.henyey_greenstein_g = 5.0e-4
    same(s, some)
    # backward scattering, and a strongly forward lobe
    for g in (-0.6, 0.95):
        s = scene.copy(); s.volume_parameters.henyey_greenstein_g = g
        same(s, some)
    # max_depth == 0: the bounce loop never runs, every path is the infinite light (worker.cpp:130,198-200)
    s = scene.copy(); s.worker_parameters.max_depth = 0
    same(s, some)
    film = renderer.run(s, renderer.new_film(s))
    inf = np.array(s.worker_parameters.infinite_light.xyz, np.float32) * np.float32(s.worker_parameters.infinite_light.multiplier)
    assert np.allclose(film[..., :3], (np.float32(s.camera_parameters.imaging_ratio) * inf * np.float32(s.num_waves))[None, None, :], rtol=1e-6)
    # max_depth == 1: one bounce, then out of depth
    s = scene.copy(); s.worker_parameters.max_depth = 1
    same(s, some)
    # purely absorbing medium (sigma_s = 0): every real collision terminates the path (worker.cpp:163-166)
    s = scene.copy(); s.volume_parameters.sigma_s = 0.0
    same(s, some)
    # camera INSIDE the volume: rays start in the middle of the grid (Ray::clip with t0 inside the box)
    s = scene.copy()
    s.camera_parameters.position[0], s.camera_parameters.position[1], s.camera_parameters.position[2] = 0.5, 20.0, 0.25
    same(s, some)
    # a tile larger than the image: one job is the whole frame, clipped (tile_provider.cpp:69-85)
    s = scene.copy()
    s.output_size[0], s.output_size[1] = 21, 13
    s.tile_size[0], s.tile_size[1] = 32, 32
    assert s.ntiles == 1
    same(s, np.arange(3, dtype=np.uint64))
    film = renderer.run(s, renderer.new_film(s))
    assert (film[..., 3] == s.num_waves).all()
    # an axis-parallel view direction: zero direction components, the HDDA's dir == 0 arms (HDDA.h:60-75)
    s = scene.copy()
    s.camera_parameters.position[0], s.camera_parameters.position[1], s.camera_parameters.position[2] = 120.0, 30.0, 0.0
    s.camera_parameters.look[0], s.camera_parameters.look[1], s.camera_parameters.look[2] = 0.0, 30.0, 0.0
    s.worker_parameters.use_jitter = 0
    s.output_size[0], s.output_size[1] = 33, 33  # odd size: the centre pixel looks exactly along -x
    s.tile_size[0], s.tile_size[1] = 8, 8
    same(s, np.arange(s.ntiles, dtype=np.uint64))


def test_tiny_volume_mostly_missed(renderer, oracle, api, scenes, loaded):
    """A volume of a handful of leaves far smaller than the view: most rays miss the grid's bounding box
    (Volume::intersect -> nullopt, volume.cpp:84-86), the rest cross one or two nodes."""
    scene, spec, grids = loaded("fire")
    tiny = api.synth_volume("torus", (12, 12, 12), 1.0, 1, None)
    try:
        renderer.set_volume(tiny)
        vol = oracle.volume(tiny)
        s = scene.copy()
        s.camera_parameters.look[0], s.camera_parameters.look[1], s.camera_parameters.look[2] = 0.0, 0.0, 0.0
        s.camera_parameters.position[0], s.camera_parameters.position[1], s.camera_parameters.position[2] = 60.0, 10.0, 5.0
        jids = np.arange(0, s.ntiles * 2, dtype=np.uint64)
        xo, no, co = vol.trace_jobs(s, jids); xg, ng = renderer.trace_jobs(s, jids)
        assert np.array_equal(no, ng) and rel_err(xo, xg).max() <= REL_TOL
        assert co.density_evals > 0 and (no[:, 1:] - no[:, :-1] == 2).any()  # some paths sample the medium; many draw only their 2 jitters
    finally:
        renderer.set_volume(grids)


def test_density_only_volume_skips_emission(renderer, oracle, api, loaded):
    """No temperature grid => the emission term is skipped entirely (worker.cpp:152)."""
    scene, spec, grids = loaded("fire")
    only_density = api.VolumeGrids(grids.density, None)
    renderer.set_volume(only_density)
    try:
        vol = oracle.volume(only_density)
        jids = np.arange(0, scene.ntiles * 2, 2, dtype=np.uint64)
        xo, no, co = vol.trace_jobs(scene, jids); xg, ng = renderer.trace_jobs(scene, jids)
        assert co.temperature_evals == 0
        ok = (no == ng).all(axis=1)
        assert ok.mean() > 0.97 and rel_err(xo[ok], xg[ok]).max() <= REL_TOL
    finally:
        renderer.set_volume(grids)


def test_film_to_image_device_byte_exact(renderer, oracle, gold):
    """XYZ/N -> sRGB8 on the device == film_to_image on the host (main.cpp:12-24, color.hpp:8-30), BYTE FOR BYTE: the OETF's
    powf(x, 1/2.4f) is glibc's scheme on the device (libm_exact.cuh, exhaustively checked against glibc over its whole domain)."""
    for name in SMALL:
        film = gold[f"{name}_film"].copy()
        assert np.array_equal(renderer.film_to_image(film), oracle.film_to_image(film))
    rs = np.random.default_rng(11)
    film = np.zeros((256, 256, 4), np.float32)
    film[..., :3] = (rs.gamma(0.4, 0.5, (256, 256, 3)) * 16).astype(np.float32)
    film[..., 3] = 16
    film[0, :, 3] = 0                        # never sampled: 0/0 = NaN -> 0
    film[1, :, :3] = 0                       # black
    film[2, :, :3] = 3.0e38                  # overflows to +inf in the matrix product
    film[3, :, :3] = np.linspace(0.0031308 * 15.9, 0.0031308 * 16.1, 256, dtype=np.float32)[:, None]  # around the OETF's knee
    film[4, :, :3] = np.linspace(15.99, 16.01, 256, dtype=np.float32)[:, None]                          # around 1.0 (254 / 255)
    film[5, :, 0] = -1.0                     # out-of-gamut negatives
    got, want = renderer.film_to_image(film), oracle.film_to_image(film)
    assert np.array_equal(got, want)
    assert got[0].max() == 0 and got[2].min() == 255 and got[4, :, 0].min() == 255


def test_blackbody_device_bit_exact(renderer, oracle, gold):
    """The DEVICE's blackbody_radiation_xyz (precompute_blackbody.cpp:24-52) == the oracle's, bit for bit: table lerp, exact hits, the pad
    entry for 49 800 K < T < 49 900 K, T <= 0, non-finite, and the direct 471-term integration (spectral.cpp:7-20) for T >= 49 900 K —
    glibc-exact expf on the device, lambda^5 from the host's libm."""
    rs = np.random.default_rng(2)
    T = np.concatenate([gold["bb_temps"], rs.uniform(1, 49800, 20000), np.arange(-100, 50100, 100), rs.uniform(49900, 2e5, 4000), 10 ** rs.uniform(4.7, 38.4, 4000),
                        [49800.5, 49899.9, 49899.996, 49900, 49900.004, 6e4, 1e5, 1e6, 3.0e38, np.inf, -np.inf, np.nan, -1, 0, 1e-30, 99.99, 100, 100.01]]).astype(np.float32)
    got, want = renderer.blackbody(T), oracle.blackbody(T)
    assert np.array_equal(got, want, equal_nan=True)
    assert np.isnan(got[np.isinf(T) | np.isnan(T)]).all() and (got[(T <= 0) & np.isfinite(T)] == 0).all()
    assert (got[(T >= 49900) & (T < 1e6)] > 0).all()


def test_emission_through_the_direct_blackbody_branch(renderer, oracle, loaded):
    """temperature_scale = 2000 lifts the fire's temperatures (0..~50) to 300..100 000 K: collisions take the table, the pad entry and the
    direct-integration branch inside the render kernel; per-job radiance and draw counts stay bit-equal to the oracle."""
    scene, spec, grids = loaded("fire")
    hot = scene.copy()
    hot.volume_parameters.temperature_scale = 2000.0
    hot.volume_parameters.le_scale = 1e-13
    jids = np.arange(0, scene.ntiles * 2, 3, dtype=np.uint64)
    xo, no, co = oracle.volume(grids).trace_jobs(hot, jids)
    xg, ng = renderer.trace_jobs(hot, jids)
    assert co.temperature_evals > 0
    assert np.array_equal(no, ng) and np.array_equal(xo, xg)
    cold_x, _, _ = oracle.volume(grids).trace_jobs(scene, jids)
    assert not np.array_equal(cold_x, xo), "the hot scene must actually change the emission"


def test_errors(renderer, api, scenes):
    scene, _ = scenes.small_variant("fire", (16, 16), 1, 0.2)
    r2 = api.Renderer(0)
    with pytest.raises(api.VptError) as e:
        r2.run(scene, r2.new_film(scene))
    assert e.value.code == -4  # VPTC_ERR_NO_VOLUME
    bad = scene.copy(); bad.tile_size[0] = 0
    with pytest.raises(api.VptError):
        renderer.run(bad, renderer.new_film(scene))
    huge = scene.copy(); huge.output_size[0] = huge.output_size[1] = 2 ** 31  # W*H*16 wraps to 0 in 64 bits
    with pytest.raises(api.VptError) as e:
        renderer.run_device(huge, 16, 1, 1, 1)
    assert e.value.code == -1
    junk = np.zeros(4096, np.uint8)
    assert api.lib().vptc_set_volume(r2._ctx, junk.ctypes.data, junk.nbytes, None, 0) == -2
    r2.close()


def test_failed_set_volume_keeps_the_previous_volume(renderer, oracle, loaded):
    """A rejected second vptc_set_volume (corrupt buffer) must leave the context rendering the FIRST volume (round-1 advisor finding:
    the old volume used to be freed before the new one was validated and staged)."""
    scene, spec, grids = loaded("fire")
    jids = np.arange(8, dtype=np.uint64)
    before = renderer.trace_jobs(scene, jids)
    bad = grids.density.view().copy()
    bad[296 + 36:296 + 72] = 0  # singular inverse map
    import ctypes
    rc = renderer_api_set_volume(renderer, bad)
    assert rc == -2
    after = renderer.trace_jobs(scene, jids)
    assert np.array_equal(before[0], after[0]) and np.array_equal(before[1], after[1])


def renderer_api_set_volume(renderer, density_bytes: np.ndarray) -> int:
    from volume_path_tracer_b200 import api
    return api.lib().vptc_set_volume(renderer._ctx, density_bytes.ctypes.data, density_bytes.nbytes, None, 0)


@pytest.mark.parametrize("name,waves", [("fire", 48), ("wdas_cloud", 48)])
def test_image_within_monte_carlo_noise(renderer, oracle, loaded, name, waves):
    """Converged-image criterion at test size: the GPU film of waves 1000.. against the CPU oracle film of waves
    1.. (INDEPENDENT samples): per-pixel difference inside 3 sigma, relMSE stated. sigma is estimated per pixel from
    the per-wave luminance of both sets. Path radiance is heavy-tailed (rare bright emission / sun hits), so the
    fraction inside 3 estimated sigma sits below the Gaussian 0.9973; the CONTROL is the CPU oracle rendering waves
    1000.. itself: the GPU must score like the control (it is the same estimator), and both above 0.95."""
    scene, spec, grids = loaded(name)
    vol = oracle.volume(grids)
    s = scene.copy()
    H, W = s.height, s.width

    def per_wave_cpu(first):
        pw = np.zeros((waves, H, W), np.float64)
        for k in range(waves):
            f = np.zeros((H, W, 4), np.float32)
            vol.render(s, f, first_wave=first + k, n_waves=1)
            pw[k] = f[..., 1]
        return pw
    a, c = per_wave_cpu(1), per_wave_cpu(1000)
    gpu = renderer.run(s, renderer.new_film(s), first_wave=1000, n_waves=waves)
    assert (gpu[..., 3] == waves).all()
    ma, mc, mg = a.mean(axis=0), c.mean(axis=0), gpu[..., 1].astype(np.float64) / waves
    var = np.concatenate([a, c]).var(axis=0, ddof=1)
    sigma = np.sqrt(2 * var / waves)
    floor = 1e-3 * max(ma.mean(), 1e-12)  # pixels with ~zero sample variance (all paths identical)

    def score(m):
        z = np.abs(m - ma) / np.maximum(sigma, floor)
        return (z <= 3).mean(), np.mean((m - ma) ** 2 / (ma ** 2 + 1e-4 * ma.mean() ** 2))
    (f_gpu, r_gpu), (f_cpu, r_cpu) = score(mg), score(mc)
    print(f"\n[{name}] {W}x{H}, {waves} vs {waves} independent waves: GPU within 3 sigma {f_gpu:.4f} (CPU control {f_cpu:.4f}), "
          f"relMSE GPU {r_gpu:.4g} (control {r_cpu:.4g}), mean cpu {ma.mean():.5f} gpu {mg.mean():.5f}")
    assert f_gpu >= 0.95 and abs(f_gpu - f_cpu) <= 0.003
    assert abs(r_gpu - r_cpu) <= 0.02 * max(r_cpu, 1e-6)
    # same waves on both sides: the GPU film IS the CPU film up to fp32 add order
    assert np.allclose(mg, mc, rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("name", ["fire", "wdas_cloud", "dense512"])
def test_path_event_log_bit_exact(renderer, oracle, loaded, gold, name):
    """vptc_trace_path_log = Logger<true> + single_pixel (worker.cpp:15-49,113-119): the device's event list of one camera path
    (new_ray / sampled_point / null / scatter / absorbed with their arguments) equals the oracle's, float for float."""
    scene, spec, grids = loaded(name)
    vol = oracle.volume(grids)
    ntx = -(-scene.width // int(scene.tile_size[0]))
    rs = np.random.default_rng(11)
    total = 0
    for _ in range(24):
        px, py, wave = int(rs.integers(0, scene.width)), int(rs.integers(0, scene.height)), int(rs.integers(1, scene.num_waves + 1))
        jid = (wave - 1) * scene.ntiles + (py // int(scene.tile_size[1])) * ntx + px // int(scene.tile_size[0])
        want = vol.path_log(scene, jid, px, py)
        got = renderer.trace_path_log(scene, jid, px, py)
        assert got.shape == want.shape and np.array_equal(got, want)
        total += len(want)
    assert total > 24  # some paths entered the volume
    off = 0  # and the committed fixtures minted from the vendored-NanoVDB oracle build
    for gj, gx, gy, n in gold[f"{name}_pathlog_args"]:
        assert np.array_equal(renderer.trace_path_log(scene, int(gj), int(gx), int(gy)), gold[f"{name}_pathlog"][off:off + n])
        off += int(n)
    # capacity smaller than the log: the true count is still reported (here through the clipped array length == cap)
    assert len(renderer.trace_path_log(scene, jid, px, py, cap=1)) == 1


@pytest.mark.parametrize("tile", [(1, 1), (2, 2), (4, 4), (3, 5), (16, 2), (96, 54)])
@pytest.mark.parametrize("name", ["fire", "dense512"])
def test_tile_sizes_bit_exact(renderer, oracle, loaded, name, tile):
    """tile_size is a free scene parameter (tile_provider.cpp:95-105): per-pixel jobs (1x1), small tiles (the drain-free
    settings of DESIGN.md 4.3), non-square ragged tiles and one tile covering the whole frame all reproduce the oracle's
    per-pixel radiance and draw counts bit for bit, in every kernel variant."""
    scene0, spec, grids = loaded(name)
    scene = type(scene0).from_buffer_copy(scene0)
    scene.tile_size[0], scene.tile_size[1] = tile
    jids = np.arange(scene.ntiles * scene.num_waves, dtype=np.uint64)
    xo, no, co = oracle.volume(grids).trace_jobs(scene, jids)
    film_o = np.zeros((scene.height, scene.width, 4), np.float32)
    oracle.volume(grids).render(scene, film_o)
    try:
        for v in (2, 3):
            renderer.set_kernel_variant(v)
            xg, ng = renderer.trace_jobs(scene, jids)
            assert np.array_equal(ng, no) and np.array_equal(xg, xo)
            f = renderer.run(scene, renderer.new_film(scene))
            assert np.array_equal(f[..., 3], film_o[..., 3]) and np.allclose(f, film_o, rtol=1e-5, atol=1e-7)
    finally:
        renderer.set_kernel_variant(2)
    assert (film_o[..., 3] == scene.num_waves).all()


@pytest.mark.parametrize("name", ["fire", "wdas_cloud", "dense512"])
def test_kernel_variants_agree(renderer, loaded, name):
    """The nested (0), flat (1), stage-scheduled (2) and job-pool (3) megakernels are four schedules of the same
    arithmetic: per-job radiance, RNG draw counts and work counters must be bit-identical."""
    scene, spec, grids = loaded(name)
    jids = np.arange(scene.ntiles * scene.num_waves, dtype=np.uint64)
    res = []
    try:
        renderer.enable_counters(True)
        for v in (0, 1, 2, 3):
            renderer.set_kernel_variant(v)
            x, n = renderer.trace_jobs(scene, jids)
            f = renderer.run(scene, renderer.new_film(scene))
            res.append((x, n, f, renderer.counters().as_dict()))
    finally:
        renderer.enable_counters(False)
        renderer.set_kernel_variant(2)
    for x, n, f, c in res[1:]:
        assert np.array_equal(n, res[0][1]) and np.array_equal(x, res[0][0])
        assert c == res[0][3]
        assert np.array_equal(f[..., 3], res[0][2][..., 3]) and np.allclose(f, res[0][2], rtol=1e-5, atol=1e-7)


# ---------------------------------------------------------------- asynchronous calls, job slices (device-side range ring)
def test_async_calls_equal_blocking_calls(renderer, loaded):
    """vptc_render_async publishes each call into the device ring and the running kernel carries on into it; whatever the interleaving,
    call k's film must hold exactly call k's waves. One wave per call has no add-order slack: every film is bit-equal to a blocking call's."""
    scene, spec, grids = loaded("dense512")
    waves = [1, 2, 3, 5, 8, 13, 21, 34]
    blocking = [renderer.run(scene, renderer.new_film(scene), first_wave=w, n_waves=1) for w in waves]
    films = [renderer.new_film(scene) for _ in waves]
    for w, f in zip(waves, films):
        renderer.run_async(scene, f, w, 1)
    renderer.wait()
    for a, b in zip(films, blocking):
        assert np.array_equal(a, b)
    # several waves per call into ONE accumulating host film, like the CLI's progressive loop
    total = renderer.new_film(scene)
    for first in range(1, 25, 4):
        renderer.run_async(scene, total, first, 4)
    renderer.wait()
    want = renderer.run(scene, renderer.new_film(scene), first_wave=1, n_waves=24)
    assert np.array_equal(total[..., 3], want[..., 3]) and (total[..., 3] == 24).all()
    assert np.allclose(total, want, rtol=1e-5, atol=1e-7)
    # a different scene in the middle of a series: the series is finished first, results stay per call
    other = scene.copy(); other.seed = scene.seed + 1
    fa, fb, fc = renderer.new_film(scene), renderer.new_film(scene), renderer.new_film(scene)
    renderer.run_async(scene, fa, 1, 1); renderer.run_async(other, fb, 1, 1); renderer.run_async(scene, fc, 2, 1)
    renderer.wait()
    assert np.array_equal(fa, blocking[0]) and np.array_equal(fc, blocking[1])
    assert np.array_equal(fb, renderer.run(other, renderer.new_film(other), first_wave=1, n_waves=1)) and not np.array_equal(fb, fa)
    renderer.wait()  # idempotent


def test_device_async_series_and_tile_stripes(renderer, loaded):
    """vptc_render_device_async into one device film, and vptc_job_range tile stripes: any partition of the (tile, wave) job space — by
    waves, by stripes of tiles, across calls — renders exactly the jobs of the whole (job ids are (wave-1)*ntiles + tile whatever the slice)."""
    import torch
    scene, spec, grids = loaded("fire")
    H, W, nt = scene.height, scene.width, scene.ntiles
    whole = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda:0")
    renderer.run_device(scene, whole.data_ptr(), 1, 1, 1)
    parts = torch.zeros_like(whole)
    cuts = [0, 1, nt // 3, nt // 3, nt - 5, nt]  # includes an empty stripe
    for a, b in zip(cuts[:-1], cuts[1:]):
        if b > a:
            renderer.run_device_range(scene, parts.data_ptr(), 1, 1, 1, tile_begin=a, tile_count=b - a, wait=False)
    renderer.wait()
    torch.cuda.synchronize()
    assert torch.equal(parts, whole)  # one wave: bit-equal
    # a stripe touches only its own pixels
    stripe = torch.zeros_like(whole)
    ntx = -(-W // int(scene.tile_size[0]))
    renderer.run_device_range(scene, stripe.data_ptr(), 1, 2, 1, tile_begin=ntx, tile_count=2 * ntx)  # tile rows 1 and 2
    th = int(scene.tile_size[1])
    assert bool((stripe[th:3 * th, :, 3] == 2).all()) and float(stripe[:th].abs().sum()) == 0 and float(stripe[3 * th:].abs().sum()) == 0
    # interleaved stripes (tile_row_stride): three "devices" taking every third row of tiles cover the frame exactly once
    nty = -(-H // th)
    inter = torch.zeros_like(whole)
    for d in range(3):
        rows = len(range(d, nty, 3))
        renderer.run_device_range(scene, inter.data_ptr(), 1, 1, 1, tile_begin=d * ntx, tile_count=rows * ntx, tile_row_stride=3, wait=False)
    renderer.wait()
    torch.cuda.synchronize()
    assert torch.equal(inter, whole)
    with pytest.raises(Exception):
        renderer.run_device_range(scene, inter.data_ptr(), 1, 1, 1, tile_begin=2 * ntx, tile_count=nty * ntx, tile_row_stride=3)
    # waves as a series of async calls, interleaved strides
    series = torch.zeros_like(whole)
    for first in (1, 2, 3):
        renderer.run_device_range(scene, series.data_ptr(), first, 4, 3, wait=False)  # waves first, first+3, ... : 12 waves in all
    renderer.wait()
    ref = torch.zeros_like(whole)
    renderer.run_device(scene, ref.data_ptr(), 1, 12, 1)
    torch.cuda.synchronize()
    assert torch.equal(series[..., 3], ref[..., 3]) and torch.allclose(series, ref, rtol=1e-5, atol=1e-7)
    with pytest.raises(Exception):
        renderer.run_device_range(scene, series.data_ptr(), 1, 1, 1, tile_begin=nt - 1, tile_count=5)


# ---------------------------------------------------------------- adaptive wave count
def test_adaptive_render_contract(renderer, loaded):
    """vptc_render_adaptive: rel_error 0 is the plain render plus second moments; a huge rel_error stops every tile after the first eligible
    batch; in between, a tile's film is the plain render truncated at the wave the tile stopped (job ids are unchanged)."""
    scene, spec, grids = loaded("fire")
    ntx = -(-scene.width // int(scene.tile_size[0]))
    # one wave, never stopping: the film is the plain wave, and sum_y2 is exactly the square of the one luminance sample
    f1, sq1, tw1 = renderer.run_adaptive(scene, renderer.new_film(scene), max_waves=1, batch_waves=1, rel_error=0.0)
    plain1 = renderer.run(scene, renderer.new_film(scene), first_wave=1, n_waves=1)
    assert np.array_equal(f1, plain1) and (tw1 == 1).all()
    y2 = plain1[..., 1] * plain1[..., 1]
    normal = y2 >= np.float32(1.2e-38)  # atomic float adds flush subnormal operands to zero
    assert np.array_equal(sq1[normal], y2[normal]) and (sq1[~normal] <= y2[~normal]).all()
    # never stopping, many waves
    f, sq, tw = renderer.run_adaptive(scene, renderer.new_film(scene), max_waves=24, batch_waves=8, rel_error=0.0)
    plain = renderer.run(scene, renderer.new_film(scene), first_wave=1, n_waves=24)
    assert (tw == 24).all() and np.array_equal(f[..., 3], plain[..., 3]) and np.allclose(f, plain, rtol=1e-5, atol=1e-7)
    n = f[..., 3]
    var = (sq - f[..., 1] ** 2 / n) / (n - 1)
    assert (var > -1e-6 * np.maximum(sq / n, 1e-30)).all(), "second moments must dominate the squared mean"
    # everything stops at the first eligible batch
    f, sq, tw = renderer.run_adaptive(scene, renderer.new_film(scene), max_waves=64, batch_waves=4, rel_error=1e9, min_waves=8)
    assert (tw == 8).all() and (f[..., 3] == 8).all()
    # a real threshold: per-pixel counts are the tile's wave count, tiles differ, and each tile is the truncated plain render
    f, sq, tw = renderer.run_adaptive(scene, renderer.new_film(scene), max_waves=96, batch_waves=8, rel_error=0.06, min_waves=8)
    th, twid = int(scene.tile_size[1]), int(scene.tile_size[0])
    per_pixel = np.repeat(np.repeat(tw.reshape(-1, ntx), th, axis=0), twid, axis=1)[: scene.height, : scene.width]
    assert np.array_equal(f[..., 3], per_pixel.astype(np.float32))
    assert tw.min() == 8 and tw.max() > tw.min() and (tw % 8 == 0).all(), "noise-free sky tiles stop at once, the flame keeps going"
    for w in np.unique(tw)[:3]:
        ref = renderer.run(scene, renderer.new_film(scene), first_wave=1, n_waves=int(w))
        m = per_pixel == w
        assert np.allclose(f[m], ref[m], rtol=1e-5, atol=1e-7)
    saved = 1.0 - tw.astype(np.float64).sum() / (96.0 * len(tw))
    print(f"\n[adaptive fire 70x100, rel_error 0.06] waves per tile min {tw.min()} / mean {tw.mean():.1f} / max {tw.max()}: {100 * saved:.0f} % of the 96-wave work skipped")


def test_adaptive_image_within_monte_carlo_noise(renderer, oracle, loaded):
    """The adaptive film's per-pixel mean against an independent CPU-oracle render: inside 3 sigma of the combined Monte-Carlo noise (sigma from
    the GPU's own second-moment plane), at the rate a plain GPU render achieves against the same CPU render."""
    scene, spec, grids = loaded("fire")
    f, sq, tw = renderer.run_adaptive(scene, renderer.new_film(scene), max_waves=256, batch_waves=16, rel_error=0.03, min_waves=16)
    cpu = np.zeros((scene.height, scene.width, 4), np.float32)
    n_c = 48
    oracle.volume(grids).render(scene, cpu, first_wave=1001, n_waves=n_c)  # independent waves
    plain = renderer.run(scene, renderer.new_film(scene), first_wave=1, n_waves=256)
    _, sq_p, _ = renderer.run_adaptive(scene, renderer.new_film(scene), max_waves=256, batch_waves=256, rel_error=0.0)

    def inside(film, sq2):
        n = film[..., 3].astype(np.float64)
        m = film[..., 1] / n
        var = np.maximum(sq2 / n - m * m, 0) * n / np.maximum(n - 1, 1)
        d = m - cpu[..., 1] / n_c
        tol = 3.0 * np.sqrt(var * (1.0 / n + 1.0 / n_c)) + 1e-7
        return float((np.abs(d) <= tol).mean())
    got, control = inside(f, sq), inside(plain, sq_p)
    print(f"\n[adaptive vs CPU oracle, fire 70x100] inside 3 sigma: adaptive {got:.4f}, plain 256-wave control {control:.4f}; mean waves/tile {tw.mean():.1f} of 256")
    assert got > 0.985 and got > control - 0.01


@pytest.mark.parametrize("name", ["fire", "wdas_cloud", "dense512"])
def test_device_bake_equals_host_bake(api, small_case, monkeypatch, name):
    """The brick stage of the baker on the device (csrc/device/vpt_bake.cu: apron bricks + fix_majorants_for_interpolation, volume.cpp:104-160)
    against the host baker (csrc/host/bake.cpp, VPT_HOST_BAKE=1): same bytes on the device, probed through every lookup the kernels make —
    values, getDim and patched majorants at points around every kind of boundary, trilinear samples through a rendered film."""
    scene, spec, grids = small_case(name)
    monkeypatch.delenv("VPT_HOST_BAKE", raising=False)
    dev = api.Renderer(0); dev.set_volume(grids)
    monkeypatch.setenv("VPT_HOST_BAKE", "1")
    host = api.Renderer(0); host.set_volume(grids)
    monkeypatch.delenv("VPT_HOST_BAKE", raising=False)
    try:
        assert dev.device_volume_bytes() == host.device_volume_bytes()
        rs = np.random.default_rng(3)
        d = np.asarray(spec.dims, np.float64)
        pts = np.concatenate([
            rs.random((20000, 3)) * (d + 24) - 12,                                   # anywhere, incl. outside the grid
            np.round(rs.random((20000, 3)) * (d / 8)) * 8 + rs.integers(-1, 2, (20000, 3)),  # on and next to leaf faces (where the apron and the halo patch matter)
        ]).astype(np.float32)
        for grid_id in ((0, 1) if grids.temperature is not None else (0,)):
            a, b = dev.sample_grid(grid_id, pts), host.sample_grid(grid_id, pts)
            for x, y in zip(a, b):
                assert np.array_equal(x, y, equal_nan=True)
        fa = dev.run(scene, dev.new_film(scene), 1, 1)
        fb = host.run(scene, host.new_film(scene), 1, 1)
        assert np.array_equal(fa, fb) and fa[..., :3].any()  # one wave: one sample per pixel, no add-order slack
    finally:
        dev.close(); host.close()


def test_tile_order_keeps_every_sample(api, small_case, monkeypatch):
    """Longest-tiles-first (DESIGN.md "launch drain"): once a scene's per-tile job times are known, full-frame ranges are walked tile-major in
    cost order by the ORD = 2 instantiation. Job ids are unchanged, so: one wave rendered ordered is BIT-equal to the same wave rendered in
    natural order (one sample per pixel: no add-order slack), multi-wave films agree to add order with exact counts, and a series that mixes
    ordered full-frame calls with tile-stripe calls (plain launches) still renders every job exactly once."""
    import torch
    scene, spec, grids = small_case("dense512")
    monkeypatch.setenv("VPT_TILE_ORDER", "0")
    plain = api.Renderer(0); plain.set_volume(grids)
    monkeypatch.delenv("VPT_TILE_ORDER")
    r = api.Renderer(0); r.set_volume(grids)
    try:
        H, W = scene.height, scene.width
        r.run(scene, r.new_film(scene), 1, 2)          # learning calls: natural order, per-tile times recorded
        a = r.run(scene, r.new_film(scene), 7, 1)      # ordered from here on
        b = plain.run(scene, plain.new_film(scene), 7, 1)
        assert np.array_equal(a, b) and (a[..., 3] == 1).all()
        a = r.run(scene, r.new_film(scene), 3, 6)
        b = plain.run(scene, plain.new_film(scene), 3, 6)
        assert np.array_equal(a[..., 3], b[..., 3]) and (a[..., 3] == 6).all() and np.allclose(a, b, rtol=1e-5, atol=1e-7)
        # asynchronous series on a device film: ordered full frames, then a stripe of tiles (a plain launch), then full frames again
        d = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda:0")
        ntx = -(-W // int(scene.tile_size[0])); nty = -(-H // int(scene.tile_size[1]))
        r.run_device_range(scene, d.data_ptr(), 1, 2, wait=False)
        r.run_device_range(scene, d.data_ptr(), 3, 1, tile_begin=0, tile_count=ntx * (nty // 2), wait=False)
        r.run_device_range(scene, d.data_ptr(), 3, 1, tile_begin=ntx * (nty // 2), tile_count=ntx * (nty - nty // 2), wait=False)
        r.run_device_range(scene, d.data_ptr(), 4, 3, wait=False)
        r.wait()
        torch.cuda.synchronize()
        got = d.cpu().numpy()
        want = plain.run(scene, plain.new_film(scene), 1, 6)
        assert np.array_equal(got[..., 3], want[..., 3]) and (got[..., 3] == 6).all() and np.allclose(got, want, rtol=1e-5, atol=1e-7)
        # stripes of a scene whose order is known are ordered too (their own tiles out of the sorted list): interleaved rows, then contiguous halves
        for stride in (2, 1):
            d.zero_()
            for g in range(2):
                rows = len(range(g, nty, 2)) if stride == 2 else (nty // 2 if g == 0 else nty - nty // 2)
                begin = g * ntx if stride == 2 else (0 if g == 0 else ntx * (nty // 2))
                r.run_device_range(scene, d.data_ptr(), 2, 3, tile_begin=begin, tile_count=ntx * rows, tile_row_stride=stride)
            torch.cuda.synchronize()
            got = d.cpu().numpy()
            want = plain.run(scene, plain.new_film(scene), 2, 3)
            assert np.array_equal(got[..., 3], want[..., 3]) and (got[..., 3] == 3).all() and np.allclose(got, want, rtol=1e-5, atol=1e-7)
        one = torch.zeros_like(d)
        r.run_device_range(scene, one.data_ptr(), 9, 1, tile_begin=ntx, tile_count=ntx * len(range(1, nty, 2)), tile_row_stride=2)
        torch.cuda.synchronize()
        ref = plain.run(scene, plain.new_film(scene), 9, 1)
        got = one.cpu().numpy()
        mask = got[..., 3] > 0
        assert mask.any() and not mask.all() and np.array_equal(got[mask], ref[mask])  # one wave of a stripe: bit-equal where it renders
    finally:
        r.close(); plain.close()
