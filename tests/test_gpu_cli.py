"""The headless `vpt scene.json out.png` CLI (csrc/vpt_main.cpp): the reference's command line, JSON schema, volume_path
relative to the scene file, density-required / temperature-optional grids, [FATAL] + exit(1) on errors (main.cpp:26-56,
volume_grids.cpp:36-67, logging.hpp:16) — and the image it writes equals film_to_image of the CPU oracle's film."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

VPT = ROOT / "volume_path_tracer_b200" / "vpt"


def _scene(scenes, tmp_path, volume_rel, size=(80, 60), waves=6):
    d = json.loads(scenes.config_json("fire"))
    d["output_size"] = list(size)
    d["num_waves"] = waves
    d["volume_path"] = volume_rel
    (tmp_path / "scenes").mkdir(exist_ok=True)
    p = tmp_path / "scenes" / "fire.json"
    p.write_text(json.dumps(d))
    return p


@pytest.mark.gpu
def test_cli_renders_the_oracle_image(api, scenes, oracle, tmp_path):
    from PIL import Image
    grids = api.synth_volume("fire", (50, 60, 50), 1.0, 1, None)
    (tmp_path / "volumes").mkdir()
    grids.write(tmp_path / "volumes" / "fire.nvdb")
    scene_path = _scene(scenes, tmp_path, "../volumes/fire.nvdb")
    out = tmp_path / "out.png"
    r = subprocess.run([str(VPT), str(scene_path), str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Rendering complete in" in r.stderr
    img = np.asarray(Image.open(out))
    scene, _ = api.read_configuration(scene_path)
    assert img.shape == (scene.height, scene.width, 3)
    film = np.zeros((scene.height, scene.width, 4), np.float32)
    oracle.volume(grids).render(scene, film)
    want = oracle.film_to_image(film)
    diff = np.abs(img.astype(int) - want.astype(int))
    assert diff.max() <= 1 and (diff == 0).mean() > 0.995  # several waves: fp32 add order of the film sums at an 8-bit threshold (the resolve itself is byte-exact)
    assert img.max() > 30


@pytest.mark.gpu
def test_cli_one_wave_png_is_byte_identical(api, scenes, oracle, tmp_path):
    """One wave = one sample per pixel: no add-order slack, so the PNG the CLI writes must equal film_to_image of the oracle's film
    in every byte (integrator bit-exact + resolve byte-exact)."""
    from PIL import Image
    grids = api.synth_volume("fire", (50, 60, 50), 1.0, 1, None)
    (tmp_path / "volumes").mkdir()
    grids.write(tmp_path / "volumes" / "fire.nvdb")
    scene_path = _scene(scenes, tmp_path, "../volumes/fire.nvdb", size=(97, 61), waves=1)
    out = tmp_path / "one.png"
    r = subprocess.run([str(VPT), str(scene_path), str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    scene, _ = api.read_configuration(scene_path)
    film = np.zeros((scene.height, scene.width, 4), np.float32)
    oracle.volume(grids).render(scene, film)
    assert np.array_equal(np.asarray(Image.open(out)), oracle.film_to_image(film))


def test_cli_errors_are_fatal(api, scenes, tmp_path):
    """No GPU needed: argument, configuration and grid errors exit 1 with the reference's [FATAL] convention."""
    r = subprocess.run([str(VPT)], capture_output=True, text=True)
    assert r.returncode == 1 and "[FATAL]: Usage:" in r.stderr
    r = subprocess.run([str(VPT), str(tmp_path / "missing.json"), "o.png"], capture_output=True, text=True)
    assert r.returncode == 1 and "[FATAL]" in r.stderr
    bad = tmp_path / "bad.json"
    d = json.loads(scenes.config_json("fire")); del d["seed"]
    bad.write_text(json.dumps(d))
    r = subprocess.run([str(VPT), str(bad), "o.png"], capture_output=True, text=True)
    assert r.returncode == 1 and "seed" in r.stderr
    # a volume file without the "density" grid (volume_grids.cpp:48-56)
    grids = api.synth_volume("fire", (24, 24, 24), 1.0, 1, None)
    (tmp_path / "volumes").mkdir()
    api.VolumeGrids(grids.temperature, None).write(tmp_path / "volumes" / "fire.nvdb")
    scene_path = _scene(scenes, tmp_path, "../volumes/fire.nvdb")
    r = subprocess.run([str(VPT), str(scene_path), "o.png"], capture_output=True, text=True)
    assert r.returncode == 1 and 'does not contain the "density" grid' in r.stderr


@pytest.mark.gpu
def test_cli_progressive_save_and_resume(api, scenes, tmp_path):
    """VPT_SAVE_EVERY rewrites the image while rendering; VPT_CHECKPOINT lets a second run with a larger
    num_waves continue where the first stopped, and the result is the uninterrupted render bit for bit (one wave per
    call: exactly one sample per pixel per call, so the film sums are order-independent)."""
    from PIL import Image
    grids = api.synth_volume("fire", (50, 60, 50), 1.0, 1, None)
    (tmp_path / "volumes").mkdir()
    grids.write(tmp_path / "volumes" / "fire.nvdb")
    env = {**os.environ, "VPT_WAVES_PER_CALL": "1"}
    full = tmp_path / "full.png"
    p6 = _scene(scenes, tmp_path, "../volumes/fire.nvdb", waves=6)
    r = subprocess.run([str(VPT), str(p6), str(full)], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stderr
    # first leg: 4 waves, image rewritten every 2 waves, checkpoint kept
    ck = tmp_path / "render.film"
    out = tmp_path / "out.png"
    p4 = _scene(scenes, tmp_path, "../volumes/fire.nvdb", waves=4)
    r = subprocess.run([str(VPT), str(p4), str(out)], capture_output=True, text=True, timeout=300, env={**env, "VPT_SAVE_EVERY": "2", "VPT_CHECKPOINT": str(ck)})
    assert r.returncode == 0, r.stderr
    scene4, _ = api.read_configuration(p4)
    film, nxt = api.film_load(ck, scene4)
    assert nxt == 5 and (film[..., 3] == 4).all()
    # second leg: the scene now asks for 6 waves; the run resumes at wave 5
    p6 = _scene(scenes, tmp_path, "../volumes/fire.nvdb", waves=6)
    r = subprocess.run([str(VPT), str(p6), str(out)], capture_output=True, text=True, timeout=300, env={**env, "VPT_CHECKPOINT": str(ck)})
    assert r.returncode == 0, r.stderr
    assert "Resuming from checkpoint" in r.stderr and "Starting wave 5/6" in r.stderr and "Starting wave 4/6" not in r.stderr
    scene6, _ = api.read_configuration(p6)
    film, nxt = api.film_load(ck, scene6)
    assert nxt == 7 and (film[..., 3] == 6).all()
    assert np.array_equal(np.asarray(Image.open(out)), np.asarray(Image.open(full)))
    # a checkpoint of another scene is refused the reference's way
    d = json.loads(p6.read_text()); d["seed"] += 1; p6.write_text(json.dumps(d))
    r = subprocess.run([str(VPT), str(p6), str(out)], capture_output=True, text=True, timeout=300, env={**env, "VPT_CHECKPOINT": str(ck)})
    assert r.returncode == 1 and "[FATAL]" in r.stderr and "different scene" in r.stderr


@pytest.mark.gpu
def test_cli_sigint_stops_at_a_wave_boundary_and_saves(api, scenes, tmp_path):
    """Closing the reference's window stops at the next wave and still saves the image (tile_provider.cpp:107-110,
    main.cpp:133-143); headless, SIGINT does the same and leaves a checkpoint to resume from."""
    import signal, time
    grids = api.synth_volume("fire", (50, 60, 50), 1.0, 1, None)
    (tmp_path / "volumes").mkdir()
    grids.write(tmp_path / "volumes" / "fire.nvdb")
    p = _scene(scenes, tmp_path, "../volumes/fire.nvdb", size=(320, 240), waves=100000)
    ck = tmp_path / "render.film"
    out = tmp_path / "out.png"
    proc = subprocess.Popen([str(VPT), str(p), str(out)], stderr=subprocess.PIPE, text=True,
                            env={**os.environ, "VPT_WAVES_PER_CALL": "4", "VPT_CHECKPOINT": str(ck)})
    seen = ""
    t0 = time.time()
    while "Starting wave 9/" not in seen and time.time() - t0 < 120:  # a few calls in
        seen += proc.stderr.readline()
    proc.send_signal(signal.SIGINT)
    rest = proc.communicate(timeout=120)[1]
    assert proc.returncode == 0, seen + rest
    assert "Stopped at wave" in rest and out.exists()
    scene, _ = api.read_configuration(p)
    film, nxt = api.film_load(ck, scene)
    assert nxt > 9 and (nxt - 1) % 4 == 0 and (film[..., 3] == nxt - 1).all()


@pytest.mark.gpu
def test_cli_trace_pixel_writes_the_majorant_trace_csv(api, scenes, oracle, tmp_path):
    """VPT_TRACE_PIXEL=x,y = Volume::log_majorant_trace (volume.cpp:176-192) for the pixel-centre camera ray: the
    reference's columns, T0/T1/Majorant equal to the oracle's segment iterator (times scaled to world units), end
    points = world_to_density_index(ray.eval(t)), segments contiguous; a pixel that misses the volume gives the header only."""
    grids = api.synth_volume("fire", (50, 60, 50), 1.0, 1, None)
    (tmp_path / "volumes").mkdir()
    grids.write(tmp_path / "volumes" / "fire.nvdb")
    scene_path = _scene(scenes, tmp_path, "../volumes/fire.nvdb", waves=1)
    scene, _ = api.read_configuration(scene_path)
    px, py = scene.width // 2, scene.height // 2
    env = dict(os.environ, VPT_TRACE_PIXEL=f"{px},{py}")
    r = subprocess.run([str(VPT), str(scene_path), str(tmp_path / "o.png")], capture_output=True, text=True, timeout=300, env=env, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    lines = (tmp_path / "majorant_trace.csv").read_text().strip().split("\n")
    assert lines[0] == "X0,Y0,Z0,X1,Y1,Z1,T0,T1,Majorant"
    rows = np.array([[float(v) for v in ln.split(",")] for ln in lines[1:]], np.float64)
    o, d = oracle.camera_rays(scene, [px], [py], [0.0], [0.0])
    vol = oracle.volume(grids)
    segs, n = vol.trace_segments(o, d, 256)
    assert n[0] >= 2 and len(rows) == n[0]
    # idx_to_world_scale of this (uniform-voxel) grid = voxel size; compare world-space times and majorants
    ratio = rows[:, 6] / segs[0, : n[0], 0].astype(np.float64)
    assert np.allclose(ratio, ratio[0], rtol=1e-6) and np.allclose(rows[:, 7], segs[0, : n[0], 1].astype(np.float64) * ratio[0], rtol=1e-6)
    assert np.array_equal(rows[:, 8].astype(np.float32), segs[0, : n[0], 2])
    assert np.allclose(rows[1:, 0:3], rows[:-1, 3:6], atol=1e-3)  # each segment starts where the previous one ended
    # the end points lie on the index-space ray: trilinear probes there agree with the oracle's at the same points
    mid = (rows[:, 0:3] + rows[:, 3:6]) / 2
    lo, hi = np.array([0, 0, 0]) - 1.0, np.array([50, 60, 50]) + 1.0
    assert ((mid > lo - 64) & (mid < hi + 64)).all()
    # log.csv: the path event log of wave 1 through that pixel alone (Logger<true> + single_pixel), equal to the oracle's
    want = vol.path_log(scene, (py // 8) * (-(-scene.width // 8)) + px // 8, px, py)
    names = ["new_ray", "sampled_point", "null", "scatter", "absorbed", "scatter_terminated"]
    got = (tmp_path / "log.csv").read_text().strip().split("\n")
    assert len(got) == len(want) and [g.split(",")[0] for g in got] == [names[int(e[0])] for e in want]
    first = [np.float32(v) for v in got[0].split(",")[1:]]
    assert np.array_equal(np.array(first, np.float32), want[0, 1:7])
    env["VPT_TRACE_PIXEL"] = "0,0"  # a corner pixel misses this small volume
    r = subprocess.run([str(VPT), str(scene_path), str(tmp_path / "o.png")], capture_output=True, text=True, timeout=300, env=env, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    o, d = oracle.camera_rays(scene, [0], [0], [0.0], [0.0])
    assert vol.trace_segments(o, d, 256)[1][0] == 0
    assert (tmp_path / "majorant_trace.csv").read_text().strip() == "X0,Y0,Z0,X1,Y1,Z1,T0,T1,Majorant"
    env["VPT_TRACE_PIXEL"] = "nonsense"
    r = subprocess.run([str(VPT), str(scene_path), str(tmp_path / "o.png")], capture_output=True, text=True, timeout=300, env=env, cwd=tmp_path)
    assert r.returncode == 1 and "[FATAL]" in r.stderr
