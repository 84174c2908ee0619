"""The headless `vpt scene.json out.png` CLI (csrc/vpt_main.cpp): the reference's command line, JSON schema, volume_path
relative to the scene file, density-required / temperature-optional grids, [FATAL] + exit(1) on errors (main.cpp:26-56,
volume_grids.cpp:36-67, logging.hpp:16) — and the image it writes equals film_to_image of the CPU oracle's film."""
import json
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
    assert diff.max() <= 1 and (diff == 0).mean() > 0.995  # fp32 add order / powf ulp at an 8-bit threshold
    assert img.max() > 30


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
