"""Several GPUs behind the C ABI (vptc_group_*, include/vptc.h) and behind the `vpt` CLI (VPT_DEVICES): the film of a sharded render
equals the single-GPU film — sample counts exactly; radiance bit for bit where no add order is involved (one wave per device, or disjoint
tile stripes of one wave), to 1e-5 otherwise. Needs >= 2 visible GPUs for the real thing (`gpurun --gpus 2`); the one-device group runs anywhere."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def test_group_of_one_device_equals_the_context(api, renderer, small_case):
    scene, spec, grids = small_case("fire")
    renderer.set_volume(grids)
    want = renderer.run(scene, renderer.new_film(scene), first_wave=1, n_waves=1)
    g = api.Group([0])
    try:
        g.set_volume(grids)
        for shard in (api.SHARD_WAVES, api.SHARD_TILES):
            got = g.run(scene, np.zeros_like(want), first_wave=1, n_waves=1, sharding=shard)
            assert np.array_equal(got, want)
        with pytest.raises(api.VptError):
            api.Group([0, 0])
    finally:
        g.close()


@pytest.mark.parametrize("name", ["fire", "dense512"])
def test_group_film_equals_single_gpu_film(api, renderer, small_case, name):
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 visible GPUs")
    scene, spec, grids = small_case(name)
    renderer.set_volume(grids)
    devices = list(range(min(n, 8)))
    G = len(devices)
    g = api.Group(devices)
    try:
        g.set_volume(grids)
        assert g.size() == G
        # waves: G waves = one per device -> each partial film holds ONE sample per pixel; the sum over devices is in device order
        single = [renderer.run(scene, renderer.new_film(scene), first_wave=w, n_waves=1) for w in range(1, G + 1)]
        want = single[0].copy()
        for f in single[1:]:
            want += f
        got = g.run(scene, np.zeros_like(want), first_wave=1, n_waves=G, sharding=api.SHARD_WAVES)
        assert np.array_equal(got[..., 3], want[..., 3]) and (got[..., 3] == G).all()
        assert np.array_equal(got, want), "one wave per device: the reduce adds the partial films in device order"
        # tile stripes, one wave: disjoint pixels, no arithmetic
        got = g.run(scene, np.zeros_like(want), first_wave=3, n_waves=1, sharding=api.SHARD_TILES)
        assert np.array_equal(got, renderer.run(scene, renderer.new_film(scene), first_wave=3, n_waves=1))
        # many waves, both shardings, asynchronous series collected once
        ref = renderer.run(scene, renderer.new_film(scene), first_wave=1, n_waves=20)
        for shard in (api.SHARD_WAVES, api.SHARD_TILES):
            acc = np.zeros_like(ref)
            for first in range(1, 21, 5):
                g.run_async(scene, first, 5, shard)
            g.wait(acc)
            assert np.array_equal(acc[..., 3], ref[..., 3]) and np.allclose(acc, ref, rtol=1e-5, atol=1e-7)
        # a ragged frame whose tile rows do not divide by the device count
        odd = scene.copy(); odd.output_size[0], odd.output_size[1] = 61, 29
        want = renderer.run(odd, renderer.new_film(odd), first_wave=2, n_waves=1)
        assert np.array_equal(g.run(odd, np.zeros_like(want), first_wave=2, n_waves=1, sharding=api.SHARD_TILES), want)
    finally:
        g.close()


def test_cli_on_several_devices(api, scenes, tmp_path):
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 visible GPUs")
    from PIL import Image
    vpt = ROOT / "volume_path_tracer_b200" / "vpt"
    grids = api.synth_volume("fire", (50, 60, 50), 1.0, 1, None)
    (tmp_path / "volumes").mkdir()
    grids.write(tmp_path / "volumes" / "fire.nvdb")
    d = json.loads(scenes.config_json("fire"))
    d["output_size"] = [96, 72]; d["num_waves"] = 2; d["volume_path"] = "../volumes/fire.nvdb"
    (tmp_path / "scenes").mkdir()
    p = tmp_path / "scenes" / "fire.json"
    p.write_text(json.dumps(d))
    outs = {}
    for label, env in (("one", {}), ("waves", {"VPT_DEVICES": "0,1"}), ("tiles", {"VPT_DEVICES": "0,1", "VPT_SHARD": "tiles"})):
        out = tmp_path / f"{label}.png"
        r = subprocess.run([str(vpt), str(p), str(out)], capture_output=True, text=True, timeout=300, env={**os.environ, **env})
        assert r.returncode == 0, r.stderr
        if env:
            assert "Rendering on 2 devices" in r.stderr
        outs[label] = np.asarray(Image.open(out))
    assert np.array_equal(outs["one"], outs["waves"])  # 2 waves: w1 + w2 commutes
    assert np.array_equal(outs["one"], outs["tiles"])  # two samples per pixel: a + b commutes, whichever lane adds first
