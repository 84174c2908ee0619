"""Host-side logic and the C-ABI surface — no GPU needed.

* libvptcuda.so loads and exports every symbol include/vptc.h declares; struct layouts match the header.
* JSON scene reader = read_configuration (src/configuration.cpp:8-22): all keys mandatory; the reference's own
  shipped scene files parse to the values volume_path_tracer_b200/scenes.py carries.
* `.nvdb` writer/reader round trip; density required / temperature optional (src/volume_grids.cpp:36-67).
* NanoVDB-format builder: the serialized grids are accepted by the reference's vendored NanoVDB (via oracle/_ref).
* PNG writer (src/image_io.cpp:18-58) round trip through PIL; film_to_image oracle known answers.
* No GPU => vptc_create fails loudly (no CPU fallback).
"""
import ctypes as C
import os
import json
import re
import sys
from pathlib import Path

import numpy as np
import pytest

from conftest import ROOT


def test_abi_exports_every_declared_symbol(api):
    header = (ROOT / "include" / "vptc.h").read_text()
    declared = re.findall(r"VPTC_API\s+[\w\s\*]+?\b(vptc_\w+)\s*\(", header)
    assert len(declared) >= 20 and len(set(declared)) == len(declared)
    L = api.lib()
    for name in declared:
        assert hasattr(L, name), f"libvptcuda.so does not export {name}"
    assert sorted(declared) == sorted(api.EXPORTS), "api.EXPORTS out of date with include/vptc.h"
    assert b"sm_100a" in L.vptc_version()


def test_host_library_has_no_cuda_dependency(api):
    """libvpthost.so = the host-side helpers alone: exports them all, and maps no CUDA library (CPU-only tools load this one)."""
    import subprocess
    H = api.host_lib()
    assert [s for s in api.HOST_EXPORTS if not hasattr(H, s)] == []
    deps = subprocess.run(["ldd", str(api.HOST_LIB_PATH)], capture_output=True, text=True).stdout
    assert "cuda" not in deps.lower() and "nccl" not in deps.lower()
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from volume_path_tracer_b200 import api, scenes\n"
            "s, spec = scenes.small_variant('fire', (16, 16), 1, 0.1)\n"
            "g = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)\n"
            "maps = open('/proc/self/maps').read()\n"
            "assert 'libvpthost.so' in maps and 'libvptcuda.so' not in maps, maps\n"
            "print('ok')\n") % str(ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-1500:]


def test_abi_struct_layout(api):
    # natural alignment of the C structs in include/vptc.h
    assert C.sizeof(api.CameraParameters) == 44
    assert C.sizeof(api.InfiniteLight) == 16 and C.sizeof(api.DistantLight) == 28
    assert C.sizeof(api.VolumeParameters) == 24
    assert C.sizeof(api.WorkerParameters) == 80
    assert C.sizeof(api.Scene) == 200
    assert C.sizeof(api.Counters) == 56


def test_abi_struct_layout_matches_c_compiler(api, tmp_path):
    """sizeof/offsetof as gcc sees include/vptc.h == the ctypes mirror."""
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "vptc.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(vptc_scene), offsetof(vptc_scene, camera_parameters), offsetof(vptc_scene, worker_parameters),'
                   'offsetof(vptc_scene, volume_parameters), offsetof(vptc_worker_parameters, distant_light),'
                   'offsetof(vptc_worker_parameters, max_depth), sizeof(vptc_counters));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", str(ROOT / "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    S, Wp = api.Scene, api.WorkerParameters
    assert got == [C.sizeof(S), S.camera_parameters.offset, S.worker_parameters.offset, S.volume_parameters.offset,
                   Wp.distant_light.offset, Wp.max_depth.offset, C.sizeof(api.Counters)]


def test_library_has_sm100a_code_only():
    import subprocess
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "--list-elf", str(ROOT / "volume_path_tracer_b200" / "libvptcuda.so")],
                         capture_output=True, text=True)
    if out.returncode != 0:
        pytest.skip("cuobjdump unavailable")
    archs = set(re.findall(r"sm_\d+a?", out.stdout))
    assert archs == {"sm_100a"}, archs


def test_no_gpu_fails_loudly(api):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.VptError) as e:
        api.Renderer(0)
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)


def test_null_arguments_are_errors_not_crashes(api):
    L = api.lib()
    assert L.vptc_set_volume(None, None, 0, None, 0) == -1
    assert L.vptc_render(None, None, 1, 1, None) == -1
    assert L.vptc_scene_from_json(None, None, None, 0) == -1
    assert L.vptc_nvdb_read_grid(b"/nonexistent/file.nvdb", b"density", C.byref(C.c_void_p()), C.byref(C.c_size_t())) == -5
    L.vptc_destroy(None)
    L.vptc_free(None)


# ---------------------------------------------------------------- scene JSON
def test_scene_json_roundtrip(api, scenes):
    s, vol = api.scene_from_json(scenes.config_json("fire"))
    assert vol == "synthetic"
    assert (s.seed, s.width, s.height, s.num_waves, s.num_workers) == (500, 700, 1000, 1024, 12)
    assert tuple(s.tile_size) == (8, 8) and s.ntiles == 11000
    assert tuple(s.camera_parameters.position) == (120.0, 30.0, 0.0) and s.camera_parameters.vfov_deg == 37.0
    assert s.camera_parameters.imaging_ratio == np.float32(0.1)
    wp, vp = s.worker_parameters, s.volume_parameters
    assert wp.max_depth == 1000000 and wp.use_jitter == 1 and wp.single_pixel_enabled == 0 and tuple(wp.single_pixel_coord) == (239, 879)
    assert tuple(wp.distant_light.inv_direction) == (0.5, 1.0, 0.0) and wp.distant_light.multiplier == 20.0
    assert (vp.sigma_s, vp.sigma_a, vp.henyey_greenstein_g) == (np.float32(0.9), 2.0, np.float32(0.7))
    assert vp.le_scale == np.float32(4e-8) and vp.temperature_offset == 300.0 and vp.temperature_scale == 43.0


def _drop(d, path):
    d = json.loads(json.dumps(d))
    node = d
    for k in path[:-1]:
        node = node[k]
    del node[path[-1]]
    return d


def test_scene_json_every_key_is_mandatory(api, scenes):
    """glaze error_on_missing_keys = true (src/configuration.cpp:13-15)."""
    base = json.loads(scenes.config_json("wdas_cloud"))

    def paths(d, prefix=()):
        for k, v in d.items():
            yield prefix + (k,)
            if isinstance(v, dict):
                yield from paths(v, prefix + (k,))
    all_paths = list(paths(base))
    assert len(all_paths) == 32
    for p in all_paths:
        with pytest.raises(api.VptError) as e:
            api.scene_from_json(json.dumps(_drop(base, p)))
        assert e.value.code == -6 and p[-1] in str(e.value)


def test_scene_sizes_are_bounded_before_they_multiply(api, scenes):
    """A [2^31, 2^31] frame used to wrap W*H*16 to 0 (round-1 advisor finding); integers beyond 2^53 / non-finite are not integers."""
    d = json.loads(scenes.config_json("fire"))
    for bad in ("1e300", "1e19", "9007199254740993", "-1e300"):
        txt = json.dumps(d).replace('"output_size": [700, 1000]', f'"output_size": [{bad}, 8]')
        assert bad in txt
        with pytest.raises(api.VptError):
            api.scene_from_json(txt)
    s, _ = api.scene_from_json(json.dumps(d))
    film = np.zeros((4, 4, 4), np.float32)
    rgb = np.zeros((4, 4, 3), np.uint8)
    L = api.lib()
    # no context needed to see the argument check: a null ctx is refused first, so go through film_to_image's own size check with a dummy ctx-less call
    assert L.vptc_film_to_image(None, film.ctypes.data, 4, 4, rgb.ctypes.data) == -1


def test_scene_json_rejects_bad_input(api, scenes):
    base = json.loads(scenes.config_json("fire"))
    for mutate in (lambda d: d.__setitem__("output_size", [700]), lambda d: d.__setitem__("seed", "x"),
                   lambda d: d["camera_parameters"].__setitem__("up", [0, 1]), lambda d: d.__setitem__("unknown_key", 1)):
        d = json.loads(json.dumps(base)); mutate(d)
        with pytest.raises(api.VptError):
            api.scene_from_json(json.dumps(d))
    with pytest.raises(api.VptError):
        api.scene_from_json("{ not json")
    with pytest.raises(api.VptError):
        api.read_configuration("/nonexistent/scene.json")


@pytest.mark.parametrize("name", ["fire", "fire_lowscattering", "wdas_cloud"])
def test_reference_scene_files_parse_to_our_configs(api, scenes, reference_tree, name):
    """The reference's shipped scenes go through our reader unchanged and equal scenes.py's C1-C3 (Appendix B)."""
    ref_scene, vol = api.read_configuration(reference_tree / "scenes" / f"{name}.json")
    ours, _ = scenes.load_config(name)
    assert bytes(ref_scene) == bytes(ours)
    assert vol.startswith("../volumes/")


# ---------------------------------------------------------------- .nvdb I/O
def test_nvdb_roundtrip_and_grid_contract(api, tmp_path):
    grids = api.synth_volume("fire", (40, 48, 40), 0.5, 2, None)
    assert grids.has_temperature()
    p = tmp_path / "fire.nvdb"
    grids.write(p)
    back = api.VolumeGrids.read_from_file(p)
    assert back.has_temperature()
    assert bytes(back.density.view()) == bytes(grids.density.view())
    assert bytes(back.temperature.view()) == bytes(grids.temperature.view())
    # density only: temperature optional
    q = tmp_path / "cloud.nvdb"
    cloud = api.synth_volume("cloud", (64, 48, 64), 1.0, 2, None)
    assert not cloud.has_temperature()
    cloud.write(q)
    assert not api.VolumeGrids.read_from_file(q).has_temperature()
    # a file without "density" is fatal (volume_grids.cpp:48-56)
    only_t = api.VolumeGrids(grids.temperature, None)
    r = tmp_path / "t.nvdb"
    only_t.write(r)
    with pytest.raises(api.VptError) as e:
        api.VolumeGrids.read_from_file(r)
    assert e.value.code == -2 and "density" in str(e.value)
    grids.temperature = grids.temperature  # keep alive
    (tmp_path / "junk.nvdb").write_bytes(b"not a nanovdb file at all" * 10)
    with pytest.raises(api.VptError):
        api.VolumeGrids.read_from_file(tmp_path / "junk.nvdb")


def test_nvdb_container_variants(api, tmp_path):
    """The loader takes what nanovdb::io::readGrid takes (NV/io/IO.h:204-320,575-618): ZIP-coded segments (u64 size + zlib stream;
    the author's real files may be compressed), several grids in ONE segment, and raw dumps of concatenated grid buffers
    (NV/NanoVDB.h:5861-5865); BLOSC and truncated files are errors, not crashes."""
    import struct
    import zlib
    grids = api.synth_volume("fire", (40, 48, 40), 0.5, 2, None)
    p = tmp_path / "fire.nvdb"
    grids.write(p)
    raw = p.read_bytes()
    # parse our own file: one segment per grid = FileHeader(16) + FileMetaData(176) + name + grid bytes
    segs, off = [], 0
    while off < len(raw):
        magic, version, count, codec = struct.unpack_from("<QIHH", raw, off)
        assert count == 1 and codec == 0
        meta = bytearray(raw[off + 16: off + 16 + 176])
        grid_size, file_size = struct.unpack_from("<QQ", meta, 0)
        name_size = struct.unpack_from("<I", meta, 8 * 4 + 8 + 48 + 24 + 24)[0]
        name = raw[off + 192: off + 192 + name_size]
        data = raw[off + 192 + name_size: off + 192 + name_size + grid_size]
        assert file_size == grid_size and len(data) == grid_size
        segs.append((magic, version, meta, name, data))
        off += 192 + name_size + grid_size
    assert [s[3].rstrip(b"\0") for s in segs] == [b"density", b"temperature"]

    def check(path):
        back = api.VolumeGrids.read_from_file(path)
        assert bytes(back.density.view()) == bytes(grids.density.view()) and bytes(back.temperature.view()) == bytes(grids.temperature.view())

    # 1. ZIP codec, one segment per grid
    out = b""
    for magic, version, meta, name, data in segs:
        z = zlib.compress(data, 6)
        m = bytearray(meta); struct.pack_into("<Q", m, 8, len(z) + 8)  # fileSize = bytes on disk
        out += struct.pack("<QIHH", magic, version, 1, 1) + bytes(m) + name + struct.pack("<Q", len(z)) + z
    (tmp_path / "zip.nvdb").write_bytes(out)
    assert len(out) < len(raw)
    check(tmp_path / "zip.nvdb")
    # 2. both grids in ONE uncompressed segment: header, all metadata + names, then all buffers
    magic, version = segs[0][0], segs[0][1]
    out = struct.pack("<QIHH", magic, version, 2, 0) + b"".join(bytes(m) + n for _, _, m, n, _ in segs) + b"".join(d for *_, d in segs)
    (tmp_path / "one_segment.nvdb").write_bytes(out)
    check(tmp_path / "one_segment.nvdb")
    # 3. raw dump of the grid buffers
    (tmp_path / "raw.nvdb").write_bytes(segs[1][4] + segs[0][4])  # temperature first: lookup is by name
    check(tmp_path / "raw.nvdb")
    # 4. BLOSC (codec 2) is refused with a message; truncation is an error
    magic, version, meta, name, data = segs[0]
    (tmp_path / "blosc.nvdb").write_bytes(struct.pack("<QIHH", magic, version, 1, 2) + bytes(meta) + name + data)
    with pytest.raises(api.VptError) as e:
        api.VolumeGrids.read_from_file(tmp_path / "blosc.nvdb")
    assert "BLOSC" in str(e.value)
    (tmp_path / "cut.nvdb").write_bytes(raw[: len(raw) // 3])
    with pytest.raises(api.VptError):
        api.VolumeGrids.read_from_file(tmp_path / "cut.nvdb")
    (tmp_path / "zipcut.nvdb").write_bytes((tmp_path / "zip.nvdb").read_bytes()[:4000])
    with pytest.raises(api.VptError):
        api.VolumeGrids.read_from_file(tmp_path / "zipcut.nvdb")


def test_grid_header_fields(api):
    g = api.synth_volume("dense", (32, 32, 32), 0.5, 1, (1.0, 2.0, 3.0))
    b = g.density.view()
    assert bytes(b[:8]) == b"NanoVDB1"
    name = bytes(b[40:40 + 8]).split(b"\0")[0]
    assert name == b"density"
    vox = np.frombuffer(bytes(b[608:632]), np.float64)
    assert np.allclose(vox, 0.5)
    assert int(np.frombuffer(bytes(b[32:40]), np.uint64)[0]) == g.density.nbytes


def test_synth_is_deterministic(api):
    a = api.synth_volume("cloud", (96, 64, 96), 1.0, 5, None)
    b = api.synth_volume("cloud", (96, 64, 96), 1.0, 5, None)
    c = api.synth_volume("cloud", (96, 64, 96), 1.0, 6, None)
    assert bytes(a.density.view()) == bytes(b.density.view())
    assert bytes(a.density.view()) != bytes(c.density.view())
    with pytest.raises(api.VptError):
        api.synth_volume("nonsense", (8, 8, 8), 1.0, 1, None)


def test_reference_nanovdb_accepts_our_files(api, oracle, oracle_ref, tmp_path):
    """A grid written by our builder + .nvdb writer, read back by our reader, is walked identically by the
    reference's vendored NanoVDB accessor (oracle/_ref) and by our own tree walker."""
    grids = api.synth_volume("torus", (120, 60, 120), 1.0, 1, None)
    p = tmp_path / "torus.nvdb"
    grids.write(p)
    back = api.VolumeGrids.read_from_file(p)
    a, b = oracle.volume(back), oracle_ref.volume(back)
    pts = (np.random.default_rng(0).random((30000, 3)).astype(np.float32) * 2 - 1) * np.array([120, 60, 120], np.float32)
    va, vb = a.sample_grid(0, pts), b.sample_grid(0, pts)
    for x, y in zip(va, vb):
        assert np.array_equal(x, y)
    assert (va[0] > 0).mean() > 0.01
    assert np.array_equal(a.leaf_maxima(), b.leaf_maxima())


# ---------------------------------------------------------------- PNG + film_to_image
def test_png_roundtrip(api, tmp_path):
    from PIL import Image
    rgb = np.random.default_rng(0).integers(0, 256, (37, 53, 3), dtype=np.uint8)
    api.write_png(tmp_path / "a.png", rgb)
    img = Image.open(tmp_path / "a.png")
    assert img.mode == "RGB" and img.size == (53, 37)
    assert np.array_equal(np.asarray(img), rgb)


def test_film_to_image_known_answers(oracle):
    """src/main.cpp:12-24 + include/vpt/color.hpp:8-30: XYZ/N -> linear sRGB -> OETF -> clamp -> *255 -> truncate."""
    film = np.zeros((1, 4, 4), np.float32)
    film[0, 0] = (0.95047 * 3, 1.0 * 3, 1.08883 * 3, 3)   # D65 white, 3 samples -> (255,255,255) up to truncation
    film[0, 1] = (0, 0, 0, 5)                               # black
    film[0, 2] = (0.4124 * 2, 0.2126 * 2, 0.0193 * 2, 2)    # sRGB red primary
    film[0, 3] = (100, 100, 100, 1)                         # clamps
    rgb = oracle.film_to_image(film)
    assert rgb[0, 0].min() >= 254
    assert rgb[0, 1].tolist() == [0, 0, 0]
    assert rgb[0, 2, 0] >= 254 and rgb[0, 2, 1] <= 2 and rgb[0, 2, 2] <= 2
    assert rgb[0, 3].tolist() == [255, 255, 255]


# ---------------------------------------------------------------- camera (src/camera.cpp:5-57)
def test_camera_rays(oracle, scenes):
    s, _ = scenes.load_config("fire")
    W, H = s.width, s.height
    px = np.array([0, W - 1, 0, W - 1, W // 2]); py = np.array([0, 0, H - 1, H - 1, H // 2])
    o, d = oracle.camera_rays(s, px, py, np.zeros(5, np.float32), np.zeros(5, np.float32))
    assert np.array_equal(o, np.tile(np.array([[120, 30, 0]], np.float32), (5, 1)))
    assert np.allclose(np.linalg.norm(d, axis=1), 1, atol=1e-6)
    # centre pixel looks (almost) along look - position = -x
    assert d[4, 0] < -0.9999
    # raster (0,0) maps toward +left and +new_up (camera.cpp:21-31): up = +y; left = up x dir = (0,1,0)x(-1,0,0) = (0,0,1)
    assert d[0, 1] > 0 and d[0, 2] > 0 and d[3, 1] < 0 and d[3, 2] < 0
    # vertical half-angle = vfov/2 at the top edge (pixel centre offset 0.5)
    o2, d2 = oracle.camera_rays(s, np.array([W // 2]), np.array([0]), np.array([-0.5], np.float32), np.array([-0.5], np.float32))
    ang = np.degrees(np.arctan2(d2[0, 1], -d2[0, 0]))
    assert abs(ang - 18.5) < 0.05


# ---------------------------------------------------------------- tile/wave job space (src/tile_provider.cpp:14-105)
def test_ragged_tiles_cover_image_once(oracle, api, scenes):
    scene, spec = scenes.small_variant("fire", (37, 29), 3, 0.2)
    scene.tile_size[0], scene.tile_size[1] = 8, 5
    assert scene.ntiles == 5 * 6
    grids = api.synth_volume(spec.kind, spec.dims, spec.voxel_size, spec.seed, spec.anchor)
    vol = oracle.volume(grids)
    film = np.zeros((29, 37, 4), np.float32)
    vol.render(scene, film, threads=4)
    assert (film[..., 3] == 3).all()
    # waves are additive: [1..3] == [1..2] + [3]
    a = np.zeros_like(film); vol.render(scene, a, first_wave=1, n_waves=2); vol.render(scene, a, first_wave=3, n_waves=1)
    assert np.array_equal(a, film)


# ---------------------------------------------------------------- film checkpoint (film: main.cpp:58-60, wave counter: tile_provider.cpp:27-38)
def test_film_checkpoint_roundtrip_and_refusals(api, scenes, tmp_path):
    scene, _ = scenes.small_variant("fire", (37, 29), 8, 0.2)
    rng = np.random.default_rng(3)
    film = rng.random((29, 37, 4), dtype=np.float32)
    film[0, 0] = [np.float32(np.inf), np.float32(np.nan), -0.0, 5.0]  # bit patterns survive, not just values
    path = tmp_path / "render.film"
    api.film_save(path, scene, film, 6)
    assert not (tmp_path / "render.film.tmp").exists()
    got, nxt = api.film_load(path, scene)
    assert nxt == 6 and got.tobytes() == film.tobytes()
    # a resumed render may ask for more waves / other worker counts: not part of the fingerprint
    more = scene.copy(); more.num_waves = 64; more.num_workers = 3
    assert api.film_load(path, more)[1] == 6
    # anything that changes the film is: seed, camera, medium, tile size
    for mutate in (lambda s: setattr(s, "seed", s.seed + 1), lambda s: s.tile_size.__setitem__(0, 4),
                   lambda s: setattr(s.volume_parameters, "sigma_s", 0.5), lambda s: s.camera_parameters.position.__setitem__(1, 31.0)):
        other = scene.copy(); mutate(other)
        with pytest.raises(api.VptError) as e:
            api.film_load(path, other)
        assert e.value.code == -6 and "different scene" in str(e.value)
    other = scene.copy(); other.output_size[0] = 36
    with pytest.raises(api.VptError):
        api.film_load(path, other)
    # truncated and missing files are I/O errors; so is a file that is not a checkpoint
    data = path.read_bytes()
    (tmp_path / "cut.film").write_bytes(data[: len(data) // 2])
    (tmp_path / "junk.film").write_bytes(b"not a film" * 10)
    for name in ("cut.film", "junk.film", "missing.film"):
        with pytest.raises(api.VptError) as e:
            api.film_load(tmp_path / name, scene)
        assert e.value.code == -5


def test_baker_rejects_corrupt_grids(built, tmp_path):
    """vptc_set_volume takes raw NanoVDB buffers across the C ABI: the baker validates every offset before following it. An
    AddressSanitizer + UBSan build of the baker is fed randomly corrupted grids (headers, root table, child offsets, truncation): every
    run must end with 'baked / rejected' counts and no sanitizer report."""
    import shutil
    import subprocess
    if not shutil.which("g++"):
        pytest.skip("g++ absent")
    csrc = ROOT / "volume_path_tracer_b200" / "csrc"
    exe = tmp_path / "fuzz_bake"
    cmd = ["g++", "-std=c++20", "-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined", "-fno-omit-frame-pointer", "-ffp-contract=off",
           f"-I{ROOT / 'include'}", f"-I{csrc}", str(ROOT / "tests" / "fuzz" / "fuzz_bake.cpp"), str(csrc / "host" / "bake.cpp"),
           str(csrc / "host" / "nvdb_build.cpp"), str(csrc / "host" / "synth.cpp"), "-o", str(exe), "-lpthread"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if r.returncode != 0 and "asan" in r.stderr.lower():
        pytest.skip("sanitizer runtime not installed")
    assert r.returncode == 0, r.stderr[-2000:]
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=0")
    for seed, iters, kind in ((1, 250, "cloud"), (2, 120, "fire"), (3, 120, "dense")):
        r = subprocess.run([str(exe), str(seed), str(iters), kind], capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0 and "runtime error" not in r.stderr and "AddressSanitizer" not in r.stderr, (r.stdout + r.stderr)[-3000:]
        baked, rejected = [int(v) for v in r.stdout.split() if v.isdigit()][-2:]
        assert baked + rejected == iters and rejected > 0


def test_host_parsers_survive_mutated_input(api, scenes, tmp_path):
    """Seeded mutation fuzz of the three host-side readers behind the C ABI — scene JSON, `.nvdb` container, film checkpoint: every
    input either parses or raises VptError (error code + message); none may crash the process (they run in-process here)."""
    import random
    rnd = random.Random(5)
    base = scenes.config_json("fire")
    alphabet = '{}[]",:0123456789.-+eE tfnul\\\n\t' + "abcxyz_"
    outcomes = {"ok": 0, "err": 0}

    def attempt(fn):
        try:
            fn(); outcomes["ok"] += 1
        except api.VptError:
            outcomes["err"] += 1
        except UnicodeError:
            pass
    for _ in range(3000):
        t = list(base)
        for _ in range(rnd.choice([1, 1, 2, 3, 8])):
            if not t:
                break
            pos, op = rnd.randrange(len(t)), rnd.random()
            if op < 0.35: t[pos] = rnd.choice(alphabet)
            elif op < 0.65: del t[pos]
            elif op < 0.9: t.insert(pos, rnd.choice(alphabet))
            else: t = t[:pos]
        attempt(lambda: api.scene_from_json("".join(t)))
    for txt in ["", "{", "[" * 100000, "{" * 100000, '{"a":' * 50000, '"' + "\\" * 99999, "1e999999", '{"seed": 1e400}', base.replace("500", "9" * 400), base + base]:
        attempt(lambda: api.scene_from_json(txt))
    assert outcomes["err"] > 2000

    def mutate(raw, header):
        b = bytearray(raw)
        for _ in range(rnd.choice([1, 2, 4])):
            pos = rnd.randrange(0, header) if rnd.random() < 0.75 else rnd.randrange(len(b))
            if rnd.random() < 0.5:
                b[pos] = rnd.randrange(256)
            else:
                v, w = rnd.choice([0, 1, 0xffff, 0xffffffff, 0x7fffffff, 2**63, 2**64 - 1, len(b), len(b) * 2]), rnd.choice([2, 4, 8])
                pos = min(pos, len(b) - 8)
                b[pos:pos + w] = (v & (2 ** (8 * w) - 1)).to_bytes(w, "little")
        return bytes(b[: rnd.randrange(len(b))] if rnd.random() < 0.1 else b)
    g = api.synth_volume("fire", (24, 24, 24), 0.5, 2, None)
    p = tmp_path / "a.nvdb"
    g.write(p)
    raw = p.read_bytes()
    for _ in range(600):
        (tmp_path / "m.nvdb").write_bytes(mutate(raw, 400))
        attempt(lambda: api.VolumeGrids.read_from_file(tmp_path / "m.nvdb"))
    scene, _ = scenes.small_variant("fire", (37, 29), 8, 0.2)
    film = np.random.default_rng(1).random((29, 37, 4)).astype(np.float32)
    api.film_save(tmp_path / "a.ckpt", scene, film, 5)
    raw = (tmp_path / "a.ckpt").read_bytes()
    for _ in range(600):
        (tmp_path / "m.ckpt").write_bytes(mutate(raw, 300))
        attempt(lambda: api.film_load(tmp_path / "m.ckpt", scene))
    assert outcomes["ok"] > 0


def test_nvdb_reader_ends_on_hostile_size_fields(api, tmp_path):
    """Size fields that point backwards, past the end of the file or at 2^64-ish values must end in an error quickly (round-1 advisor
    finding: a raw dump whose second grid header carried gridSize = -672 rewound the reader onto offset 0 forever; a segment whose
    fileSize was negative did the same). Runs in a child process under a timeout so that a regression is a failure, not a hang."""
    import struct
    import subprocess
    import sys
    g = api.synth_volume("fire", (16, 16, 16), 0.5, 2, None)
    good = tmp_path / "good.nvdb"
    g.write(good)
    seg = good.read_bytes()
    raw = bytes(g.density.view())
    hdr = bytearray(raw[:672])
    files = {}
    a = bytearray(hdr); a[32:40] = struct.pack("<Q", 672)                    # header 1: gridSize = 672 (its own header only) ...
    a[40:44] = b"zzz\0"                                                        # (a name that never matches)
    b = bytearray(a); b[32:40] = struct.pack("<Q", (1 << 64) - 672)            # ... header 2: gridSize = (uint64)-672 -> rewind to 0
    files["raw_rewind"] = bytes(a + b)
    files["raw_huge"] = bytes(a[:32] + struct.pack("<Q", 1 << 62) + a[40:])
    files["raw_zero"] = bytes(a[:32] + struct.pack("<Q", 0) + a[40:])
    meta_off = 16                                                              # FileHeader is 16 B; FileMetaData.fileSize at +8
    for label, v in (("seg_negative", (1 << 64) - (16 + 176 + 8)), ("seg_huge", 1 << 62), ("seg_zero", 0)):
        s = bytearray(seg)
        s[meta_off + 8:meta_off + 16] = struct.pack("<Q", v)
        name_size = struct.unpack_from("<I", s, meta_off + 136)[0]
        s[meta_off + 176:meta_off + 176 + name_size] = b"x" * (name_size - 1) + b"\0"  # never matches: the reader must SKIP this grid
        files[label] = bytes(s) + seg
    s = bytearray(seg); s[meta_off:meta_off + 8] = struct.pack("<Q", 1 << 61); files["seg_gridsize_huge"] = bytes(s)
    for label, data in files.items():
        (tmp_path / f"{label}.nvdb").write_bytes(data)
    code = ("import sys; sys.path.insert(0, %r)\n"
            "from volume_path_tracer_b200 import api\n"
            "import glob\n"
            "for p in sorted(glob.glob(%r)):\n"
            "    for name in ('density', 'temperature', 'nope'):\n"
            "        try:\n"
            "            api.VolumeGrids.read_from_file(p) if name == 'density' else api.lib().vptc_nvdb_read_grid(p.encode(), name.encode(), api.C.byref(api.C.c_void_p()), api.C.byref(api.C.c_size_t()))\n"
            "        except api.VptError:\n"
            "            pass\n"
            "print('done')\n") % (str(ROOT), str(tmp_path / "*.nvdb"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "done" in r.stdout, r.stderr[-2000:]
    # and the well-formed file still reads
    assert api.VolumeGrids.read_from_file(good).density.nbytes == g.density.nbytes
