"""ctypes binding of libvptcuda (include/vptc.h) + a thin mirror of the reference's objects.

The product is the C-ABI library; this module only marshals numpy / torch buffers into it. There is
no CPU fallback: importing works without a GPU (so the symbol table can be checked), but creating a
``Renderer`` without one raises, and a missing library raises at import of :func:`lib`.

Reference objects mirrored (names and argument meaning as in the reference):
  Configuration / read_configuration   include/vpt/configuration.hpp:55-67
  VolumeGrids.read_from_file            src/volume_grids.cpp:58-67
  Renderer.run(...)                     vpt::run, include/vpt/worker.hpp:11  (+ main.cpp:63-87)
  film_to_image                         src/main.cpp:12-24
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_PKG = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ["VPT_LIBVPTCUDA"]) if os.environ.get("VPT_LIBVPTCUDA") else _PKG / "libvptcuda.so"  # env: development builds only


# ---------------------------------------------------------------- POD structs (include/vptc.h)
class CameraParameters(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("look", C.c_float * 3), ("up", C.c_float * 3),
                ("vfov_deg", C.c_float), ("imaging_ratio", C.c_float)]


class InfiniteLight(C.Structure):
    _fields_ = [("xyz", C.c_float * 3), ("multiplier", C.c_float)]


class DistantLight(C.Structure):
    _fields_ = [("xyz", C.c_float * 3), ("multiplier", C.c_float), ("inv_direction", C.c_float * 3)]


class WorkerParameters(C.Structure):
    _fields_ = [("single_pixel_enabled", C.c_int32), ("single_pixel_coord", C.c_int64 * 2), ("use_jitter", C.c_int32),
                ("infinite_light", InfiniteLight), ("distant_light", DistantLight), ("max_depth", C.c_uint32)]


class VolumeParameters(C.Structure):
    _fields_ = [("henyey_greenstein_g", C.c_float), ("le_scale", C.c_float), ("sigma_a", C.c_float), ("sigma_s", C.c_float),
                ("temperature_offset", C.c_float), ("temperature_scale", C.c_float)]


class Scene(C.Structure):
    """vptc_scene == vpt::Configuration minus volume_path."""
    _fields_ = [("seed", C.c_uint32), ("output_size", C.c_int64 * 2), ("tile_size", C.c_int64 * 2), ("num_waves", C.c_uint32),
                ("num_workers", C.c_uint32), ("camera_parameters", CameraParameters), ("worker_parameters", WorkerParameters),
                ("volume_parameters", VolumeParameters)]

    @property
    def width(self) -> int: return int(self.output_size[0])

    @property
    def height(self) -> int: return int(self.output_size[1])

    @property
    def ntiles(self) -> int:
        W, H, tw, th = self.width, self.height, int(self.tile_size[0]), int(self.tile_size[1])
        return (-(-W // tw)) * (-(-H // th))

    def copy(self) -> "Scene":
        s = Scene()
        C.memmove(C.byref(s), C.byref(self), C.sizeof(Scene))
        return s


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("pixel_samples", "density_evals", "temperature_evals", "majorant_fetches", "rng_draws",
                                          "scatter_events", "shadow_rays")]

    def as_dict(self) -> dict:
        return {n: int(getattr(self, n)) for n, _ in self._fields_}

    def algorithmic_bytes(self) -> int:
        """32*N_rho + 32*N_T + 4*N_maj + 32 per pixel-sample."""
        return 32 * self.density_evals + 32 * self.temperature_evals + 4 * self.majorant_fetches + 32 * self.pixel_samples


class JobRange(C.Structure):
    """vptc_job_range: waves first_wave + k*wave_stride (k < n_waves) of the tiles [tile_begin, tile_begin + tile_count); tile_count 0 = all."""
    _fields_ = [("first_wave", C.c_uint32), ("n_waves", C.c_uint32), ("wave_stride", C.c_uint32), ("tile_row_stride", C.c_uint32),
                ("tile_begin", C.c_uint64), ("tile_count", C.c_uint64)]


class AdaptiveParams(C.Structure):
    """vptc_adaptive_params"""
    _fields_ = [("max_waves", C.c_uint32), ("batch_waves", C.c_uint32), ("min_waves", C.c_uint32), ("rel_error", C.c_float)]


class VptError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"vptc error {code}: {msg}")
        self.code = code


EXPORTS = [
    "vptc_create", "vptc_destroy", "vptc_last_error", "vptc_set_volume", "vptc_render", "vptc_render_device", "vptc_render_device_range", "vptc_render_device_async", "vptc_render_async", "vptc_wait", "vptc_render_adaptive", "vptc_last_render_ms", "vptc_device_volume_bytes",
    "vptc_get_counters", "vptc_enable_counters", "vptc_set_kernel_variant", "vptc_trace_jobs", "vptc_trace_segments", "vptc_trace_pixel_csv", "vptc_trace_path_log", "vptc_sample_grid", "vptc_film_to_image", "vptc_film_to_image_static", "vptc_blackbody",
    "vptc_scene_from_json_file", "vptc_scene_from_json", "vptc_nvdb_read_grid", "vptc_nvdb_write_grids", "vptc_free",
    "vptc_synth_volume", "vptc_write_png", "vptc_film_save", "vptc_film_load", "vptc_version", "vptc_host_last_error", "vptc_dda_trace_csv", "vptc_dda_trace_ray_csv",
    "vptc_group_create", "vptc_group_destroy", "vptc_group_last_error", "vptc_group_size", "vptc_group_peer_mapped", "vptc_group_set_volume", "vptc_group_render",
    "vptc_group_render_async", "vptc_group_wait", "vptc_group_last_render_ms", "vptc_group_context",
]

HOST_EXPORTS = ["vptc_host_last_error", "vptc_scene_from_json_file", "vptc_scene_from_json", "vptc_nvdb_read_grid", "vptc_nvdb_write_grids", "vptc_free",
                "vptc_synth_volume", "vptc_write_png", "vptc_film_save", "vptc_film_load", "vptc_version", "vptc_dda_trace_csv", "vptc_dda_trace_ray_csv"]
HOST_LIB_PATH = _PKG / "libvpthost.so"

_lib = None
_host_lib = None


def _declare_host(L):
    vp, sz, u32 = C.c_void_p, C.c_size_t, C.c_uint32
    L.vptc_host_last_error.restype = C.c_char_p
    L.vptc_version.restype = C.c_char_p
    L.vptc_scene_from_json_file.argtypes = [C.c_char_p, C.POINTER(Scene), C.c_char_p, sz]
    L.vptc_scene_from_json.argtypes = [C.c_char_p, C.POINTER(Scene), C.c_char_p, sz]
    L.vptc_nvdb_read_grid.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(vp), C.POINTER(sz)]
    L.vptc_nvdb_write_grids.argtypes = [C.c_char_p, C.POINTER(vp), sz]
    L.vptc_free.argtypes = [vp]
    L.vptc_free.restype = None
    L.vptc_synth_volume.argtypes = [C.c_char_p, C.POINTER(C.c_int32), C.c_double, u32, C.POINTER(C.c_double), C.POINTER(vp), C.POINTER(sz), C.POINTER(vp), C.POINTER(sz)]
    L.vptc_write_png.argtypes = [C.c_char_p, vp, C.c_int64, C.c_int64]
    L.vptc_film_save.argtypes = [C.c_char_p, C.POINTER(Scene), vp, u32]
    L.vptc_film_load.argtypes = [C.c_char_p, C.POINTER(Scene), vp, C.POINTER(u32)]
    L.vptc_dda_trace_csv.argtypes = [vp, sz, C.POINTER(Scene), C.c_int64, C.c_int64, C.c_char_p, C.POINTER(sz)]
    L.vptc_dda_trace_ray_csv.argtypes = [vp, sz, vp, vp, C.c_char_p, C.POINTER(sz)]


def host_lib() -> C.CDLL:
    """The host-side helpers (scene JSON, .nvdb I/O, procedural volumes, PNG, film checkpoints): libvpthost.so, which has no CUDA
    dependency — CPU-only tools (bench.py --impl reference, the oracle tests) never map the CUDA library. A development build chosen
    with VPT_LIBVPTCUDA serves its own copies of the helpers."""
    global _host_lib
    if _host_lib is not None:
        return _host_lib
    if os.environ.get("VPT_LIBVPTCUDA"):
        _host_lib = lib()
        return _host_lib
    if not HOST_LIB_PATH.exists():
        raise ImportError(f"{HOST_LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` (make -C volume_path_tracer_b200/csrc)")
    L = C.CDLL(str(HOST_LIB_PATH))
    _declare_host(L)
    _host_lib = L
    return L


def _check_host(rc: int):
    if rc != 0:
        raise VptError(rc, (host_lib().vptc_host_last_error() or b"").decode(errors="replace"))


def lib() -> C.CDLL:
    """Load the in-tree C-ABI library. Fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(make -C volume_path_tracer_b200/csrc). There is no fallback implementation.")
    L = C.CDLL(str(LIB_PATH))
    vp, sz, u32, u64p = C.c_void_p, C.c_size_t, C.c_uint32, C.POINTER(C.c_uint64)
    L.vptc_last_error.restype = C.c_char_p
    L.vptc_last_error.argtypes = [vp]
    L.vptc_create.argtypes = [C.POINTER(vp), C.c_int]
    L.vptc_destroy.argtypes = [vp]
    L.vptc_destroy.restype = None
    L.vptc_set_volume.argtypes = [vp, vp, sz, vp, sz]
    L.vptc_render.argtypes = [vp, C.POINTER(Scene), u32, u32, vp]
    L.vptc_render_device.argtypes = [vp, C.POINTER(Scene), u32, u32, u32, vp]
    L.vptc_render_device_range.argtypes = [vp, C.POINTER(Scene), C.POINTER(JobRange), vp]
    L.vptc_render_device_async.argtypes = [vp, C.POINTER(Scene), C.POINTER(JobRange), vp]
    L.vptc_render_async.argtypes = [vp, C.POINTER(Scene), u32, u32, vp]
    L.vptc_wait.argtypes = [vp]
    L.vptc_render_adaptive.argtypes = [vp, C.POINTER(Scene), u32, C.POINTER(AdaptiveParams), vp, vp, vp]
    L.vptc_last_render_ms.argtypes = [vp]
    L.vptc_last_render_ms.restype = C.c_float
    L.vptc_device_volume_bytes.argtypes = [vp]
    L.vptc_device_volume_bytes.restype = C.c_uint64
    L.vptc_get_counters.argtypes = [vp, C.POINTER(Counters)]
    L.vptc_enable_counters.argtypes = [vp, C.c_int]
    L.vptc_set_kernel_variant.argtypes = [vp, C.c_int]
    L.vptc_trace_jobs.argtypes = [vp, C.POINTER(Scene), vp, sz, vp, vp]
    L.vptc_trace_segments.argtypes = [vp, vp, vp, sz, u32, vp, vp]
    L.vptc_trace_pixel_csv.argtypes = [vp, C.POINTER(Scene), C.c_int64, C.c_int64, C.c_char_p]
    L.vptc_trace_path_log.argtypes = [vp, C.POINTER(Scene), C.c_uint64, C.c_int64, C.c_int64, vp, sz, C.POINTER(sz)]
    L.vptc_sample_grid.argtypes = [vp, C.c_int, vp, sz, vp, vp, vp]
    L.vptc_film_to_image.argtypes = [vp, vp, C.c_int64, C.c_int64, vp]
    L.vptc_film_to_image_static.argtypes = [vp, vp, C.c_int64, C.c_int64, vp]
    L.vptc_blackbody.argtypes = [vp, vp, sz, vp]
    _declare_host(L)
    L.vptc_group_create.argtypes = [C.POINTER(vp), C.POINTER(C.c_int), C.c_int]
    L.vptc_group_destroy.argtypes = [vp]
    L.vptc_group_destroy.restype = None
    L.vptc_group_last_error.argtypes = [vp]
    L.vptc_group_last_error.restype = C.c_char_p
    L.vptc_group_size.argtypes = [vp]
    L.vptc_group_peer_mapped.argtypes = [vp]
    L.vptc_group_set_volume.argtypes = [vp, vp, sz, vp, sz]
    L.vptc_group_render.argtypes = [vp, C.POINTER(Scene), u32, u32, C.c_int, vp]
    L.vptc_group_render_async.argtypes = [vp, C.POINTER(Scene), u32, u32, C.c_int]
    L.vptc_group_wait.argtypes = [vp, vp]
    L.vptc_group_last_render_ms.argtypes = [vp]
    L.vptc_group_last_render_ms.restype = C.c_float
    L.vptc_group_context.argtypes = [vp, C.c_int]
    L.vptc_group_context.restype = vp
    _lib = L
    return L


def _check(rc: int, ctx=None):
    if rc != 0:
        raise VptError(rc, (lib().vptc_last_error(ctx) or b"").decode(errors="replace"))


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


# ---------------------------------------------------------------- grid buffers
class GridBuffer:
    """A serialized NanoVDB float grid owned by the C library (freed on garbage collection)."""

    def __init__(self, ptr: int, nbytes: int):
        self.ptr, self.nbytes = ptr, nbytes

    def __del__(self):
        try:
            if getattr(self, "ptr", None):
                host_lib().vptc_free(self.ptr)
                self.ptr = None
        except Exception:  # interpreter shutdown
            pass

    def view(self) -> np.ndarray:
        return np.ctypeslib.as_array((C.c_uint8 * self.nbytes).from_address(self.ptr))

    def __len__(self): return self.nbytes


def synth_volume(kind: str, dims, voxel_size: float, seed: int = 1, anchor=None) -> "VolumeGrids":
    """Procedural volumes of SURVEY.md §8(d) as NanoVDB grids. anchor = world position of the box centre
    (centre of the base for "fire"); default the world origin."""
    d, db, t, tb = C.c_void_p(), C.c_size_t(), C.c_void_p(), C.c_size_t()
    anc = (C.c_double * 3)(*[float(x) for x in anchor]) if anchor is not None else None
    _check_host(host_lib().vptc_synth_volume(kind.encode(), (C.c_int32 * 3)(*[int(x) for x in dims]), float(voxel_size), int(seed), anc,
                                   C.byref(d), C.byref(db), C.byref(t), C.byref(tb)))
    return VolumeGrids(GridBuffer(d.value, db.value), GridBuffer(t.value, tb.value) if t.value else None)


class VolumeGrids:
    """src/volume_grids.cpp:21-67 — density required, temperature optional."""

    def __init__(self, density: GridBuffer, temperature: GridBuffer | None = None):
        self.density, self.temperature = density, temperature

    def has_temperature(self) -> bool: return self.temperature is not None

    @staticmethod
    def read_from_file(path) -> "VolumeGrids":
        def read(name, required):
            p, n = C.c_void_p(), C.c_size_t()
            rc = host_lib().vptc_nvdb_read_grid(str(path).encode(), name.encode(), C.byref(p), C.byref(n))
            if rc != 0:
                if required:
                    _check_host(rc)
                return None
            return GridBuffer(p.value, n.value)
        return VolumeGrids(read("density", True), read("temperature", False))

    def write(self, path):
        grids = [self.density] + ([self.temperature] if self.temperature else [])
        arr = (C.c_void_p * len(grids))(*[g.ptr for g in grids])
        _check_host(host_lib().vptc_nvdb_write_grids(str(path).encode(), arr, len(grids)))


def read_configuration(path):
    """-> (Scene, volume_path as written in the file).  src/configuration.cpp:8-22"""
    s, buf = Scene(), C.create_string_buffer(4096)
    _check_host(host_lib().vptc_scene_from_json_file(str(path).encode(), C.byref(s), buf, 4096))
    return s, buf.value.decode()


def scene_from_json(text: str):
    s, buf = Scene(), C.create_string_buffer(4096)
    _check_host(host_lib().vptc_scene_from_json(text.encode(), C.byref(s), buf, 4096))
    return s, buf.value.decode()


def dda_trace_csv(grids: "VolumeGrids", path, scene: "Scene" = None, px: int = 0, py: int = 0, origin=None, direction=None) -> int:
    """dda_trace.csv (Volume::log_dda_trace, volume.cpp:194-225) of the camera ray through pixel (px, py), or of an explicit world ray.
    -> number of voxel rows written (0 and no file when the ray misses the grid)."""
    n = C.c_size_t(0)
    if origin is not None:
        o, d = np.ascontiguousarray(origin, np.float32), np.ascontiguousarray(direction, np.float32)
        _check_host(host_lib().vptc_dda_trace_ray_csv(grids.density.ptr, grids.density.nbytes, _ptr(o), _ptr(d), str(path).encode(), C.byref(n)))
    else:
        _check_host(host_lib().vptc_dda_trace_csv(grids.density.ptr, grids.density.nbytes, C.byref(scene), int(px), int(py), str(path).encode(), C.byref(n)))
    return int(n.value)


def write_png(path, rgb: np.ndarray):
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    _check_host(host_lib().vptc_write_png(str(path).encode(), _ptr(rgb), rgb.shape[1], rgb.shape[0]))


def film_save(path, scene: "Scene", film: np.ndarray, next_wave: int):
    """vptc_film_save: checkpoint of a progressive render (float4 film + the next wave to render)."""
    film = np.ascontiguousarray(film, dtype=np.float32)
    assert film.shape == (scene.height, scene.width, 4)
    _check_host(host_lib().vptc_film_save(str(path).encode(), C.byref(scene), _ptr(film), int(next_wave)))


def film_load(path, scene: "Scene"):
    """vptc_film_load -> (film, next_wave); raises VptError when the file is missing/truncated or belongs to another scene."""
    film = np.zeros((scene.height, scene.width, 4), np.float32)
    nxt = C.c_uint32(0)
    _check_host(host_lib().vptc_film_load(str(path).encode(), C.byref(scene), _ptr(film), C.byref(nxt)))
    return film, int(nxt.value)


# ---------------------------------------------------------------- the renderer
class Renderer:
    """One CUDA device's context (vptc_ctx)."""

    def __init__(self, device: int = 0):
        self._ctx = C.c_void_p()
        _check(lib().vptc_create(C.byref(self._ctx), int(device)))
        self.device = device

    def close(self):
        if self._ctx:
            lib().vptc_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_volume(self, grids: VolumeGrids):
        t = grids.temperature
        _check(lib().vptc_set_volume(self._ctx, grids.density.ptr, grids.density.nbytes, t.ptr if t else None, t.nbytes if t else 0), self._ctx)

    def enable_counters(self, on: bool = True):
        _check(lib().vptc_enable_counters(self._ctx, int(on)), self._ctx)

    def set_kernel_variant(self, variant: int):
        """2 = flat + warp-level stage scheduling (default), 1 = flat state machine, 0 = first nested megakernel."""
        _check(lib().vptc_set_kernel_variant(self._ctx, int(variant)), self._ctx)

    def counters(self) -> Counters:
        c = Counters()
        _check(lib().vptc_get_counters(self._ctx, C.byref(c)), self._ctx)
        return c

    def device_volume_bytes(self) -> int:
        return int(lib().vptc_device_volume_bytes(self._ctx))

    def last_render_ms(self) -> float:
        return float(lib().vptc_last_render_ms(self._ctx))

    def new_film(self, scene: Scene) -> np.ndarray:
        return np.zeros((scene.height, scene.width, 4), np.float32)

    def run(self, scene: Scene, film: np.ndarray, first_wave: int = 1, n_waves: int | None = None):
        """vpt::run for waves [first_wave, first_wave + n_waves): adds into the HOST film (H, W, 4) float32."""
        assert film.dtype == np.float32 and film.flags.c_contiguous and film.shape == (scene.height, scene.width, 4)
        _check(lib().vptc_render(self._ctx, C.byref(scene), first_wave, scene.num_waves if n_waves is None else n_waves, _ptr(film)), self._ctx)
        return film

    def run_device(self, scene: Scene, d_film_ptr: int, first_wave: int, n_waves: int, wave_stride: int = 1):
        """Same, film resident on the device (raw device pointer, e.g. torch_tensor.data_ptr())."""
        _check(lib().vptc_render_device(self._ctx, C.byref(scene), first_wave, n_waves, wave_stride, C.c_void_p(d_film_ptr)), self._ctx)

    def run_device_range(self, scene: Scene, d_film_ptr: int, first_wave: int, n_waves: int, wave_stride: int = 1, tile_begin: int = 0, tile_count: int = 0,
                         wait: bool = True, tile_row_stride: int = 1):
        """A slice of the job space (vptc_job_range) into a device film; wait=False enqueues (vptc_render_device_async) — finish with wait()."""
        jr = JobRange(first_wave, n_waves, wave_stride, tile_row_stride, tile_begin, tile_count)
        fn = lib().vptc_render_device_range if wait else lib().vptc_render_device_async
        _check(fn(self._ctx, C.byref(scene), C.byref(jr), C.c_void_p(d_film_ptr)), self._ctx)

    def run_async(self, scene: Scene, film: np.ndarray, first_wave: int, n_waves: int):
        """vptc_render_async: enqueue waves into the HOST film; the film is complete after the next run_async or wait()."""
        assert film.dtype == np.float32 and film.flags.c_contiguous and film.shape == (scene.height, scene.width, 4)
        _check(lib().vptc_render_async(self._ctx, C.byref(scene), first_wave, n_waves, _ptr(film)), self._ctx)

    def wait(self):
        _check(lib().vptc_wait(self._ctx), self._ctx)

    def run_adaptive(self, scene: Scene, film: np.ndarray, max_waves: int, batch_waves: int, rel_error: float, min_waves: int = 0, first_wave: int = 1):
        """vptc_render_adaptive -> (film, sum_y2 (H, W), waves per tile): each tile stops when its relative standard error <= rel_error."""
        assert film.dtype == np.float32 and film.flags.c_contiguous and film.shape == (scene.height, scene.width, 4)
        sq = np.zeros((scene.height, scene.width), np.float32)
        tw = np.zeros(scene.ntiles, np.uint32)
        p = AdaptiveParams(max_waves, batch_waves, min_waves, rel_error)
        _check(lib().vptc_render_adaptive(self._ctx, C.byref(scene), first_wave, C.byref(p), _ptr(film), _ptr(sq), _ptr(tw)), self._ctx)
        return film, sq, tw

    def trace_jobs(self, scene: Scene, jids):
        jids = np.ascontiguousarray(jids, dtype=np.uint64)
        slots = int(scene.tile_size[0]) * int(scene.tile_size[1])
        xyz = np.zeros((len(jids), slots, 3), np.float32)
        nd = np.zeros((len(jids), slots), np.uint32)
        _check(lib().vptc_trace_jobs(self._ctx, C.byref(scene), _ptr(jids), len(jids), _ptr(xyz), _ptr(nd)), self._ctx)
        return xyz, nd

    def trace_segments(self, origins, dirs, max_segs: int = 256):
        o = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        segs = np.zeros((len(o), max_segs, 3), np.float32)
        n = np.zeros(len(o), np.uint32)
        _check(lib().vptc_trace_segments(self._ctx, _ptr(o), _ptr(d), len(o), max_segs, _ptr(segs), _ptr(n)), self._ctx)
        return segs, n

    def trace_path_log(self, scene, jid: int, px: int, py: int, cap: int = 1 << 16) -> np.ndarray:
        """Logger<true> + single_pixel (worker.cpp:15-49): events of the path job `jid` traces through (px, py); rows of 8 floats."""
        ev, n = np.zeros((cap, 8), np.float32), C.c_size_t(0)
        _check(lib().vptc_trace_path_log(self._ctx, C.byref(scene), jid, px, py, _ptr(ev), cap, C.byref(n)), self._ctx)
        return ev[: min(n.value, cap)].copy()

    def trace_pixel_csv(self, scene, px: int, py: int, path):
        """majorant_trace.csv of the pixel-centre camera ray (Volume::log_majorant_trace, volume.cpp:176-192)."""
        _check(lib().vptc_trace_pixel_csv(self._ctx, C.byref(scene), int(px), int(py), str(path).encode()), self._ctx)

    def sample_grid(self, grid: int, index_xyz):
        p = np.ascontiguousarray(index_xyz, dtype=np.float32).reshape(-1, 3)
        v, dm, mj = np.zeros(len(p), np.float32), np.zeros(len(p), np.uint32), np.zeros(len(p), np.float32)
        _check(lib().vptc_sample_grid(self._ctx, grid, _ptr(p), len(p), _ptr(v), _ptr(dm), _ptr(mj)), self._ctx)
        return v, dm, mj

    def blackbody(self, temps_k) -> np.ndarray:
        """The device's blackbody_radiation_xyz (precompute_blackbody.cpp:24-52) for temperatures in kelvin -> (n, 3) XYZ."""
        t = np.ascontiguousarray(temps_k, dtype=np.float32).reshape(-1)
        out = np.zeros((len(t), 3), np.float32)
        _check(lib().vptc_blackbody(self._ctx, _ptr(t), len(t), _ptr(out)), self._ctx)
        return out

    def film_to_image(self, film: np.ndarray) -> np.ndarray:
        film = np.ascontiguousarray(film, dtype=np.float32)
        rgb = np.zeros(film.shape[:2] + (3,), np.uint8)
        _check(lib().vptc_film_to_image(self._ctx, _ptr(film), film.shape[1], film.shape[0], _ptr(rgb)), self._ctx)
        return rgb

    def film_to_image_static(self, film: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        """vptc_film_to_image_static: no allocation on the device (static scratch, at most STATIC_IMAGE_MAX_PIXELS pixels)."""
        film = np.ascontiguousarray(film, dtype=np.float32)
        rgb = out if out is not None else np.zeros(film.shape[:2] + (3,), np.uint8)
        _check(lib().vptc_film_to_image_static(self._ctx, _ptr(film), film.shape[1], film.shape[0], _ptr(rgb)), self._ctx)
        return rgb


SHARD_WAVES, SHARD_TILES = 0, 1
STATIC_IMAGE_MAX_PIXELS = 1 << 21  # VPTC_STATIC_IMAGE_MAX_PIXELS (include/vptc.h)


class Group:
    """vptc_group: several GPUs behind one handle, driven by this one host thread (the reference's worker pool, src/main.cpp:63-87).
    The volume is replicated, the job space sharded by waves (films summed over peer memory) or by stripes of tile rows (films gathered)."""

    def __init__(self, devices):
        devices = [int(d) for d in devices]
        self._g = C.c_void_p()
        arr = (C.c_int * len(devices))(*devices)
        _check(lib().vptc_group_create(C.byref(self._g), arr, len(devices)))
        self.devices = devices

    def _chk(self, rc):
        if rc != 0:
            raise VptError(rc, (lib().vptc_group_last_error(self._g) or b"").decode(errors="replace"))

    def close(self):
        if self._g:
            lib().vptc_group_destroy(self._g)
            self._g = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def size(self) -> int: return int(lib().vptc_group_size(self._g))
    def peer_mapped(self) -> int: return int(lib().vptc_group_peer_mapped(self._g))
    def last_render_ms(self) -> float: return float(lib().vptc_group_last_render_ms(self._g))

    def set_volume(self, grids: VolumeGrids):
        t = grids.temperature
        self._chk(lib().vptc_group_set_volume(self._g, grids.density.ptr, grids.density.nbytes, t.ptr if t else None, t.nbytes if t else 0))

    def run(self, scene: Scene, film: np.ndarray, first_wave: int = 1, n_waves: int | None = None, sharding: int = SHARD_WAVES):
        assert film.dtype == np.float32 and film.flags.c_contiguous and film.shape == (scene.height, scene.width, 4)
        self._chk(lib().vptc_group_render(self._g, C.byref(scene), first_wave, scene.num_waves if n_waves is None else n_waves, sharding, _ptr(film)))
        return film

    def run_async(self, scene: Scene, first_wave: int, n_waves: int, sharding: int = SHARD_WAVES):
        self._chk(lib().vptc_group_render_async(self._g, C.byref(scene), first_wave, n_waves, sharding))

    def wait(self, film: np.ndarray):
        assert film.dtype == np.float32 and film.flags.c_contiguous
        self._chk(lib().vptc_group_wait(self._g, _ptr(film)))
        return film
