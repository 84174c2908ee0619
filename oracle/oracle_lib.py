"""ctypes binding of the CPU oracle — TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
Two builds of one C API (oracle/oracle_capi.cpp):
  "own": oracle/libvpt_oracle.so           self-contained restatement, (re)built on demand with g++
  "ref": oracle/_ref/libvpt_oracle_ref.so  on the reference's vendored NanoVDB/pcg headers; built only
         where /root/reference exists, travels to the GPU box as a prebuilt file
A third library, oracle/_ref/libvpt_reference.so (class Reference), is NOT a restatement: it is the reference's own
src/{worker,volume,majorant_transmittance_sampler,camera,precompute_blackbody,spectral,tile_provider,volume_grids}.cpp
compiled unchanged (Eigen stand-in: oracle/ref/eigen_shim). It pins the restatement (tests/test_reference_compiled.py)
and is the "reference" CPU baseline of bench.py.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
OWN_PATH = _DIR / "libvpt_oracle.so"
REF_PATH = _DIR / "_ref" / "libvpt_oracle_ref.so"
REFSRC_PATH = _DIR / "_ref" / "libvpt_reference.so"  # the reference's own src/*.cpp, compiled unchanged (ref/reference_capi.cpp)
REFERENCE_ROOT = Path(os.environ.get("VPT_REFERENCE_DIR", "/root/reference"))


def build(ref: bool = True) -> None:
    """make the oracle libraries (the checker). `ref` is skipped when the reference tree is absent."""
    subprocess.run(["make", "-C", str(_DIR), "-s", "libvpt_oracle.so", "libvpt_libm_check.so"], check=True)
    if ref and (REFERENCE_ROOT / "external" / "openvdb" / "nanovdb").is_dir():
        subprocess.run(["make", "-C", str(_DIR), "-s", "-j", str(min(8, os.cpu_count() or 1)), "ref", f"REF={REFERENCE_ROOT}"], check=True)


def libm_check(stride: int = 1):
    """Sweep the device's glibc-exact logf/sinf/cosf restatement (compiled for the host) against this machine's
    libm. -> (n_logf, mismatches_logf, n_sincos, mismatches_sincos); stride 1 is exhaustive (~20-40 s)."""
    path = _DIR / "libvpt_libm_check.so"
    if not path.exists():
        build(ref=False)
    L = C.CDLL(str(path))
    out = (C.c_uint64 * 4)()
    L.vptm_check(C.c_uint32(stride), out)
    return tuple(int(x) for x in out)


def libm_check2(stride: int = 1):
    """Same for expf on +-[2^-126, 88) and powf(x, 1/2.4f) on (0.0031308, FLT_MAX]. -> (n_expf, mismatches_expf, n_powf, mismatches_powf)."""
    path = _DIR / "libvpt_libm_check.so"
    if not path.exists():
        build(ref=False)
    L = C.CDLL(str(path))
    out = (C.c_uint64 * 4)()
    L.vptm_check2(C.c_uint32(stride), out)
    return tuple(int(x) for x in out)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Oracle:
    """One oracle library ("own" or "ref")."""

    def __init__(self, flavor: str = "own"):
        path = OWN_PATH if flavor == "own" else REF_PATH
        if flavor == "own" and not path.exists():
            build(ref=False)
        if not path.exists():
            raise FileNotFoundError(f"{path} not built (flavor {flavor!r})")
        self.flavor = flavor
        L = self.L = C.CDLL(str(path))
        L.vpto_substrate.restype = C.c_char_p
        vp, sz = C.c_void_p, C.c_size_t
        L.vpto_create.argtypes = [C.POINTER(vp), vp, sz, vp, sz]
        L.vpto_destroy.argtypes = [vp]
        L.vpto_destroy.restype = None
        L.vpto_leaf_maxima.argtypes = [vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
        L.vpto_render.argtypes = [vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp, C.POINTER(C.c_double)]
        L.vpto_render_sample.argtypes = [vp, vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, vp, vp, C.POINTER(C.c_double)]
        L.vpto_trace_jobs.argtypes = [vp, vp, vp, sz, vp, vp, vp]
        L.vpto_trace_segments.argtypes = [vp, vp, vp, sz, C.c_uint32, vp, vp]
        L.vpto_path_log.argtypes = [vp, vp, C.c_uint64, C.c_int64, C.c_int64, vp, sz, C.POINTER(sz)]
        L.vpto_sample_grid.argtypes = [vp, C.c_int, vp, sz, vp, vp, vp]
        L.vpto_rng.argtypes = [C.c_uint32, C.c_uint64, sz, C.POINTER(C.c_uint64), vp, vp]
        L.vpto_rng.restype = None
        L.vpto_camera_rays.argtypes = [vp, sz, vp, vp, vp, vp, vp, vp]
        L.vpto_camera_rays.restype = None
        L.vpto_blackbody_table.argtypes = [vp]
        L.vpto_blackbody_table.restype = None
        L.vpto_blackbody.argtypes = [vp, sz, vp]
        L.vpto_blackbody.restype = None
        L.vpto_hg_sample.argtypes = [vp, vp, sz, C.c_float, vp]
        L.vpto_hg_sample.restype = None
        L.vpto_hg_eval.argtypes = [vp, sz, C.c_float, vp]
        L.vpto_hg_eval.restype = None
        L.vpto_film_to_image.argtypes = [vp, C.c_int64, C.c_int64, vp]
        L.vpto_film_to_image.restype = None

    def substrate(self) -> str:
        return self.L.vpto_substrate().decode()

    # ---- stateless pieces ----
    def rng(self, seed: int, jid: int, n: int):
        s0 = C.c_uint64()
        u, f = np.zeros(n, np.uint32), np.zeros(n, np.float32)
        self.L.vpto_rng(seed, jid, n, C.byref(s0), _ptr(u), _ptr(f))
        return s0.value, u, f

    def camera_rays(self, scene, px, py, jx, jy):
        px, py = np.ascontiguousarray(px, np.int64), np.ascontiguousarray(py, np.int64)
        jx, jy = np.ascontiguousarray(jx, np.float32), np.ascontiguousarray(jy, np.float32)
        o, d = np.zeros((len(px), 3), np.float32), np.zeros((len(px), 3), np.float32)
        self.L.vpto_camera_rays(C.addressof(scene), len(px), _ptr(px), _ptr(py), _ptr(jx), _ptr(jy), _ptr(o), _ptr(d))
        return o, d

    def blackbody_table(self) -> np.ndarray:
        t = np.zeros((500, 3), np.float32)
        self.L.vpto_blackbody_table(_ptr(t))
        return t

    def blackbody(self, temps) -> np.ndarray:
        temps = np.ascontiguousarray(temps, np.float32)
        out = np.zeros((len(temps), 3), np.float32)
        self.L.vpto_blackbody(_ptr(temps), len(temps), _ptr(out))
        return out

    def hg_sample(self, w, u, g: float) -> np.ndarray:
        w, u = np.ascontiguousarray(w, np.float32), np.ascontiguousarray(u, np.float32)
        out = np.zeros_like(w)
        self.L.vpto_hg_sample(_ptr(w), _ptr(u), len(w), g, _ptr(out))
        return out

    def hg_eval(self, cos_theta, g: float) -> np.ndarray:
        c = np.ascontiguousarray(cos_theta, np.float32)
        out = np.zeros_like(c)
        self.L.vpto_hg_eval(_ptr(c), len(c), g, _ptr(out))
        return out

    def film_to_image(self, film: np.ndarray) -> np.ndarray:
        film = np.ascontiguousarray(film, np.float32)
        rgb = np.zeros(film.shape[:2] + (3,), np.uint8)
        self.L.vpto_film_to_image(_ptr(film), film.shape[1], film.shape[0], _ptr(rgb))
        return rgb

    def volume(self, grids) -> "OracleVolume":
        return OracleVolume(self, grids)


class OracleVolume:
    """The oracle's Volume (src/volume.cpp:162-170) over private copies of the grid buffers."""

    def __init__(self, oracle: Oracle, grids):
        self.o, self.L = oracle, oracle.L
        self.ctx = C.c_void_p()
        t = grids.temperature
        rc = self.L.vpto_create(C.byref(self.ctx), grids.density.ptr, grids.density.nbytes, t.ptr if t else None, t.nbytes if t else 0)
        if rc != 0:
            raise RuntimeError(f"vpto_create failed ({rc})")

    def __del__(self):
        if getattr(self, "ctx", None):
            self.L.vpto_destroy(self.ctx)
            self.ctx = None

    def leaf_maxima(self) -> np.ndarray:
        n = C.c_uint64()
        self.L.vpto_leaf_maxima(self.ctx, None, 0, C.byref(n))
        out = np.zeros(n.value, np.float32)
        self.L.vpto_leaf_maxima(self.ctx, _ptr(out), n.value, C.byref(n))
        return out

    def render(self, scene, film: np.ndarray, first_wave: int = 1, n_waves: int | None = None, threads: int | None = None):
        """-> (seconds, counters dict). Adds into `film` like the reference's workers do."""
        from volume_path_tracer_b200.api import Counters
        c, secs = Counters(), C.c_double()
        threads = threads or os.cpu_count() or 1
        rc = self.L.vpto_render(self.ctx, C.addressof(scene), first_wave, scene.num_waves if n_waves is None else n_waves, threads, _ptr(film),
                                C.addressof(c), C.byref(secs))
        if rc != 0:
            raise RuntimeError(f"vpto_render failed ({rc})")
        return secs.value, c

    def render_sample(self, scene, film: np.ndarray, first_wave: int, n_waves: int, job_stride: int, threads: int | None = None):
        """Timed CPU-baseline leg: every job_stride-th (tile, wave) job of the range. -> (seconds, counters)."""
        from volume_path_tracer_b200.api import Counters
        c, secs = Counters(), C.c_double()
        threads = threads or os.cpu_count() or 1
        rc = self.L.vpto_render_sample(self.ctx, C.addressof(scene), first_wave, n_waves, job_stride, threads, _ptr(film), C.addressof(c), C.byref(secs))
        if rc != 0:
            raise RuntimeError(f"vpto_render_sample failed ({rc})")
        return secs.value, c

    def trace_jobs(self, scene, jids):
        from volume_path_tracer_b200.api import Counters
        jids = np.ascontiguousarray(jids, np.uint64)
        slots = int(scene.tile_size[0]) * int(scene.tile_size[1])
        xyz, nd = np.zeros((len(jids), slots, 3), np.float32), np.zeros((len(jids), slots), np.uint32)
        c = Counters()
        rc = self.L.vpto_trace_jobs(self.ctx, C.addressof(scene), _ptr(jids), len(jids), _ptr(xyz), _ptr(nd), C.addressof(c))
        if rc != 0:
            raise RuntimeError(f"vpto_trace_jobs failed ({rc})")
        return xyz, nd, c

    def path_log(self, scene, jid: int, px: int, py: int, cap: int = 1 << 16) -> np.ndarray:
        """Logger<true> + single_pixel (worker.cpp:15-49): events of the path job `jid` traces through (px, py); rows of 8 floats."""
        ev, n = np.zeros((cap, 8), np.float32), C.c_size_t(0)
        rc = self.L.vpto_path_log(self.ctx, C.addressof(scene), jid, px, py, _ptr(ev), cap, C.byref(n))
        if rc != 0:
            raise RuntimeError(f"vpto_path_log failed ({rc})")
        return ev[: min(n.value, cap)].copy()

    def trace_segments(self, origins, dirs, max_segs: int = 256):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        segs, n = np.zeros((len(o), max_segs, 3), np.float32), np.zeros(len(o), np.uint32)
        rc = self.L.vpto_trace_segments(self.ctx, _ptr(o), _ptr(d), len(o), max_segs, _ptr(segs), _ptr(n))
        if rc != 0:
            raise RuntimeError(f"vpto_trace_segments failed ({rc})")
        return segs, n

    def sample_grid(self, grid: int, index_xyz):
        p = np.ascontiguousarray(index_xyz, np.float32).reshape(-1, 3)
        v, dm, mj = np.zeros(len(p), np.float32), np.zeros(len(p), np.uint32), np.zeros(len(p), np.float32)
        rc = self.L.vpto_sample_grid(self.ctx, grid, _ptr(p), len(p), _ptr(v), _ptr(dm), _ptr(mj))
        if rc != 0:
            raise RuntimeError(f"vpto_sample_grid failed ({rc})")
        return v, dm, mj


class Reference:
    """oracle/_ref/libvpt_reference.so: the reference's own translation units behind oracle/ref/reference_capi.cpp."""

    def __init__(self):
        if not REFSRC_PATH.exists():
            raise FileNotFoundError(f"{REFSRC_PATH} not built (needs the reference tree: make -C oracle ref)")
        L = self.L = C.CDLL(str(REFSRC_PATH))
        vp, sz = C.c_void_p, C.c_size_t
        L.vptr_about.restype = C.c_char_p
        L.vptr_create.argtypes = [C.POINTER(vp), vp, sz, vp, sz]
        L.vptr_destroy.argtypes = [vp]
        L.vptr_destroy.restype = None
        L.vptr_render.argtypes = [vp, vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, C.POINTER(C.c_double)]
        L.vptr_leaf_maxima.argtypes = [vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
        L.vptr_trace_segments.argtypes = [vp, vp, vp, sz, C.c_uint32, vp, vp]
        L.vptr_sample_collisions.argtypes = [vp, C.c_uint32, vp, vp, vp, sz, C.c_float, C.c_uint32, vp, vp]
        L.vptr_tile_rects.argtypes = [vp, vp, sz, C.POINTER(sz)]
        L.vptr_log_dda_trace.argtypes = [vp, vp, vp]
        L.vptr_rng.argtypes = [C.c_uint32, C.c_uint64, sz, C.POINTER(C.c_uint64), vp, vp]
        L.vptr_rng.restype = None
        L.vptr_camera_rays.argtypes = [vp, sz, vp, vp, vp, vp, vp, vp]
        L.vptr_camera_rays.restype = None
        L.vptr_blackbody.argtypes = [vp, sz, vp]
        L.vptr_blackbody.restype = None
        L.vptr_blackbody_table.argtypes = [vp]
        L.vptr_blackbody_table.restype = None
        L.vptr_hg_sample.argtypes = [vp, vp, sz, C.c_float, vp]
        L.vptr_hg_sample.restype = None
        L.vptr_hg_eval.argtypes = [vp, sz, C.c_float, vp]
        L.vptr_hg_eval.restype = None
        L.vptr_film_to_image.argtypes = [vp, C.c_int64, C.c_int64, vp]
        L.vptr_film_to_image.restype = None

    def about(self) -> str:
        return self.L.vptr_about().decode()

    def rng(self, seed: int, jid: int, n: int):
        s0 = C.c_uint64()
        u, f = np.zeros(n, np.uint32), np.zeros(n, np.float32)
        self.L.vptr_rng(seed, jid, n, C.byref(s0), _ptr(u), _ptr(f))
        return s0.value, u, f

    def camera_rays(self, scene, px, py, jx, jy):
        px, py = np.ascontiguousarray(px, np.int64), np.ascontiguousarray(py, np.int64)
        jx, jy = np.ascontiguousarray(jx, np.float32), np.ascontiguousarray(jy, np.float32)
        o, d = np.zeros((len(px), 3), np.float32), np.zeros((len(px), 3), np.float32)
        self.L.vptr_camera_rays(C.addressof(scene), len(px), _ptr(px), _ptr(py), _ptr(jx), _ptr(jy), _ptr(o), _ptr(d))
        return o, d

    def blackbody_table(self) -> np.ndarray:
        t = np.zeros((500, 3), np.float32)
        self.L.vptr_blackbody_table(_ptr(t))
        return t

    def blackbody(self, temps) -> np.ndarray:
        temps = np.ascontiguousarray(temps, np.float32)
        out = np.zeros((len(temps), 3), np.float32)
        self.L.vptr_blackbody(_ptr(temps), len(temps), _ptr(out))
        return out

    def hg_sample(self, w, u, g: float) -> np.ndarray:
        w, u = np.ascontiguousarray(w, np.float32), np.ascontiguousarray(u, np.float32)
        out = np.zeros_like(w)
        self.L.vptr_hg_sample(_ptr(w), _ptr(u), len(w), g, _ptr(out))
        return out

    def hg_eval(self, cos_theta, g: float) -> np.ndarray:
        c = np.ascontiguousarray(cos_theta, np.float32)
        out = np.zeros_like(c)
        self.L.vptr_hg_eval(_ptr(c), len(c), g, _ptr(out))
        return out

    def film_to_image(self, film: np.ndarray) -> np.ndarray:
        film = np.ascontiguousarray(film, np.float32)
        rgb = np.zeros(film.shape[:2] + (3,), np.uint8)
        self.L.vptr_film_to_image(_ptr(film), film.shape[1], film.shape[0], _ptr(rgb))
        return rgb

    def tile_rects(self, scene) -> np.ndarray:
        """TileProvider::next + token::compute_rect for wave 1 -> rows {x0, y0, w, h, jid} in hand-out order."""
        n = C.c_size_t(0)
        out = np.zeros((scene.ntiles, 5), np.int64)
        rc = self.L.vptr_tile_rects(C.addressof(scene), _ptr(out), scene.ntiles, C.byref(n))
        if rc != 0 or n.value != scene.ntiles:
            raise RuntimeError(f"vptr_tile_rects: rc {rc}, {n.value} tokens for {scene.ntiles} tiles")
        return out

    def volume(self, grids) -> "ReferenceVolume":
        return ReferenceVolume(self, grids)


class ReferenceVolume:
    """vpt::VolumeGrids + vpt::Volume of the compiled reference over private copies of the grid buffers."""

    def __init__(self, ref: Reference, grids):
        self.L = ref.L
        self.ctx = C.c_void_p()
        t = grids.temperature
        rc = self.L.vptr_create(C.byref(self.ctx), grids.density.ptr, grids.density.nbytes, t.ptr if t else None, t.nbytes if t else 0)
        if rc != 0:
            raise RuntimeError(f"vptr_create failed ({rc})")

    def __del__(self):
        if getattr(self, "ctx", None):
            self.L.vptr_destroy(self.ctx)
            self.ctx = None

    def render(self, scene, film: np.ndarray, first_wave: int = 1, n_waves: int | None = None, threads: int | None = None) -> float:
        """vpt::run on `threads` jthreads (main.cpp:63-87). Adds the reference's film into `film`; -> seconds."""
        secs = C.c_double()
        threads = threads or os.cpu_count() or 1
        rc = self.L.vptr_render(self.ctx, C.addressof(scene), first_wave, scene.num_waves if n_waves is None else n_waves, threads, _ptr(film), C.byref(secs))
        if rc != 0:
            raise RuntimeError(f"vptr_render failed ({rc})")
        return secs.value

    def leaf_maxima(self) -> np.ndarray:
        n = C.c_uint64()
        self.L.vptr_leaf_maxima(self.ctx, None, 0, C.byref(n))
        out = np.zeros(n.value, np.float32)
        self.L.vptr_leaf_maxima(self.ctx, _ptr(out), n.value, C.byref(n))
        return out

    def trace_segments(self, origins, dirs, max_segs: int = 256):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        segs, n = np.zeros((len(o), max_segs, 3), np.float32), np.zeros(len(o), np.uint32)
        rc = self.L.vptr_trace_segments(self.ctx, _ptr(o), _ptr(d), len(o), max_segs, _ptr(segs), _ptr(n))
        if rc != 0:
            raise RuntimeError(f"vptr_trace_segments failed ({rc})")
        return segs, n

    def log_dda_trace(self, origin, direction, workdir) -> str:
        """Volume::log_dda_trace: the reference writes ./dda_trace.csv; run it with `workdir` as the current directory. -> the file's text ('' if none)."""
        o, d = np.ascontiguousarray(origin, np.float32), np.ascontiguousarray(direction, np.float32)
        cwd = os.getcwd()
        os.chdir(workdir)
        try:
            p = Path("dda_trace.csv")
            if p.exists():
                p.unlink()
            if self.L.vptr_log_dda_trace(self.ctx, _ptr(o), _ptr(d)) != 0:
                raise RuntimeError("vptr_log_dda_trace failed")
            return p.read_text() if p.exists() else ""
        finally:
            os.chdir(cwd)

    def sample_collisions(self, seed: int, jids, origins, dirs, sigma_t: float, max_hits: int = 64):
        """MajorantTransmittanceSampler::next along world rays -> (hits[n, max_hits, 5] = point xyz, sigma_maj, density; n_hits)."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        j = np.ascontiguousarray(jids, np.uint64)
        hits, n = np.zeros((len(o), max_hits, 5), np.float32), np.zeros(len(o), np.uint32)
        rc = self.L.vptr_sample_collisions(self.ctx, seed, _ptr(j), _ptr(o), _ptr(d), len(o), sigma_t, max_hits, _ptr(hits), _ptr(n))
        if rc != 0:
            raise RuntimeError(f"vptr_sample_collisions failed ({rc})")
        return hits, n
