"""vptc_film_to_image_static (include/vptc.h): the resolve kernel with its device staging in a file-scope static __device__ array,
for callers that may not call cudaMalloc. No counterpart in the reference, so the
check is against the existing entry point, vptc_film_to_image, whose output tests/test_gpu_parity.py pins byte-exactly to the
reference's film_to_image: RGB8 is integer work, the tolerance is zero.

Collected last on purpose. Passes alone in a fresh process and after the other tests: it makes
its own context, and nothing else uses the static array.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _film(h, w, seed, waves=7.0):
    rs = np.random.default_rng(seed)
    f = np.empty((h, w, 4), np.float32)
    # XYZ sums covering black, mid-tones (both OETF branches), out-of-gamut negatives after the matrix and > 1 values that clamp
    f[..., :3] = (rs.random((h, w, 3)) ** 3 * 1.6 * waves).astype(np.float32)
    f[..., 3] = np.float32(waves)
    f[0, : w // 2, :3] = 0.0
    f[-1, :, 3] = 0.0  # unsampled pixels: 0/0, the NaN path of film_to_image
    return f


def test_static_resolve_equals_allocating_resolve_twice(api):
    L = api.lib()
    r = api.Renderer(0)
    try:
        # second call is SMALLER and has other contents: stale bytes of the first call stay in the array beyond it and must not leak in
        for (h, w, seed) in ((270, 480, 1), (100, 70, 2)):
            film = _film(h, w, seed)
            want = r.film_to_image(film)
            got = np.full((h, w, 3), 0x5A, np.uint8)
            rc = L.vptc_film_to_image_static(r._ctx, film.ctypes.data_as(C.c_void_p), w, h, got.ctypes.data_as(C.c_void_p))
            assert rc == 0, L.vptc_last_error(r._ctx)
            assert np.array_equal(got, want)
            assert want.max() == 255 and want.min() == 0 and len(np.unique(want)) > 200  # a real image, not a constant
        # the largest frame the scratch takes (1920 x 1080 fits under 2^21 pixels), then the first frame again
        film = _film(1080, 1920, 3)
        assert np.array_equal(r.film_to_image_static(film), r.film_to_image(film))
        film = _film(270, 480, 1)
        assert np.array_equal(r.film_to_image_static(film), r.film_to_image(film))
    finally:
        r.close()


def test_static_resolve_rejects_oversize_and_null(api):
    L = api.lib()
    r = api.Renderer(0)
    try:
        w, h = 2048, 1025  # 2^21 + 2048 pixels
        assert w * h > api.STATIC_IMAGE_MAX_PIXELS
        film = np.zeros((h, w, 4), np.float32)
        rgb = np.zeros((h, w, 3), np.uint8)
        rc = L.vptc_film_to_image_static(r._ctx, film.ctypes.data_as(C.c_void_p), w, h, rgb.ctypes.data_as(C.c_void_p))
        assert rc != 0
        msg = L.vptc_last_error(r._ctx).decode()
        assert "VPTC_STATIC_IMAGE_MAX_PIXELS" in msg and str(api.STATIC_IMAGE_MAX_PIXELS) in msg
        assert L.vptc_film_to_image_static(r._ctx, None, 4, 4, rgb.ctypes.data_as(C.c_void_p)) != 0
        assert L.vptc_film_to_image_static(r._ctx, film.ctypes.data_as(C.c_void_p), 0, 4, rgb.ctypes.data_as(C.c_void_p)) != 0
        # exactly the maximum is accepted
        w, h = 2048, 1024
        film = _film(h, w, 4)
        assert np.array_equal(r.film_to_image_static(film), r.film_to_image(film))
    finally:
        r.close()
