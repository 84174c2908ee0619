"""The device's logf / sinf / cosf (csrc/device/libm_exact.cuh) round exactly like this machine's glibc on the
integrator's argument domains — checked on the host build of the same header. The exhaustive sweep (every float in
[2^-126, 1] and in [0, 2*pi], 2.1e9 values, zero mismatches on glibc 2.39) runs with VPT_LIBM_EXHAUSTIVE=1; the
default run sweeps every 61st float (35 M values)."""
import os


def test_libm_exact_matches_glibc(built):
    from oracle import oracle_lib
    stride = 1 if os.environ.get("VPT_LIBM_EXHAUSTIVE") == "1" else 61
    n_log, bad_log, n_sc, bad_sc = oracle_lib.libm_check(stride)
    assert n_log > 1e7 and n_sc > 1e7
    assert bad_log == 0 and bad_sc == 0


def test_expf_powf_exact_match_glibc(built):
    """expf (planck_law, spectral.cpp:16) and powf(x, 1/2.4f) (linsrgb_to_srgb, color.hpp:22) of libm_exact.cuh against this machine's
    glibc: exhaustive (2.2e9 + 1.1e9 values) with VPT_LIBM_EXHAUSTIVE=1, else every 61st float. Zero mismatches on glibc 2.39 / FMA CPUs."""
    from oracle import oracle_lib
    stride = 1 if os.environ.get("VPT_LIBM_EXHAUSTIVE") == "1" else 61
    n_exp, bad_exp, n_pow, bad_pow = oracle_lib.libm_check2(stride)
    assert n_exp > 1e7 and n_pow > 1e7
    assert bad_exp == 0 and bad_pow == 0
