// libm_exact.cuh — logf / sinf / cosf that round exactly like the host's glibc (2.28+) on the arguments the
// integrator feeds them, so that delta-tracking branch decisions on the device match the CPU reference bit for
// bit instead of "within 1-2 ulp" (any ulp difference in -logf(1-u) can flip t < seg.t1
// and shift every later RNG draw of the job).
//
// glibc's float routines (sysdeps/ieee754/flt-32/e_logf.c, s_sincosf.h — the ARM "optimized routines"
// algorithms) evaluate a short table + polynomial scheme in DOUBLE and round once to float. The same scheme is
// restated here, in the same operation order, with IEEE double arithmetic (compile with --fmad=false; the
// FMA-contracted and the plain variants were both checked). Verified EXHAUSTIVELY against glibc 2.39 on the
// CPU (oracle/libm_check.cpp: every float in [2^-126, 1] for logf — 1 056 964 609 values — and every float in
// [0, 2*pi] for sinf/cosf — 1 086 918 621 values — zero mismatches); tests/test_libm_exact.py re-checks a
// strided sweep on every run. Domains: logf argument 1-u in [2^-24, 1] (include/vpt/random.hpp:19-22);
// sinf/cosf argument phi = 2*pi*u in [0, 2*pi) (random.hpp:68-71). Outside [0, 120) / subnormals the routines
// below are NOT valid and the callers never get there.
#pragma once
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define VPT_HD __host__ __device__ __forceinline__
#else
#define VPT_HD inline
#endif

namespace vptm {

VPT_HD uint32_t as_u32(float f)
{
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    uint32_t u; std::memcpy(&u, &f, 4); return u;
#endif
}
VPT_HD float as_f32(uint32_t u)
{
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    float f; std::memcpy(&f, &u, 4); return f;
#endif
}

// 16 x {1/c, log(c)} for the sub-intervals of [0x3f330000, 2*0x3f330000), as double bit patterns.
struct LogfTable { double invc_logc[32]; };
#define VPT_LOGF_TABLE_INIT                                                                                          \
    {                                                                                                                \
        0x1.661ec79f8f3bep+0, -0x1.57bf7808caadep-2, 0x1.571ed4aaf883dp+0, -0x1.2bef0a7c06ddbp-2,                    \
        0x1.49539f0f010bp+0, -0x1.01eae7f513a67p-2, 0x1.3c995b0b80385p+0, -0x1.b31d8a68224e9p-3,                     \
        0x1.30d190c8864a5p+0, -0x1.6574f0ac07758p-3, 0x1.25e227b0b8eap+0, -0x1.1aa2bc79c81p-3,                       \
        0x1.1bb4a4a1a343fp+0, -0x1.a4e76ce8c0e5ep-4, 0x1.12358f08ae5bap+0, -0x1.1973c5a611cccp-4,                    \
        0x1.0953f419900a7p+0, -0x1.252f438e10c1ep-5, 0x1p+0, 0x0p+0, 0x1.e608cfd9a47acp-1, 0x1.aa5aa5df25984p-5,     \
        0x1.ca4b31f026aap-1, 0x1.c5e53aa362eb4p-4, 0x1.b2036576afce6p-1, 0x1.526e57720db08p-3,                       \
        0x1.9c2d163a1aa2dp-1, 0x1.bc2860d22477p-3, 0x1.886e6037841edp-1, 0x1.1058bc8a07ee1p-2,                       \
        0x1.767dcf5534862p-1, 0x1.4043057b6ee09p-2                                                                   \
    }

// logf(x) for normal x in [2^-126, 1]. `tab` = the 32 doubles above (shared memory on the device).
VPT_HD float logf_exact(float x, const double* tab)
{
    const uint32_t ix = as_u32(x);
    if (ix == 0x3f800000u) return 0.0f;
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = int((tmp >> 19) & 15u);
    const int k = int32_t(tmp) >> 23;
    const uint32_t iz = ix - (tmp & 0xff800000u);
    const double invc = tab[2 * i], logc = tab[2 * i + 1];
    const double z = double(as_f32(iz));
    const double r = z * invc - 1.0;
    const double y0 = logc + double(k) * 0x1.62e42fefa39efp-1;
    const double r2 = r * r;
    double y = 0x1.5575b0be00b6ap-2 * r + -0x1.ffffef20a4123p-2;
    y = -0x1.00ea348b88334p-2 * r2 + y;
    y = y * r2 + (y0 + r);
    return float(y);
}

// Shared argument reduction + polynomials of glibc's sinf/cosf for |y| < 120.
VPT_HD float sincos_poly(double x, double x2, bool negate_cos, int n)
{
    if ((n & 1) == 0) { // sine polynomial
        const double x3 = x * x2;
        const double s1 = 0x1.1107605230bc4p-7 + x2 * -0x1.994eb3774cf24p-13;
        const double x7 = x3 * x2;
        const double s = x + x3 * -0x1.555545995a603p-3;
        return float(s + x7 * s1);
    }
    // table[1] = table[0] with the cosine coefficients negated; round-to-nearest is sign-symmetric, so the
    // negated polynomial is exactly the negated result.
    const double x4 = x2 * x2;
    const double c2 = -0x1.6c087e89a359dp-10 + x2 * 0x1.99343027bf8c3p-16;
    const double c1 = 0x1p0 + x2 * -0x1.ffffffd0c621cp-2;
    const double x6 = x4 * x2;
    const double c = c1 + x4 * 0x1.55553e1068f19p-5;
    const float res = float(c + x6 * c2);
    return negate_cos ? -res : res;
}
VPT_HD uint32_t abstop12(float x) { return (as_u32(x) >> 20) & 0x7ffu; }

// sinf(y) and cosf(y) together, 0 <= y < 120 (one shared reduction).
VPT_HD void sincosf_exact(float y, float& s_out, float& c_out)
{
    double x = double(y);
    // Branch-free form for SIMT: glibc's |y| < pi/4 shortcut is the general path with n = 0 (x - 0 * pi/2 == x, sign +1,
    // table 0), and sincosf_poly evaluates BOTH polynomials and swaps them when n is odd — so one straight-line sequence
    // serves every quadrant; only the |y| < 2^-12 early-out stays, as a select.
    const double r = x * 0x1.45F306DC9C883p+23; // 2/pi * 2^24
    const int n = abstop12(y) < abstop12(0x1.921FB6p-1f) ? 0 : (int32_t(r) + 0x800000) >> 24;
    x = x - double(n) * 0x1.921FB54442D18p0;
    const double sgn = ((n + 1) & 2) ? -1.0 : 1.0; // sign[n & 3] = {1, -1, -1, 1}
    const bool neg = (n & 2) != 0;
    const double xs = x * sgn, x2 = x * x;
    const float sp = sincos_poly(xs, x2, neg, 0), cp = sincos_poly(xs, x2, neg, 1);
    const bool odd = (n & 1) != 0, tiny = abstop12(y) < abstop12(0x1p-12f);
    s_out = tiny ? y : (odd ? cp : sp);
    c_out = tiny ? 1.0f : (odd ? sp : cp);
}

} // namespace vptm
