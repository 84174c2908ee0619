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
#include <cmath>
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

// ---------------------------------------------------------------- expf / powf (glibc e_expf.c, e_powf.c: the same ARM "optimized routines")
// Needed where the reference calls them on the HOST side of the hot path and the device must produce the same bytes:
//   * powf(x, 1/2.4f) in linsrgb_to_srgb (include/vpt/color.hpp:18-30) -> vpt_resolve_kernel (film_to_image on the device);
//   * expf in planck_law (src/spectral.cpp:16) for the direct 471-term blackbody integration at T >= 49 900 K
//     (src/precompute_blackbody.cpp:34-36), reachable from the render kernel through temperature_scale.
// Tables: glibc 2.39's __exp2f_data.tab (2^(i/32) bit patterns minus i<<47) and __powf_log2_data.tab ({1/c, log2 c} x 16), read out of
// this image's libm.so.6 (tools/extract_libm_tables.py prints them). On x86-64 glibc dispatches these routines to builds compiled with
// -mfma (sysdeps/x86_64/fpu/multiarch/e_expf-fma.c, e_powf-fma.c) on every CPU that has FMA — the GPU box's host included — so the
// double-precision a*b+c steps below are fused exactly where that build fuses them. Which steps those are was settled empirically, by
// exhaustive comparison with this machine's glibc (oracle/libm_check.cpp): expf with every step fused, INCLUDING kd = fma(InvLn2N, x, Shift)
// and r = fma(InvLn2N, x, -kd): 0 mismatches over every float in [2^-126, 88) (1 mismatch if kd is not fused); powf(x, 1/2.4f) with the
// polynomial steps fused and y*log2(x) NOT fused into the shift: 0 mismatches over every float in (0.0031308, FLT_MAX] (8 if it is fused).
#define VPT_EXP2F_TABLE_INIT                                                                                          \
    {                                                                                                                 \
        0x3ff0000000000000ULL, 0x3fefd9b0d3158574ULL, 0x3fefb5586cf9890fULL, 0x3fef9301d0125b51ULL,                    \
        0x3fef72b83c7d517bULL, 0x3fef54873168b9aaULL, 0x3fef387a6e756238ULL, 0x3fef1e9df51fdee1ULL,                    \
        0x3fef06fe0a31b715ULL, 0x3feef1a7373aa9cbULL, 0x3feedea64c123422ULL, 0x3feece086061892dULL,                    \
        0x3feebfdad5362a27ULL, 0x3feeb42b569d4f82ULL, 0x3feeab07dd485429ULL, 0x3feea47eb03a5585ULL,                    \
        0x3feea09e667f3bcdULL, 0x3fee9f75e8ec5f74ULL, 0x3feea11473eb0187ULL, 0x3feea589994cce13ULL,                    \
        0x3feeace5422aa0dbULL, 0x3feeb737b0cdc5e5ULL, 0x3feec49182a3f090ULL, 0x3feed503b23e255dULL,                    \
        0x3feee89f995ad3adULL, 0x3feeff76f2fb5e47ULL, 0x3fef199bdd85529cULL, 0x3fef3720dcef9069ULL,                    \
        0x3fef5818dcfba487ULL, 0x3fef7c97337b9b5fULL, 0x3fefa4afa2a490daULL, 0x3fefd0765b6e4540ULL                     \
    }
#define VPT_POWF_LOG2_TABLE_INIT                                                                                      \
    {                                                                                                                 \
        0x1.661ec79f8f3bep+0, -0x1.efec65b963019p-2, 0x1.571ed4aaf883dp+0, -0x1.b0b6832d4fca4p-2,                      \
        0x1.49539f0f010b0p+0, -0x1.7418b0a1fb77bp-2, 0x1.3c995b0b80385p+0, -0x1.39de91a6dcf7bp-2,                      \
        0x1.30d190c8864a5p+0, -0x1.01d9bf3f2b631p-2, 0x1.25e227b0b8ea0p+0, -0x1.97c1d1b3b7af0p-3,                      \
        0x1.1bb4a4a1a343fp+0, -0x1.2f9e393af3c9fp-3, 0x1.12358f08ae5bap+0, -0x1.960cbbf788d5cp-4,                      \
        0x1.0953f419900a7p+0, -0x1.a6f9db6475fcep-5, 0x1.0000000000000p+0, 0x0.0p+0,                                   \
        0x1.e608cfd9a47acp-1, 0x1.338ca9f24f53dp-4, 0x1.ca4b31f026aa0p-1, 0x1.476a9543891bap-3,                        \
        0x1.b2036576afce6p-1, 0x1.e840b4ac4e4d2p-3, 0x1.9c2d163a1aa2dp-1, 0x1.40645f0c6651cp-2,                        \
        0x1.886e6037841edp-1, 0x1.88e9c2c1b9ff8p-2, 0x1.767dcf5534862p-1, 0x1.ce0a44eb17bccp-2                         \
    }

VPT_HD uint64_t as_u64(double d)
{
#if defined(__CUDA_ARCH__)
    return (uint64_t)__double_as_longlong(d);
#else
    uint64_t u; std::memcpy(&u, &d, 8); return u;
#endif
}
VPT_HD double as_f64(uint64_t u)
{
#if defined(__CUDA_ARCH__)
    return __longlong_as_double((long long)u);
#else
    double d; std::memcpy(&d, &u, 8); return d;
#endif
}

// expf(x) for finite |x| < 88 (no overflow / underflow handling: callers stay inside). `tab` = VPT_EXP2F_TABLE_INIT.
VPT_HD float expf_exact(float x, const uint64_t* tab)
{
    const double xd = double(x);
    const double InvLn2N = 0x1.71547652b82fep+5, Shift = 0x1.8p+52; // N = 32
    double kd = fma(InvLn2N, xd, Shift);
    const uint64_t ki = as_u64(kd);
    kd -= Shift;
    const double r = fma(InvLn2N, xd, -kd);
    uint64_t t = tab[ki & 31u];
    t += ki << 47;
    const double s = as_f64(t);
    const double z = fma(0x1.c6af84b912394p-20, r, 0x1.ebfce50fac4f3p-13); // poly_scaled
    const double r2 = r * r;
    double y = fma(0x1.62e42ff0c52d6p-6, r, 1.0);
    y = fma(z, r2, y);
    return float(y * s);
}

// powf(x, y) for positive NORMAL finite x and |y * log2(x)| < 126 (no special cases). `e2` = VPT_EXP2F_TABLE_INIT, `l2` = VPT_POWF_LOG2_TABLE_INIT.
VPT_HD float powf_exact(float x, float yf, const uint64_t* e2, const double* l2)
{
    const uint32_t ix = as_u32(x);
    const uint32_t tmp = ix - 0x3f330000u;
    const int i = int((tmp >> 19) & 15u);
    const uint32_t top = tmp & 0xff800000u;
    const int k = int32_t(top) >> 23;
    const double invc = l2[2 * i], logc = l2[2 * i + 1];
    const double z = double(as_f32(ix - top));
    const double r = fma(z, invc, -1.0);
    const double y0 = logc + double(k);
    const double r2 = r * r;
    double y = fma(0x1.27616c9496e0bp-2, r, -0x1.71969a075c67ap-2);
    const double p = fma(0x1.ec70a6ca7baddp-2, r, -0x1.7154748bef6c8p-1);
    const double r4 = r2 * r2;
    double q = fma(0x1.71547652ab82bp0, r, y0);
    q = fma(p, r2, q);
    y = fma(y, r4, q); // log2(x)
    const double xd = double(yf) * y;
    const double Shift = 0x1.8p+47; // 0x1.8p52 / 32
    double kd = xd + Shift;
    const uint64_t ki = as_u64(kd);
    kd -= Shift;
    const double rr = xd - kd;
    uint64_t t = e2[ki & 31u];
    t += ki << 47;
    const double s = as_f64(t);
    const double zz = fma(0x1.c6af84b912394p-5, rr, 0x1.ebfce50fac4f3p-3);
    const double rr2 = rr * rr;
    double v = fma(0x1.62e42ff0c52d6p-1, rr, 1.0);
    v = fma(zz, rr2, v);
    return float(v * s);
}

} // namespace vptm
