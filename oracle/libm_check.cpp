// oracle/libm_check.cpp — TEST INFRASTRUCTURE. Sweeps the device's exact-rounding logf / sinf / cosf restatement
// (volume_path_tracer_b200/csrc/device/libm_exact.cuh, compiled here for the host) against THIS machine's glibc.
//   g++ -O2 -ffp-contract=off -shared -fPIC oracle/libm_check.cpp -o oracle/libvpt_libm_check.so
// vptm_check(stride): stride 1 = exhaustive (every float of the domains, ~40 s); returns the mismatch counts.
#include <cmath>
#include <cstdint>

#include "../volume_path_tracer_b200/csrc/device/libm_exact.cuh"

extern "C" void vptm_check(uint32_t stride, uint64_t out[4])
{
    static const double tab[32] = VPT_LOGF_TABLE_INIT;
    uint64_t n_log = 0, bad_log = 0, n_sc = 0, bad_sc = 0;
    for (uint64_t u = 0x00800000u; u <= 0x3f800000u; u += stride) { // normal floats in [2^-126, 1]
        const float x = vptm::as_f32(uint32_t(u));
        ++n_log;
        if (vptm::as_u32(std::log(x)) != vptm::as_u32(vptm::logf_exact(x, tab))) ++bad_log;
    }
    const uint32_t top = vptm::as_u32(6.2831860f);
    for (uint64_t u = 0; u <= top; u += stride) { // [0, 2*pi]
        const float x = vptm::as_f32(uint32_t(u));
        float s, c;
        vptm::sincosf_exact(x, s, c);
        ++n_sc;
        if (vptm::as_u32(std::sin(x)) != vptm::as_u32(s) || vptm::as_u32(std::cos(x)) != vptm::as_u32(c)) ++bad_sc;
    }
    out[0] = n_log; out[1] = bad_log; out[2] = n_sc; out[3] = bad_sc;
}
