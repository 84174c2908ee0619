// oracle/libm_check.cpp — TEST INFRASTRUCTURE. Sweeps the device's exact-rounding logf / sinf / cosf / expf / powf restatement
// (volume_path_tracer_b200/csrc/device/libm_exact.cuh, compiled here for the host) against THIS machine's glibc.
//   g++ -O2 -ffp-contract=off -shared -fPIC -pthread oracle/libm_check.cpp -o oracle/libvpt_libm_check.so
// vptm_check(stride)  : logf on [2^-126, 1], sinf/cosf on [0, 2*pi]
// vptm_check2(stride) : expf on [2^-126, 88) and its negatives, powf(x, 1/2.4f) on (0.0031308, FLT_MAX] (the sRGB OETF's whole domain)
// stride 1 = exhaustive (every float of the domains); both return mismatch counts. 8 threads.
#include <atomic>
#include <cmath>
#include <cstdint>
#include <thread>
#include <vector>

#include "../volume_path_tracer_b200/csrc/device/libm_exact.cuh"

namespace {
template <class F>
void sweep(uint64_t lo, uint64_t hi_inclusive, uint32_t stride, F body)
{
    const unsigned nt = 8;
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; ++t)
        th.emplace_back([=] { for (uint64_t u = lo + uint64_t(t) * stride; u <= hi_inclusive; u += uint64_t(nt) * stride) body(uint32_t(u)); });
    for (auto& x : th) x.join();
}
} // namespace

extern "C" void vptm_check(uint32_t stride, uint64_t out[4])
{
    static const double tab[32] = VPT_LOGF_TABLE_INIT;
    std::atomic<uint64_t> n_log{0}, bad_log{0}, n_sc{0}, bad_sc{0};
    sweep(0x00800000u, 0x3f800000u, stride, [&](uint32_t u) { // normal floats in [2^-126, 1]
        const float x = vptm::as_f32(u);
        n_log.fetch_add(1, std::memory_order_relaxed);
        if (vptm::as_u32(std::log(x)) != vptm::as_u32(vptm::logf_exact(x, tab))) ++bad_log;
    });
    sweep(0, vptm::as_u32(6.2831860f), stride, [&](uint32_t u) { // [0, 2*pi]
        const float x = vptm::as_f32(u);
        float s, c;
        vptm::sincosf_exact(x, s, c);
        n_sc.fetch_add(1, std::memory_order_relaxed);
        if (vptm::as_u32(std::sin(x)) != vptm::as_u32(s) || vptm::as_u32(std::cos(x)) != vptm::as_u32(c)) ++bad_sc;
    });
    out[0] = n_log; out[1] = bad_log; out[2] = n_sc; out[3] = bad_sc;
}

extern "C" void vptm_check2(uint32_t stride, uint64_t out[4])
{
    static const uint64_t e2[32] = VPT_EXP2F_TABLE_INIT;
    static const double l2[32] = VPT_POWF_LOG2_TABLE_INIT;
    std::atomic<uint64_t> n_exp{0}, bad_exp{0}, n_pow{0}, bad_pow{0};
    sweep(0x00800000u, vptm::as_u32(88.0f) - 1, stride, [&](uint32_t u) { // +-[2^-126, 88)
        const float x = vptm::as_f32(u);
        n_exp.fetch_add(2, std::memory_order_relaxed);
        if (vptm::as_u32(std::exp(x)) != vptm::as_u32(vptm::expf_exact(x, e2))) ++bad_exp;
        if (x < 87.0f && vptm::as_u32(std::exp(-x)) != vptm::as_u32(vptm::expf_exact(-x, e2))) ++bad_exp; // below -87.3 the result is subnormal: outside the contract
    });
    const float y = 1.0f / 2.4f;
    sweep(vptm::as_u32(0.0031308f), 0x7f7fffffu, stride, [&](uint32_t u) {
        const float x = vptm::as_f32(u);
        n_pow.fetch_add(1, std::memory_order_relaxed);
        if (vptm::as_u32(std::pow(x, y)) != vptm::as_u32(vptm::powf_exact(x, y, e2, l2))) ++bad_pow;
    });
    out[0] = n_exp; out[1] = bad_exp; out[2] = n_pow; out[3] = bad_pow;
}
