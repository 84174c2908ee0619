// oracle/vpt_restated.hpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may build or call it.
//
// Statement-by-statement CPU restatement of the reference's hot path with Eigen replaced by vec3.hpp,
// templated on a substrate S that supplies the NanoVDB / pcg-cpp arithmetic:
//   * substrate_own.hpp      — our restatement of that arithmetic (builds anywhere)
//   * ref/substrate_nanovdb.hpp — the reference's vendored headers themselves (`oracle/_ref` build)
// Each block cites the reference file:line it follows. Compile with -ffp-contract=off: the reference's
// Release build targets baseline x86-64 (no FMA), so only NanoVDB's explicit fmaf calls fuse.
//
// Pinning: the reference has no tests or golden outputs; this restatement is pinned by
// running it on the reference's own vendored arithmetic (oracle/_ref) and comparing with the
// self-contained build bit for bit (tests/test_oracle_ref.py), plus NanoVDB unit-test known answers — and, since round 2, by THE
// REFERENCE'S OWN src/*.cpp compiled unchanged into oracle/_ref/libvpt_reference.so (oracle/Makefile `ref`, oracle/ref/reference_capi.cpp):
// tests/test_reference_compiled.py asserts this restatement equals that library bit for bit (films of vpt::run, segments, collisions,
// patched maxima, camera, tiles, RNG, blackbody, HG, film_to_image, dda_trace.csv). Parity is pinned; the one thing no build on this
// machine can check is Eigen's own association order in 3-term reductions (Eigen is absent: the stand-in header and vec3.hpp use a + (b + c)).
#pragma once
#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numbers>
#include <numeric>
#include <optional>
#include <thread>
#include <vector>

#include "vec3.hpp"
#include "vptc.h"

namespace vpto {

struct Counters {
    uint64_t pixel_samples = 0, density_evals = 0, temperature_evals = 0, majorant_fetches = 0, rng_draws = 0,
             scatter_events = 0, shadow_rays = 0;
    // Logger<true> of src/worker.cpp:15-49 (the reference ships Logger<false>, :102): when set, every camera-path event is appended as
    // 8 floats {type, 6 arguments, 0}: 0 new_ray(o, d)  1 sampled_point(point, density)  2 null  3 scatter(o, d)  4 absorbed  5 scatter_terminated
    std::vector<float>* log = nullptr;
    void event(float type, float a = 0, float b = 0, float c = 0, float d = 0, float e = 0, float f = 0)
    {
        if (log) log->insert(log->end(), {type, a, b, c, d, e, f, 0.0f});
    }
    void add(const Counters& o)
    {
        pixel_samples += o.pixel_samples; density_evals += o.density_evals; temperature_evals += o.temperature_evals;
        majorant_fetches += o.majorant_fetches; rng_draws += o.rng_draws; scatter_events += o.scatter_events; shadow_rays += o.shadow_rays;
    }
};

// ---------------------------------------------------------------- include/vpt/hash.hpp:20-67
inline uint64_t murmur64a_padded(const void* key, int len, uint64_t seed)
{
    constexpr uint64_t m = 0xc6a4a7935bd1e995ULL;
    constexpr int r = 47;
    uint64_t h = seed ^ (uint64_t(len) * m);
    const uint64_t* data = static_cast<const uint64_t*>(key);
    const uint64_t* end = data + (len / 8);
    while (data != end) {
        uint64_t k = *data++;
        k *= m; k ^= k >> r; k *= m;
        h ^= k; h *= m;
    }
    h ^= h >> r; h *= m; h ^= h >> r;
    return h;
}
inline uint64_t hash_jid(uint64_t seed, uint64_t jid) // hash(m_seed, jid), one size_t argument
{
    uint64_t key[1] = {jid};
    return murmur64a_padded(key, 8, seed);
}

// ---------------------------------------------------------------- include/vpt/random.hpp:86-115
template <class Engine>
struct Rng {
    explicit Rng(uint32_t seed) : m_seed(seed) {}
    void begin_job(uint64_t jid) { m_engine.seed(hash_jid(m_seed, jid)); }
    uint32_t uniform_u32() { ++draws; return m_engine(); }
    float uniform()
    {
        constexpr float ONE_MINUS_EPS = float(0x1.fffffep-1);
        return std::min<float>(ONE_MINUS_EPS, uniform_u32() * 0x1p-32f);
    }
    Engine m_engine;
    uint32_t m_seed;
    uint64_t draws = 0; // instrumentation only
};

// ---------------------------------------------------------------- include/vpt/random.hpp:19-47
inline float sample_exponential(float u, float a) { return -std::log(1 - u) / a; }

enum class ScatterEvent { Null, Absorption, Scatter };
struct Choice { ScatterEvent value; float weight; };
inline ScatterEvent sample_discrete(std::initializer_list<Choice> choices, float u)
{
    float total_w = std::accumulate(choices.begin(), choices.end(), 0.0f, [](float acc, Choice x) { return acc + x.weight; });
    u = u * total_w;
    auto end = choices.end();
    for (auto it = choices.begin(); it != end; ++it) {
        u -= it->weight;
        if (u <= 0) return it->value;
    }
    return std::prev(end)->value;
}

// ---------------------------------------------------------------- include/vpt/utils.hpp:39-66
inline void coordinate_system(V3 v1, V3& v2, V3& v3)
{
    float sign = std::copysign(1.0f, v1.z);
    float a = -1.0f / (sign + v1.z);
    float b = v1.x * v1.y * a;
    v2.x = 1.0f + sign * a * std::pow(v1.x, 2.0f);
    v2.y = sign * b;
    v2.z = -sign * v1.x;
    v3.x = b;
    v3.y = sign + a * std::pow(v1.y, 2.0f);
    v3.z = -v1.y;
}
inline float henyey_greenstein(float cos_theta, float g)
{
    float den = 1.0f + g * g + 2.0f * g * cos_theta;
    constexpr float inv_4_pi = static_cast<float>(std::numbers::inv_pi / 4.0);
    return inv_4_pi * (1.0f - g * g) / (den * std::sqrt(std::max(0.0f, den)));
}

// ---------------------------------------------------------------- include/vpt/random.hpp:56-84
inline V3 sample_henyey_greenstein(const V3& w, float ux, float uy, float g)
{
    float g2 = std::pow(g, 2.0f);
    float cos_theta;
    if (std::abs(g) < 1e-3f)
        cos_theta = 1 - 2 * ux;
    else
        cos_theta = 1.0f / (2.0f * g) * (1.0f + g2 - std::pow((1.0f - g2) / (1.0f + g - 2.0f * g * ux), 2.0f));
    float sin_theta = std::sqrt(std::max(0.0f, 1.0f - std::pow(cos_theta, 2.0f)));
    float phi = 2.0f * static_cast<float>(std::numbers::pi) * uy;
    V3 local(std::clamp(sin_theta, -1.0f, 1.0f) * std::cos(phi), std::clamp(sin_theta, -1.0f, 1.0f) * std::sin(phi),
             std::clamp(cos_theta, -1.0f, 1.0f));
    local = normalized(local);
    V3 x, y;
    V3 z = w;
    coordinate_system(z, x, y);
    return local.x * x + local.y * y + local.z * z;
}

// ---------------------------------------------------------------- src/camera.cpp:5-57, include/vpt/camera.hpp:14-23
struct Camera {
    Camera(const vptc_camera_parameters& p, int64_t W, int64_t H) : params(p)
    {
        float ar = static_cast<float>(W) / static_cast<float>(H);
        float vfov_rad = float(std::numbers::pi) * p.vfov_deg / 180.0f;
        // camera_to_world
        V3 pos(p.position[0], p.position[1], p.position[2]), look(p.look[0], p.look[1], p.look[2]), up(p.up[0], p.up[1], p.up[2]);
        V3 dir = normalized(look - pos);
        V3 left = cross(normalized(up), dir);
        V3 new_up = cross(dir, left);
        Affine c2w; c2w.set_identity();
        const V3 cols[3] = {left, new_up, dir};
        for (int c = 0; c < 3; ++c) { c2w.m[0][c] = cols[c].x; c2w.m[1][c] = cols[c].y; c2w.m[2][c] = cols[c].z; }
        c2w.t[0] = pos.x; c2w.t[1] = pos.y; c2w.t[2] = pos.z;
        // screen_to_camera
        float tan_vfov2 = std::tan(vfov_rad / 2);
        Affine s2c; s2c.set_identity();
        s2c.m[0][0] = ar * tan_vfov2; s2c.m[1][1] = tan_vfov2; s2c.m[2][2] = 0.0f;
        s2c.t[0] = 0.0f; s2c.t[1] = 0.0f; s2c.t[2] = 1.0f;
        // raster_to_screen
        Affine r2s; r2s.set_identity();
        r2s.m[0][0] = -(1.0f / (static_cast<float>(W) / 2.0f)); r2s.m[1][1] = -(1.0f / (static_cast<float>(H) / 2.0f)); r2s.m[2][2] = 0.0f;
        r2s.t[0] = 1.0f; r2s.t[1] = 1.0f; r2s.t[2] = 0.0f;
        // m_screen_to_world_dir = c2w.linear() * s2c ; m_raster_to_world_dir = that * r2s
        Affine lin = c2w; lin.t[0] = lin.t[1] = lin.t[2] = 0.0f;
        Affine s2w = compose(lin, s2c);
        raster_to_world_dir = compose(s2w, r2s);
    }
    void generate_ray(int64_t px, int64_t py, float jx, float jy, V3& origin, V3& direction) const
    {
        float rx = (static_cast<float>(px) + 0.5f) + jx, ry = (static_cast<float>(py) + 0.5f) + jy;
        origin = V3(params.position[0], params.position[1], params.position[2]);
        direction = normalized(raster_to_world_dir.apply(V3(rx, ry, 0.0f)));
    }
    vptc_camera_parameters params;
    Affine raster_to_world_dir;
};

// ---------------------------------------------------------------- src/spectral.cpp:7-20, include/vpt/spectral.hpp:16-75,
//                                                                  src/precompute_blackbody.cpp:7-52
namespace bb {
static const float CIE[471][3] = {
#include "cie1931_xyz_1nm.inc"
};
constexpr float Y_integral = VPT_CIE_Y_INTEGRAL;

inline float planck_law(float lambda_m, float temperature_k)
{
    if (temperature_k <= 0.0f) return 0.0f;
    constexpr float c = 299792458.f;
    constexpr float h = 6.62606957e-34f;
    constexpr float kb = 1.3806488e-23f;
    constexpr float num = 2 * h * c * c;
    float lambda5 = std::pow(lambda_m, 5); // float,int overload: evaluated in double, rounded on assignment
    float exp = std::exp((h * c) / (lambda_m * kb * temperature_k));
    float den = lambda5 * (exp - 1);
    return num / den;
}
inline V3 spectrum_to_xyz(float temperature)
{
    float acc[3];
    for (int ch = 0; ch < 3; ++ch) {
        float integral = 0.0f;
        for (unsigned int i = 0; i < 471; ++i) integral += CIE[i][ch] * planck_law(static_cast<float>(360 + i) * 1e-9f, temperature);
        acc[ch] = integral;
    }
    return V3(acc[0], acc[1], acc[2]) / Y_integral;
}
constexpr unsigned int N_BREAKPOINTS = 500;
constexpr float RESOLUTION = 50000.0f / N_BREAKPOINTS;
constexpr float TEMP_MAX = (N_BREAKPOINTS - 1) * RESOLUTION;
inline float idx_to_temp(int idx) { return (idx - 1) * RESOLUTION; }

struct Table {
    std::array<V3, N_BREAKPOINTS + 1> xyz; // +1: the reference reads bb_xyz[dn_idx+1] one past the end for
                                           // 49800 K < T < 49900 K (latent bug); we pad with 0.
    Table() { for (unsigned int i = 0; i < N_BREAKPOINTS; ++i) xyz[i] = spectrum_to_xyz(idx_to_temp(int(i))); xyz[N_BREAKPOINTS] = V3(); }
};
inline const Table& table() { static const Table t; return t; }

inline V3 blackbody_radiation_xyz(float temperature_k)
{
    if (!std::isfinite(temperature_k)) { const float n = std::numeric_limits<float>::quiet_NaN(); return V3(n, n, n); }
    if (temperature_k <= 0.0f) return V3();
    if (temperature_k >= TEMP_MAX) return spectrum_to_xyz(temperature_k);
    const auto& T = table().xyz;
    int dn_idx = static_cast<int>(temperature_k / RESOLUTION);
    while (temperature_k <= idx_to_temp(dn_idx - 1)) --dn_idx;
    while (temperature_k >= idx_to_temp(dn_idx + 1)) ++dn_idx;
    float dn_temp = idx_to_temp(dn_idx);
    if (temperature_k == dn_temp) return T[dn_idx];
    float t = (temperature_k - idx_to_temp(dn_idx)) / RESOLUTION;
    return T[dn_idx] + (T[dn_idx + 1] - T[dn_idx]) * t; // lerp(a, b, t) = a + (b - a) * t (utils.hpp:26-29)
}
} // namespace bb

} // namespace vpto

#include "vpt_restated_volume.inl"
