// synth.cpp — seeded procedural volumes of the shapes BASELINE.json names, emitted as
// serialized NanoVDB float grids via nvdb_build. The reference ships no volumes (`volumes/` is
// git-ignored, .gitignore:3); its only built-in procedural source is VolumeGrids::generate_donut
// (src/volume_grids.cpp:35-37). Everything here is ours.
//
// Determinism: integer lattice hashes and + - * / sqrt floor only (no libm transcendentals), so the
// same (kind, dims, voxel_size, seed) yields the same bytes on every machine.
#include "synth.hpp"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <thread>

namespace vpt_host {
namespace {

inline uint32_t hash4(int32_t x, int32_t y, int32_t z, uint32_t s)
{
    uint32_t h = uint32_t(x) * 0x8da6b343u ^ uint32_t(y) * 0xd8163841u ^ uint32_t(z) * 0xcb1ab31fu ^ (s * 0x9e3779b9u + 0x7f4a7c15u);
    h ^= h >> 16; h *= 0x7feb352du; h ^= h >> 15; h *= 0x846ca68bu; h ^= h >> 16;
    return h;
}
inline float hash01(int32_t x, int32_t y, int32_t z, uint32_t s) { return float(hash4(x, y, z, s) >> 8) * (1.0f / 16777216.0f); }

// Sum of value-noise octaves evaluated on one 8^3 leaf. Lattice cell sizes are powers of two so a leaf
// touches (8/cell+1)^3 lattice points (2^3 when cell >= 8): the hashes are hoisted out of the voxel loop.
struct LeafNoise {
    float v[512];
    void clear() { std::memset(v, 0, sizeof(v)); }
    void add_octave(const int32_t o[3], int cell, float amp, uint32_t seed)
    {
        const int nc = cell >= 8 ? 1 : 8 / cell, np = nc + 1;
        float lat[9 * 9 * 9];
        int32_t base[3];
        for (int a = 0; a < 3; ++a) base[a] = int32_t(std::floor(double(o[a]) / cell));
        for (int i = 0; i < np; ++i) for (int j = 0; j < np; ++j) for (int k = 0; k < np; ++k)
            lat[(i * np + j) * np + k] = hash01(base[0] + i, base[1] + j, base[2] + k, seed);
        int   ci[3][8]; float w[3][8];
        for (int a = 0; a < 3; ++a) for (int i = 0; i < 8; ++i) {
            const int32_t g = o[a] + i;
            const int32_t c = int32_t(std::floor(double(g) / cell));
            const float t = float(g - c * cell) / float(cell);
            ci[a][i] = c - base[a];
            w[a][i] = t * t * (3.0f - 2.0f * t);
        }
        for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) {
            const float* l00 = &lat[((ci[0][i]) * np + ci[1][j]) * np];
            const float* l01 = l00 + np;
            const float* l10 = l00 + np * np;
            const float* l11 = l10 + np;
            const float wx = w[0][i], wy = w[1][j];
            for (int k = 0; k < 8; ++k) {
                const int c = ci[2][k]; const float wz = w[2][k];
                const float a00 = l00[c] + wz * (l00[c + 1] - l00[c]);
                const float a01 = l01[c] + wz * (l01[c + 1] - l01[c]);
                const float a10 = l10[c] + wz * (l10[c + 1] - l10[c]);
                const float a11 = l11[c] + wz * (l11[c + 1] - l11[c]);
                const float b0 = a00 + wy * (a01 - a00), b1 = a10 + wy * (a11 - a10);
                v[(i << 6) | (j << 3) | k] += amp * (b0 + wx * (b1 - b0));
            }
        }
    }
    // fbm in [0,1): cells top, top/2, ... (octaves), amplitudes 1/2, 1/4, ... renormalised
    void fbm(const int32_t o[3], int top_cell, int octaves, uint32_t seed)
    {
        clear();
        float amp = 0.5f, total = 0.0f;
        int cell = top_cell;
        for (int n = 0; n < octaves && cell >= 1; ++n, cell /= 2, amp *= 0.5f) { add_octave(o, cell, amp, seed + 101u * n); total += amp; }
        const float inv = 1.0f / total;
        for (float& x : v) x *= inv;
    }
};

struct Field {
    virtual ~Field() = default;
    // false => the block [o, o+dim)^3 is certainly all-background (conservative)
    virtual bool maybe(const int32_t o[3], int dim) const = 0;
    virtual void leaf(const int32_t o[3], float out[512]) const = 0;
};

inline float clamp01(float x) { return x < 0.0f ? 0.0f : (x > 1.0f ? 1.0f : x); }
// distance from point c to the box [o, o+dim]
inline float box_dist(const float c[3], const int32_t o[3], int dim, const float scale[3])
{
    float d2 = 0;
    for (int a = 0; a < 3; ++a) {
        const float lo = float(o[a]), hi = float(o[a] + dim);
        const float d = c[a] < lo ? lo - c[a] : (c[a] > hi ? c[a] - hi : 0.0f);
        d2 += (d * scale[a]) * (d * scale[a]);
    }
    return std::sqrt(d2);
}

// ---- fire: turbulent plume, density 0..~2 and a dimensionless temperature 0..50 (T_K = 43 t + 300) ----
struct FireField : Field {
    int32_t dims[3]; uint32_t seed; bool temperature;
    FireField(const int32_t d[3], uint32_t s, bool t) : seed(s), temperature(t) { std::memcpy(dims, d, sizeof(dims)); }
    float radius_at(float h) const { return (0.16f + 0.30f * h) * float(std::min(dims[0], dims[2])); }
    bool maybe(const int32_t o[3], int dim) const override
    {
        if (o[1] + dim <= 0 || o[1] >= dims[1]) return false;
        const float h1 = clamp01(float(o[1] + dim) / float(dims[1]));
        const float rmax = radius_at(h1) * 1.45f + 2.0f; // noise can widen the plume by < 45 %
        const float c[3] = {0.5f * dims[0], 0.0f, 0.5f * dims[2]};
        const int32_t oo[3] = {o[0], 0, o[2]};
        const float sc[3] = {1, 0, 1};
        return box_dist(c, oo, dim, sc) <= rmax;
    }
    void leaf(const int32_t o[3], float out[512]) const override
    {
        LeafNoise shape, detail;
        shape.fbm(o, 32, 3, seed);
        detail.fbm(o, 8, 3, seed + 7777u);
        for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) for (int k = 0; k < 8; ++k) {
            const int n = (i << 6) | (j << 3) | k;
            const float x = float(o[0] + i) - 0.5f * dims[0], z = float(o[2] + k) - 0.5f * dims[2];
            const float y = float(o[1] + j);
            float val = 0.0f;
            if (y >= 0.0f && y < float(dims[1])) {
                const float h = y / float(dims[1]);
                const float r = radius_at(h) * (0.55f + 0.9f * shape.v[n]);
                const float d = std::sqrt(x * x + z * z) / r;
                const float core = 1.0f - d * d;
                if (core > 0.0f) {
                    const float fade = (1.0f - h) * (1.0f - h) * (3.0f - 2.0f * (1.0f - h)); // smooth fade to the top
                    const float wisps = 0.25f + 1.5f * detail.v[n];
                    if (!temperature) {
                        val = 2.0f * core * fade * wisps - 0.08f;
                    } else {
                        // hot in the core and low in the plume; present wherever the density is
                        val = (2.0f * core * fade * wisps - 0.08f) > 0.0f ? 4.0f + 46.0f * clamp01(core * (0.35f + 0.65f * fade) * (0.5f + detail.v[n])) : 0.0f;
                    }
                    if (val < 0.0f) val = 0.0f;
                }
            }
            out[n] = val;
        }
    }
};

// ---- cloud: fBm-thresholded cumulus inside an ellipsoid; density 0..1, sparse ----
struct CloudField : Field {
    int32_t dims[3]; uint32_t seed; int top;
    CloudField(const int32_t d[3], uint32_t s) : seed(s)
    {
        std::memcpy(dims, d, sizeof(dims));
        const int m = std::min(dims[0], std::min(dims[1], dims[2]));
        top = 8; while (top * 6 <= m && top < 256) top *= 2; // coarsest lattice ~ 1/6 of the short side
    }
    float envelope(float x, float y, float z) const
    {
        const float ex = (x - 0.5f * dims[0]) / (0.5f * dims[0]), ey = (y - 0.42f * dims[1]) / (0.56f * dims[1]), ez = (z - 0.5f * dims[2]) / (0.5f * dims[2]);
        const float r2 = ex * ex + ey * ey + ez * ez;
        return r2 >= 1.0f ? 0.0f : (1.0f - r2);
    }
    bool maybe(const int32_t o[3], int dim) const override
    {
        const float c[3] = {0.5f * dims[0], 0.42f * dims[1], 0.5f * dims[2]};
        const float sc[3] = {1.0f / (0.5f * dims[0]), 1.0f / (0.56f * dims[1]), 1.0f / (0.5f * dims[2])};
        const float d = box_dist(c, o, dim, sc);
        // density > 0 needs envelope * 1.9 * fbm > 0.62 with fbm < 1  =>  envelope > 0.326
        return d < 1.0f && (1.0f - d * d) > 0.32f;
    }
    void leaf(const int32_t o[3], float out[512]) const override
    {
        LeafNoise nz;
        nz.fbm(o, top, 5, seed);
        for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) for (int k = 0; k < 8; ++k) {
            const int n = (i << 6) | (j << 3) | k;
            const float e = envelope(float(o[0] + i), float(o[1] + j), float(o[2] + k));
            const float v = (e * 1.9f * nz.v[n] - 0.62f) * 4.0f;
            out[n] = clamp01(v);
        }
    }
};

// ---- dense: strictly positive, non-constant noise over the whole box (every brick is a leaf) ----
struct DenseField : Field {
    int32_t dims[3]; uint32_t seed;
    DenseField(const int32_t d[3], uint32_t s) : seed(s) { std::memcpy(dims, d, sizeof(dims)); }
    bool maybe(const int32_t o[3], int dim) const override
    {
        for (int a = 0; a < 3; ++a) if (o[a] + dim <= 0 || o[a] >= dims[a]) return false;
        return true;
    }
    void leaf(const int32_t o[3], float out[512]) const override
    {
        LeafNoise nz;
        nz.fbm(o, 32, 4, seed);
        for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) for (int k = 0; k < 8; ++k) {
            const int n = (i << 6) | (j << 3) | k;
            const bool in = o[0] + i >= 0 && o[0] + i < dims[0] && o[1] + j >= 0 && o[1] + j < dims[1] && o[2] + k >= 0 && o[2] + k < dims[2];
            out[n] = in ? 0.04f + 0.96f * nz.v[n] : 0.0f;
        }
    }
};

// ---- torus: soft fog torus centred in the box ----
struct TorusField : Field {
    int32_t dims[3]; uint32_t seed;
    TorusField(const int32_t d[3], uint32_t s) : seed(s) { std::memcpy(dims, d, sizeof(dims)); }
    bool maybe(const int32_t o[3], int dim) const override
    {
        for (int a = 0; a < 3; ++a) if (o[a] + dim <= 0 || o[a] >= dims[a]) return false;
        return true;
    }
    void leaf(const int32_t o[3], float out[512]) const override
    {
        const float R = 0.33f * float(std::min(dims[0], dims[2])), r = 0.4f * float(dims[1]);
        for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) for (int k = 0; k < 8; ++k) {
            const float x = float(o[0] + i) - 0.5f * dims[0], y = float(o[1] + j) - 0.5f * dims[1], z = float(o[2] + k) - 0.5f * dims[2];
            const float q = std::sqrt(x * x + z * z) - R;
            const float d = std::sqrt(q * q + y * y) / r;
            out[(i << 6) | (j << 3) | k] = d < 1.0f ? clamp01(1.5f * (1.0f - d)) : 0.0f;
        }
    }
};

std::vector<uint8_t> build_from_field(const Field& f, const int32_t lo[3], const int32_t hi[3], const BuildParams& P,
                                      const std::vector<BuildTile>& tiles, const std::vector<std::array<int32_t, 3>>& skipLeaves)
{
    // block loop at 128^3, parallel; leaves with no active voxel are dropped (like build::Grid, which
    // never creates nodes for background-valued voxels, NV/tools/GridBuilder.h:1924-1990)
    int32_t b0[3], nb[3];
    for (int a = 0; a < 3; ++a) { b0[a] = lo[a] & ~127; nb[a] = ((hi[a] | 127) + 1 - b0[a]) / 128; }
    const size_t nBlocks = size_t(nb[0]) * nb[1] * nb[2];
    std::vector<BuildLeaf> leaves;
    std::mutex mtx;
    std::atomic<size_t> next{0};
    auto worker = [&]() {
        std::vector<BuildLeaf> local;
        for (;;) {
            const size_t b = next.fetch_add(1);
            if (b >= nBlocks) break;
            const int32_t bo[3] = {b0[0] + int32_t(b / (size_t(nb[1]) * nb[2])) * 128,
                                   b0[1] + int32_t((b / nb[2]) % nb[1]) * 128, b0[2] + int32_t(b % nb[2]) * 128};
            if (!f.maybe(bo, 128)) continue;
            for (int i = 0; i < 16; ++i) for (int j = 0; j < 16; ++j) for (int k = 0; k < 16; ++k) {
                const int32_t o[3] = {bo[0] + 8 * i, bo[1] + 8 * j, bo[2] + 8 * k};
                if (o[0] > hi[0] || o[1] > hi[1] || o[2] > hi[2] || o[0] + 7 < lo[0] || o[1] + 7 < lo[1] || o[2] + 7 < lo[2]) continue;
                if (!f.maybe(o, 8)) continue;
                bool skip = false;
                for (const auto& s : skipLeaves) skip |= (s[0] == o[0] && s[1] == o[1] && s[2] == o[2]);
                if (skip) continue;
                BuildLeaf L;
                std::memcpy(L.origin, o, sizeof(o));
                f.leaf(o, L.values);
                std::memset(L.mask, 0, sizeof(L.mask));
                bool any = false;
                for (uint32_t n = 0; n < 512; ++n) {
                    const int32_t c[3] = {o[0] + int32_t(n >> 6), o[1] + int32_t((n >> 3) & 7), o[2] + int32_t(n & 7)};
                    const bool inside = c[0] >= lo[0] && c[0] <= hi[0] && c[1] >= lo[1] && c[1] <= hi[1] && c[2] >= lo[2] && c[2] <= hi[2];
                    if (!inside) L.values[n] = P.background;
                    if (L.values[n] != P.background) { nvdbf::mask_set(L.mask, n); any = true; }
                }
                if (any) local.push_back(L);
            }
            if (local.size() > 4096) { std::lock_guard<std::mutex> g(mtx); leaves.insert(leaves.end(), local.begin(), local.end()); local.clear(); }
        }
        std::lock_guard<std::mutex> g(mtx);
        leaves.insert(leaves.end(), local.begin(), local.end());
    };
    unsigned nt = std::max(1u, std::thread::hardware_concurrency());
    nt = std::min<unsigned>(nt, 64u);
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; ++t) th.emplace_back(worker);
    for (auto& t : th) t.join();
    return build_grid(leaves, tiles, P);
}

} // namespace

SynthResult synth_volume(const std::string& kind, const int32_t dims[3], double voxel_size, uint32_t seed, const double* anchor)
{
    const double zero3[3] = {0, 0, 0};
    if (!anchor) anchor = zero3;
    if (dims[0] <= 0 || dims[1] <= 0 || dims[2] <= 0 || !(voxel_size > 0)) throw std::runtime_error("synth_volume: dims and voxel_size must be positive");
    SynthResult R;
    BuildParams P;
    P.voxel_size = voxel_size;
    const std::vector<BuildTile> noTiles;
    const std::vector<std::array<int32_t, 3>> noSkip;
    int32_t lo[3] = {0, 0, 0}, hi[3] = {dims[0] - 1, dims[1] - 1, dims[2] - 1};

    if (kind == "fire") {
        // world bbox [-W/2, W/2] x [0, H] x [-D/2, D/2] like a plume standing on the ground plane
        P.translation[0] = anchor[0] - 0.5 * dims[0] * voxel_size; P.translation[1] = anchor[1]; P.translation[2] = anchor[2] - 0.5 * dims[2] * voxel_size;
        P.name = "density";
        R.density = build_from_field(FireField(dims, seed, false), lo, hi, P, noTiles, noSkip);
        P.name = "temperature"; P.grid_class = nvdbf::GRIDCLASS_UNKNOWN;
        R.temperature = build_from_field(FireField(dims, seed, true), lo, hi, P, noTiles, noSkip);
    } else if (kind == "cloud") {
        // centred on the world origin, index origin at the lower corner shifted so the box straddles
        // index 0 (as the real WDAS cloud does): exercises negative coordinates and 8 root entries
        for (int a = 0; a < 3; ++a) { lo[a] = -((dims[a] / 2 + 7) & ~7); hi[a] = lo[a] + dims[a] - 1; P.translation[a] = anchor[a]; }
        struct Shifted : Field {
            CloudField f; int32_t s[3];
            Shifted(const int32_t d[3], uint32_t sd, const int32_t l[3]) : f(d, sd) { std::memcpy(s, l, sizeof(s)); }
            bool maybe(const int32_t o[3], int dim) const override { const int32_t p[3] = {o[0] - s[0], o[1] - s[1], o[2] - s[2]}; return f.maybe(p, dim); }
            void leaf(const int32_t o[3], float out[512]) const override
            {
                // lattice noise is evaluated in shifted coordinates; shift is NOT a multiple of 8 in
                // general, so evaluate per voxel through the 8-aligned leaves it overlaps
                const int32_t p[3] = {o[0] - s[0], o[1] - s[1], o[2] - s[2]};
                const int32_t a[3] = {p[0] & ~7, p[1] & ~7, p[2] & ~7};
                float tmp[8][512]; bool have[8] = {false, false, false, false, false, false, false, false};
                for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) for (int k = 0; k < 8; ++k) {
                    const int32_t q[3] = {p[0] + i, p[1] + j, p[2] + k};
                    const int c = (((q[0] & ~7) != a[0]) << 2) | (((q[1] & ~7) != a[1]) << 1) | int((q[2] & ~7) != a[2]);
                    if (!have[c]) { const int32_t lo2[3] = {q[0] & ~7, q[1] & ~7, q[2] & ~7}; f.leaf(lo2, tmp[c]); have[c] = true; }
                    out[(i << 6) | (j << 3) | k] = tmp[c][((q[0] & 7) << 6) | ((q[1] & 7) << 3) | (q[2] & 7)];
                }
            }
        };
        P.name = "density";
        R.density = build_from_field(Shifted(dims, seed, lo), lo, hi, P, noTiles, noSkip);
    } else if (kind == "dense") {
        for (int a = 0; a < 3; ++a) P.translation[a] = anchor[a] - 0.5 * dims[a] * voxel_size;
        P.name = "density";
        R.density = build_from_field(DenseField(dims, seed), lo, hi, P, noTiles, noSkip);
    } else if (kind == "torus") {
        for (int a = 0; a < 3; ++a) { lo[a] = -(dims[a] / 2); hi[a] = lo[a] + dims[a] - 1; P.translation[a] = anchor[a]; }
        struct Centered : Field {
            TorusField f; int32_t s[3];
            Centered(const int32_t d[3], uint32_t sd, const int32_t l[3]) : f(d, sd) { std::memcpy(s, l, sizeof(s)); }
            bool maybe(const int32_t o[3], int dim) const override { const int32_t p[3] = {o[0] - s[0], o[1] - s[1], o[2] - s[2]}; return f.maybe(p, dim); }
            void leaf(const int32_t o[3], float out[512]) const override
            {
                // the torus is analytic: evaluate with per-voxel shifted coordinates
                TorusField g = f;
                const float R0 = 0.33f * float(std::min(g.dims[0], g.dims[2])), r = 0.4f * float(g.dims[1]);
                for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) for (int k = 0; k < 8; ++k) {
                    const float x = float(o[0] + i), y = float(o[1] + j), z = float(o[2] + k);
                    const float q = std::sqrt(x * x + z * z) - R0;
                    const float d = std::sqrt(q * q + y * y) / r;
                    out[(i << 6) | (j << 3) | k] = d < 1.0f ? clamp01(1.5f * (1.0f - d)) : 0.0f;
                }
            }
        };
        P.name = "density";
        R.density = build_from_field(Centered(dims, seed, lo), lo, hi, P, noTiles, noSkip);
    } else if (kind == "tiles") {
        // Test volume: noise leaves straddling the index origin plus constant tiles at every level, an
        // inactive-but-nonzero tile and an upper node without any active value (mFlags bit0 path).
        for (int a = 0; a < 3; ++a) { lo[a] = -(dims[a] / 2); hi[a] = lo[a] + dims[a] - 1; P.translation[a] = anchor[a] + 0.37 * (a + 1); }
        struct Off : Field {
            DenseField f; int32_t s[3];
            Off(const int32_t d[3], uint32_t sd, const int32_t l[3]) : f(d, sd) { std::memcpy(s, l, sizeof(s)); }
            bool maybe(const int32_t o[3], int dim) const override { const int32_t p[3] = {o[0] - s[0], o[1] - s[1], o[2] - s[2]}; return f.maybe(p, dim); }
            void leaf(const int32_t o[3], float out[512]) const override
            {
                LeafNoise nz; nz.fbm(o, 16, 3, f.seed);
                for (int n = 0; n < 512; ++n) { const float v = nz.v[n] * 1.6f - 0.55f; out[n] = v > 0.0f ? v : 0.0f; }
            }
        };
        std::vector<BuildTile> tiles;
        std::vector<std::array<int32_t, 3>> skip;
        // level-1 tiles: replace two leaf slots inside the noise region, one active one inactive-nonzero
        const int32_t t1a[3] = {8, 8, 8}, t1b[3] = {-16, 8, 0}, t1c[3] = {16, -8, -24};
        tiles.push_back({1, {t1a[0], t1a[1], t1a[2]}, 0.8f, true});  skip.push_back({t1a[0], t1a[1], t1a[2]});
        tiles.push_back({1, {t1b[0], t1b[1], t1b[2]}, 0.6f, false}); skip.push_back({t1b[0], t1b[1], t1b[2]});
        tiles.push_back({1, {t1c[0], t1c[1], t1c[2]}, 0.0f, true});  skip.push_back({t1c[0], t1c[1], t1c[2]});
        // level-2 tiles next to the leaf region (outside [lo,hi] rounded to 128)
        const int32_t edge = ((hi[0] | 127) + 1);
        tiles.push_back({2, {edge, 0, 0}, 0.3f, true});
        tiles.push_back({2, {edge, -128, 0}, 0.9f, false});
        // an upper node holding only an inactive tile: no active value below it -> flag bit0 -> getDim 4096
        tiles.push_back({2, {4096, 0, 0}, 0.5f, false});
        // an inactive root tile with a value (getValue returns it; majorant 0)
        tiles.push_back({3, {0, 4096, 0}, 0.25f, false});
        P.name = "density";
        R.density = build_from_field(Off(dims, seed, lo), lo, hi, P, tiles, skip);
        P.name = "temperature"; P.grid_class = nvdbf::GRIDCLASS_UNKNOWN;
        P.voxel_size = voxel_size * 1.5; // a different map: the temperature lookup must use its own
        P.translation[0] -= 1.0;
        struct Temp : Field {
            DenseField f; int32_t s[3];
            Temp(const int32_t d[3], uint32_t sd, const int32_t l[3]) : f(d, sd) { std::memcpy(s, l, sizeof(s)); }
            bool maybe(const int32_t o[3], int dim) const override { const int32_t p[3] = {o[0] - s[0], o[1] - s[1], o[2] - s[2]}; return f.maybe(p, dim); }
            void leaf(const int32_t o[3], float out[512]) const override
            {
                LeafNoise nz; nz.fbm(o, 16, 2, f.seed + 99u);
                for (int n = 0; n < 512; ++n) out[n] = 60.0f * nz.v[n];
            }
        };
        R.temperature = build_from_field(Temp(dims, seed, lo), lo, hi, P, noTiles, noSkip);
    } else {
        throw std::runtime_error("synth_volume: unknown kind \"" + kind + "\" (fire|cloud|dense|torus|tiles)");
    }
    return R;
}

} // namespace vpt_host
