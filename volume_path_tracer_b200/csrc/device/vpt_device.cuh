// vpt_device.cuh — device-side building blocks of the path integrator (sm_100a).
//
// Floating-point contract: this translation unit is compiled with --fmad=false, IEEE division and
// square root (nvcc defaults -prec-div=true -prec-sqrt=true, no -use_fast_math, denormals kept), so every
// + - * / sqrt floor rounds exactly as the reference's baseline-x86-64 Release build does; the only fused
// operations are the explicit fmaf calls NanoVDB itself makes (NV/math/Ray.h:154-159, Math.h:905-939).
// logf / sinf / cosf are NOT CUDA's: libm_exact.cuh restates glibc's double-precision table+polynomial schemes,
// verified exhaustively on the argument domains used here, so branch decisions match the CPU reference.
#pragma once
#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>

#include "layout.hpp"
#include "libm_exact.cuh"

// Rarely-taken helpers (general trilinear fetch, blackbody lookup): inlined by default; -DVPT_COLD_NOINLINE keeps them
// out of line (an A/B switch for instruction-cache experiments).
#ifdef VPT_COLD_NOINLINE
#define VPT_COLD_INLINE __noinline__
#else
#define VPT_COLD_INLINE __forceinline__
#endif

namespace vptd {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ float sum3(float a, float b, float c) { return a + (b + c); } // Eigen order (oracle/vec3.hpp)
__device__ __forceinline__ float dot3(V3 a, V3 b) { return sum3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 normalized(V3 a)
{
    const float n = sqrtf(dot3(a, a));
    return mk(a.x / n, a.y / n, a.z / n);
}

// ---------------------------------------------------------------- RNG: include/vpt/random.hpp:86-115, hash.hpp:20-67
struct Rng {
    uint64_t state;
    uint32_t draws;
    __device__ __forceinline__ void begin_job(uint32_t seed, uint64_t jid)
    {
        const uint64_t m = 0xc6a4a7935bd1e995ULL;
        uint64_t h = uint64_t(seed) ^ (8ULL * m);
        uint64_t k = jid;
        k *= m; k ^= k >> 47; k *= m;
        h ^= k; h *= m;
        h ^= h >> 47; h *= m; h ^= h >> 47;
        state = h | 3ULL; // pcg32_fast (MCG): state = seed | 3
        draws = 0;
    }
    __device__ __forceinline__ uint32_t next_u32()
    {
        const uint64_t old = state;
        state = old * 6364136223846793005ULL;
        ++draws;
        const uint32_t rshift = uint32_t(old >> 61);
        const uint64_t x = old ^ (old >> 22);
        return uint32_t(x >> (22u + rshift));
    }
    __device__ __forceinline__ float uniform() { return fminf(0x1.fffffep-1f, __uint2float_rn(next_u32()) * 0x1p-32f); }
};

// ---------------------------------------------------------------- grid lookups
__device__ __forceinline__ uint32_t cell_code(const GridDev& G, int x, int y, int z)
{
    const int cx = (x - G.cell_org[0]) >> 7, cy = (y - G.cell_org[1]) >> 7, cz = (z - G.cell_org[2]) >> 7;
    if (unsigned(cx) >= unsigned(G.cell_dim[0]) || unsigned(cy) >= unsigned(G.cell_dim[1]) || unsigned(cz) >= unsigned(G.cell_dim[2]))
        return CELL_KIND_TILE | (2u << CELL_DIM_SHIFT); // outside every root entry: background, dim 4096
    return __ldg(&G.cells[(size_t(cx) * size_t(G.cell_dim[1]) + size_t(cy)) * size_t(G.cell_dim[2]) + size_t(cz)]);
}
__device__ __forceinline__ uint32_t slot_index(uint32_t cell, int x, int y, int z)
{
    return (cell & CELL_INDEX_MASK) * 4096u + ((uint32_t(x & 127) >> 3) << 8 | (uint32_t(y & 127) >> 3) << 4 | (uint32_t(z & 127) >> 3));
}
__device__ __forceinline__ Slot load_slot(const GridDev& G, uint32_t idx)
{
    const uint2 raw = __ldg(reinterpret_cast<const uint2*>(G.slots) + idx);
    Slot s; s.code = raw.x; s.majorant = __uint_as_float(raw.y);
    return s;
}
// max(8, acc.getDim(ijk, ray))   src/volume.cpp:11-14
__device__ __forceinline__ int hdda_dim(const GridDev& G, int x, int y, int z)
{
    const uint32_t c = cell_code(G, x, y, z) & CELL_DIM_MASK;
    return c == 0u ? 8 : (c == (1u << CELL_DIM_SHIFT) ? 128 : 4096);
}
// RayMajorantIterator::update_current_majorant   src/volume.cpp:18-36
__device__ __forceinline__ float majorant_at(const GridDev& G, int x, int y, int z)
{
    const uint32_t c = cell_code(G, x, y, z);
    if (c & CELL_KIND_TILE) return __ldg(&G.tile_majorant[c & CELL_INDEX_MASK]);
    return load_slot(G, slot_index(c, x, y, z)).majorant;
}
// ReadAccessor::getValue
__device__ __forceinline__ float value_at(const GridDev& G, int x, int y, int z)
{
    const uint32_t c = cell_code(G, x, y, z);
    if (c & CELL_KIND_TILE) return __ldg(&G.tile_value[c & CELL_INDEX_MASK]);
    const Slot s = load_slot(G, slot_index(c, x, y, z));
    if (s.code & SLOT_KIND_TILE) return __ldg(&G.tile_value[s.code & SLOT_INDEX_MASK]);
    return __ldg(&G.bricks[size_t(s.code) * BRICK_STRIDE + brick_offset(x & 7, y & 7, z & 7)]);
}
// The 8 corners of the trilinear stencil whose base voxel (x, y, z) lies in brick `code`: all inside the 9^3 apron brick.
__device__ __forceinline__ void brick_corners(const GridDev& G, uint32_t code, int x, int y, int z, float& v000, float& v001, float& v010, float& v011,
                                              float& v100, float& v101, float& v110, float& v111)
{
    const float* b = G.bricks + size_t(code) * BRICK_STRIDE + brick_offset(x & 7, y & 7, z & 7);
#if VPT_BRICK_PAIRS == 2
    const float4 lo = __ldg(reinterpret_cast<const float4*>(b)), hi = __ldg(reinterpret_cast<const float4*>(b) + 1);
    v000 = lo.x; v001 = lo.y; v010 = lo.z; v011 = lo.w;
    v100 = hi.x; v101 = hi.y; v110 = hi.z; v111 = hi.w;
#elif VPT_BRICK_PAIRS
    const float2 p00 = __ldg(reinterpret_cast<const float2*>(b)), p01 = __ldg(reinterpret_cast<const float2*>(b + BRICK_ROW));
    const float2 p10 = __ldg(reinterpret_cast<const float2*>(b + 9 * BRICK_ROW)), p11 = __ldg(reinterpret_cast<const float2*>(b + 10 * BRICK_ROW));
    v000 = p00.x; v001 = p00.y; v010 = p01.x; v011 = p01.y;
    v100 = p10.x; v101 = p10.y; v110 = p11.x; v111 = p11.y;
#else
    v000 = __ldg(b);      v001 = __ldg(b + 1);
    v010 = __ldg(b + 9);  v011 = __ldg(b + 10);
    v100 = __ldg(b + 81); v101 = __ldg(b + 82);
    v110 = __ldg(b + 90); v111 = __ldg(b + 91);
#endif
}
// SampleFromVoxels<Acc,1>::operator()(Vec3f)   NV/math/SampleFromVoxels.h:62-69,202-242
__device__ __forceinline__ float trilinear(const GridDev& G, float px, float py, float pz)
{
    const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
    const float u = px - fx, v = py - fy, w = pz - fz;
    const int x = int(fx), y = int(fy), z = int(fz);
    float v000, v001, v010, v011, v100, v101, v110, v111;
    bool fast = false;
    const uint32_t c = cell_code(G, x, y, z);
    if (!(c & CELL_KIND_TILE)) {
        const Slot s = load_slot(G, slot_index(c, x, y, z));
        if (!(s.code & SLOT_KIND_TILE)) { // base voxel in a leaf: the 9^3 apron brick holds all 8 corners
            brick_corners(G, s.code, x, y, z, v000, v001, v010, v011, v100, v101, v110, v111);
            fast = true;
        }
    }
    if (!fast) { // base voxel in a tile / background: corners may still reach into leaves
        v000 = value_at(G, x, y, z);         v001 = value_at(G, x, y, z + 1);
        v010 = value_at(G, x, y + 1, z);     v011 = value_at(G, x, y + 1, z + 1);
        v100 = value_at(G, x + 1, y, z);     v101 = value_at(G, x + 1, y, z + 1);
        v110 = value_at(G, x + 1, y + 1, z); v111 = value_at(G, x + 1, y + 1, z + 1);
    }
    // lerp(a, b, w) = a + w * (b - a), nested z -> y -> x
    const float a00 = v000 + w * (v001 - v000), a01 = v010 + w * (v011 - v010);
    const float a10 = v100 + w * (v101 - v100), a11 = v110 + w * (v111 - v110);
    const float b0 = a00 + v * (a01 - a00), b1 = a10 + v * (a11 - a10);
    return b0 + u * (b1 - b0);
}

// Map::applyInverseMapF / applyInverseJacobianF / applyMapF   NV/NanoVDB.h:1439-1497, Math.h:905-939
__device__ __forceinline__ V3 mat3(const float* m, V3 p)
{
    return mk(fmaf(p.x, m[0], fmaf(p.y, m[1], p.z * m[2])), fmaf(p.x, m[3], fmaf(p.y, m[4], p.z * m[5])), fmaf(p.x, m[6], fmaf(p.y, m[7], p.z * m[8])));
}
__device__ __forceinline__ V3 mat3t(const float* m, const float* t, V3 p)
{
    return mk(fmaf(p.x, m[0], fmaf(p.y, m[1], fmaf(p.z, m[2], t[0]))), fmaf(p.x, m[3], fmaf(p.y, m[4], fmaf(p.z, m[5], t[1]))),
              fmaf(p.x, m[6], fmaf(p.y, m[7], fmaf(p.z, m[8], t[2]))));
}
__device__ __forceinline__ V3 world_to_index(const GridDev& G, V3 w) { return mat3(G.invMat, mk(w.x - G.vec[0], w.y - G.vec[1], w.z - G.vec[2])); }
__device__ __forceinline__ V3 index_to_world(const GridDev& G, V3 p) { return mat3t(G.mat, G.vec, p); }

// ---------------------------------------------------------------- ray march state
// nanovdb::math::Ray<float> (index space) + HDDA + RayMajorantIterator + the sampler's current segment.
struct March {
    float ex, ey, ez, dx, dy, dz, ix, iy, iz; // eye, dir, 1/dir
    float scale;                               // idx_to_world_scale   volume.cpp:91
    int   dim;                                 // HDDA mDim
    float t0, t1;                              // HDDA mT0 (time), mT1 (maxTime)
    int   vx, vy, vz;                          // HDDA mVoxel
    float nx, ny, nz;                          // HDDA mNext
    float majorant;                            // NaN until first fetched
    float seg_t0, seg_t1, seg_dmaj;            // sampler's current segment
    bool  has_seg;
    __device__ __forceinline__ V3 at(float t) const { return mk(fmaf(t, dx, ex), fmaf(t, dy, ey), fmaf(t, dz, ez)); } // Ray::operator()
};

struct Stats { // per-thread work counters (COUNT builds only)
    uint32_t density_evals, temperature_evals, majorant_fetches, scatter_events, shadow_rays, pixel_samples;
};

// HDDA::init / update: voxel = floor(ray(t0)) & ~(dim-1) and the three "next" plane times.  HDDA.h:49-89
__device__ __forceinline__ void hdda_init(March& m, int dim)
{
    m.dim = dim;
    const V3 pos = m.at(m.t0);
    m.vx = int(floorf(pos.x)) & ~(dim - 1); m.vy = int(floorf(pos.y)) & ~(dim - 1); m.vz = int(floorf(pos.z)) & ~(dim - 1);
    // dir == 0: disabled (FLT_MAX); inv > 0: far plane; else near plane
    m.nx = m.dx == 0.0f ? FLT_MAX : (m.ix > 0.0f ? m.t0 + (float(m.vx + dim) - pos.x) * m.ix : m.t0 + (float(m.vx) - pos.x) * m.ix);
    m.ny = m.dy == 0.0f ? FLT_MAX : (m.iy > 0.0f ? m.t0 + (float(m.vy + dim) - pos.y) * m.iy : m.t0 + (float(m.vy) - pos.y) * m.iy);
    m.nz = m.dz == 0.0f ? FLT_MAX : (m.iz > 0.0f ? m.t0 + (float(m.vz + dim) - pos.z) * m.iz : m.t0 + (float(m.vz) - pos.z) * m.iz);
}
__device__ __forceinline__ void hdda_update(March& m, int dim)
{
    if (m.dim == dim) return;
    m.dim = dim;
    const V3 pos = m.at(m.t0);
    m.vx = int(floorf(pos.x)) & ~(dim - 1); m.vy = int(floorf(pos.y)) & ~(dim - 1); m.vz = int(floorf(pos.z)) & ~(dim - 1);
    const float fd = float(dim);
    if (m.dx != 0.0f) { m.nx = m.t0 + (float(m.vx) - pos.x) * m.ix; if (m.ix > 0.0f) m.nx += fd * m.ix; }
    if (m.dy != 0.0f) { m.ny = m.t0 + (float(m.vy) - pos.y) * m.iy; if (m.iy > 0.0f) m.ny += fd * m.iy; }
    if (m.dz != 0.0f) { m.nz = m.t0 + (float(m.vz) - pos.z) * m.iz; if (m.iz > 0.0f) m.nz += fd * m.iz; }
}
// HDDA::step  HDDA.h:95-160 (MinIndex tie-break Math.h:285-298; ENFORCE_FORWARD_STEPPING nudge)
__device__ __forceinline__ bool hdda_step(March& m)
{
    const float fd = float(m.dim);
    if (m.nx < m.ny && m.nx < m.nz) {
        if (m.nx <= m.t0) m.nx += m.t0 - 0.999999f * m.nx + 1.0e-6f;
        m.t0 = m.nx; m.nx += fd * fabsf(m.ix); m.vx += m.ix > 0.0f ? m.dim : -m.dim;
    } else if (m.ny < m.nz) {
        if (m.ny <= m.t0) m.ny += m.t0 - 0.999999f * m.ny + 1.0e-6f;
        m.t0 = m.ny; m.ny += fd * fabsf(m.iy); m.vy += m.iy > 0.0f ? m.dim : -m.dim;
    } else {
        if (m.nz <= m.t0) m.nz += m.t0 - 0.999999f * m.nz + 1.0e-6f;
        m.t0 = m.nz; m.nz += fd * fabsf(m.iz); m.vz += m.iz > 0.0f ? m.dim : -m.dim;
    }
    return m.t0 <= m.t1;
}

// hdda_step without the three-way axis branch (the x / y / z arms diverge inside a warp): same operations on the
// selected axis, selected with predicated moves.
__device__ __forceinline__ bool hdda_step_bf(March& m)
{
    const bool ax = m.nx < m.ny && m.nx < m.nz, ay = !ax && m.ny < m.nz, az = !ax && !ay;
    float n = ax ? m.nx : (ay ? m.ny : m.nz);
    const float inv = ax ? m.ix : (ay ? m.iy : m.iz);
    if (n <= m.t0) n += m.t0 - 0.999999f * n + 1.0e-6f;
    m.t0 = n;
    const float n2 = n + float(m.dim) * fabsf(inv);
    const int step = inv > 0.0f ? m.dim : -m.dim;
    m.nx = ax ? n2 : m.nx; m.ny = ay ? n2 : m.ny; m.nz = az ? n2 : m.nz;
    m.vx += ax ? step : 0; m.vy += ay ? step : 0; m.vz += az ? step : 0;
    return m.t0 <= m.t1;
}

// Volume::intersect + RayMajorantIterator ctor   src/volume.cpp:78-98, NV/math/Ray.h:64-74,258-268,398-424
// The direction-only part of Ray::worldToIndexF + the iterator's idx_to_world scale: constant for all shadow rays
// of a scene (one distant light), so the kernels evaluate it once per block for those.
struct RayDir { float dx, dy, dz, ix, iy, iz, length, scale; };
__device__ __forceinline__ RayDir make_ray_dir(const GridDev& G, V3 d)
{
    RayDir r;
    const V3 di = mat3(G.invMat, d);
    r.length = sqrtf(di.x * di.x + di.y * di.y + di.z * di.z);
    const float invLength = 1.0f / r.length;
    r.dx = invLength * di.x; r.dy = invLength * di.y; r.dz = invLength * di.z;
    r.ix = 1.0f / r.dx; r.iy = 1.0f / r.dy; r.iz = 1.0f / r.dz;
    // m_scale = 1 / density.worldToIndexDirF(ray.dir()).length(), with the INDEX-space direction (volume.cpp:91)
    const V3 q = mat3(G.invMat, mk(r.dx, r.dy, r.dz));
    r.scale = 1.0f / sqrtf(q.x * q.x + q.y * q.y + q.z * q.z);
    return r;
}
// Everything of march_begin except the HDDA init (the caller picks the cached or uncached getDim lookup).
__device__ __forceinline__ bool march_begin_dir(March& m, const GridDev& G, V3 o, const RayDir& r)
{
    const V3 eye = world_to_index(G, o);
    m.ex = eye.x; m.ey = eye.y; m.ez = eye.z;
    m.dx = r.dx; m.dy = r.dy; m.dz = r.dz;
    m.ix = r.ix; m.iy = r.iy; m.iz = r.iz;
    float t0 = r.length * 1e-5f, t1 = FLT_MAX; // Delta<float>, host Maximum<float>
    // clip against indexBBox with max + 1
    {
        float a = float(G.bbox_min[0]), b = float(G.bbox_max[0] + 1);
        if (a >= b) return false;
        a = (a - m.ex) * m.ix; b = (b - m.ex) * m.ix;
        if (a > b) { const float s = a; a = b; b = s; }
        if (a > t0) t0 = a;
        if (b < t1) t1 = b;
        if (t0 > t1) return false;
        a = float(G.bbox_min[1]); b = float(G.bbox_max[1] + 1);
        if (a >= b) return false;
        a = (a - m.ey) * m.iy; b = (b - m.ey) * m.iy;
        if (a > b) { const float s = a; a = b; b = s; }
        if (a > t0) t0 = a;
        if (b < t1) t1 = b;
        if (t0 > t1) return false;
        a = float(G.bbox_min[2]); b = float(G.bbox_max[2] + 1);
        if (a >= b) return false;
        a = (a - m.ez) * m.iz; b = (b - m.ez) * m.iz;
        if (a > b) { const float s = a; a = b; b = s; }
        if (a > t0) t0 = a;
        if (b < t1) t1 = b;
        if (t0 > t1) return false;
    }
    m.t0 = t0; m.t1 = t1;
    m.scale = r.scale;
    m.majorant = __int_as_float(0x7fc00000); // NaN: not fetched yet
    m.has_seg = false;
    return true;
}
__device__ __forceinline__ bool march_begin(March& m, const GridDev& G, V3 o, V3 d)
{
    if (!march_begin_dir(m, G, o, make_ray_dir(G, d))) return false;
    const V3 s = m.at(m.t0);
    hdda_init(m, hdda_dim(G, int(floorf(s.x)), int(floorf(s.y)), int(floorf(s.z))));
    return true;
}

// RayMajorantIterator::next   src/volume.cpp:38-76. Returns false when the ray has left the bbox.
template <bool COUNT>
__device__ __forceinline__ bool iterator_next(March& m, const GridDev& G, Stats& st)
{
    if (m.t0 >= m.t1) return false;
    m.seg_t0 = m.t0;
    if (isnan(m.majorant)) { m.majorant = majorant_at(G, m.vx, m.vy, m.vz); if (COUNT) ++st.majorant_fetches; }
    do {
        m.seg_dmaj = m.majorant;
        if (!hdda_step(m)) { m.seg_t1 = m.t1; return true; }
        const V3 p = m.at(m.t0 + 1.0001f);
        hdda_update(m, hdda_dim(G, int(floorf(p.x)), int(floorf(p.y)), int(floorf(p.z))));
        m.majorant = majorant_at(G, m.vx, m.vy, m.vz);
        if (COUNT) ++st.majorant_fetches;
    } while (m.majorant == m.seg_dmaj);
    m.seg_t1 = m.t0;
    return true;
}

struct Collision { V3 point; float sigma_maj, density; };

// MajorantTransmittanceSampler::next   src/majorant_transmittance_sampler.cpp:21-81
template <bool COUNT>
__device__ __forceinline__ bool sampler_next(March& m, const GridDev& G, Rng& rng, float sigma_t, const double* logtab, Collision& out, Stats& st)
{
    for (;;) {
        if (!m.has_seg) {
            if (!iterator_next<COUNT>(m, G, st)) return false;
            if (m.seg_dmaj <= 0.0f) continue; // empty space: next segment
            m.has_seg = true;
        }
        const float sigma_maj = m.seg_dmaj * sigma_t;
        const float dt_m = -vptm::logf_exact(1.0f - rng.uniform(), logtab) / sigma_maj; // sample_exponential  random.hpp:19-22
        const float t = m.seg_t0 + dt_m / m.scale;
        if (t < m.seg_t1) {
            m.seg_t0 = t;
            const V3 p = m.at(t);
            if (COUNT) ++st.density_evals;
            const float density = trilinear(G, p.x, p.y, p.z);
            if (density <= 0.0f) continue;
            out.point = index_to_world(G, p);
            out.sigma_maj = sigma_maj;
            out.density = density;
            return true;
        }
        m.has_seg = false;
    }
}

// ---------------------------------------------------------------- phase function  utils.hpp:39-66, random.hpp:56-84
__device__ __forceinline__ float henyey_greenstein(float cos_theta, float g)
{
    const float den = 1.0f + g * g + 2.0f * g * cos_theta;
    const float inv_4_pi = 0.07957747154594767f; // float(inv_pi / 4.0)
    return inv_4_pi * (1.0f - g * g) / (den * sqrtf(fmaxf(0.0f, den)));
}
__device__ __forceinline__ V3 sample_henyey_greenstein(V3 w, float ux, float uy, float g)
{
    const float g2 = g * g; // powf(g, 2.0f): glibc returns the correctly rounded square in practice
    float cos_theta;
    if (fabsf(g) < 1e-3f) cos_theta = 1.0f - 2.0f * ux;
    else {
        const float q = (1.0f - g2) / (1.0f + g - 2.0f * g * ux);
        cos_theta = 1.0f / (2.0f * g) * (1.0f + g2 - q * q);
    }
    const float sin_theta = sqrtf(fmaxf(0.0f, 1.0f - cos_theta * cos_theta));
    const float phi = 6.2831854820251465f * uy; // 2.0f * float(pi) * u.y
    const float st = fminf(fmaxf(sin_theta, -1.0f), 1.0f), ct = fminf(fmaxf(cos_theta, -1.0f), 1.0f);
    float sphi, cphi;
    vptm::sincosf_exact(phi, sphi, cphi);
    V3 local = normalized(mk(st * cphi, st * sphi, ct));
    // coordinate_system(z = w, x, y)
    const float sign = copysignf(1.0f, w.z);
    const float a = -1.0f / (sign + w.z);
    const float b = w.x * w.y * a;
    const V3 x = mk(1.0f + sign * a * (w.x * w.x), sign * b, -sign * w.x);
    const V3 y = mk(b, sign + a * (w.y * w.y), -w.y);
    return mk(local.x * x.x + local.y * y.x + local.z * w.x, local.x * x.y + local.y * y.y + local.z * w.y,
              local.x * x.z + local.y * y.z + local.z * w.z);
}

} // namespace vptd
