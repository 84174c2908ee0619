// vpt_render_flat.cuh — the throughput megakernel: vpt::run (src/worker.cpp:92-208) + sample_Ld (:52-90) +
// TileProvider::next (src/tile_provider.cpp:27-67) as ONE flat per-thread state machine.
//
// Why flat. One thread = one (tile, wave) job, because the job's pcg32_fast stream is consumed serially across its
// pixels (worker.cpp:107-110) — that is what makes the result bit-identical to the reference. The nested form of
// the algorithm (pixel loop > bounce loop > sampler loop > HDDA loop, with a second copy of the sampler inside
// sample_Ld) leaves the 32 jobs of a warp in 32 different loop nests: the first ncu capture showed 2.8 active
// threads per issued instruction (profiles/r01_nested_*). Here every thread runs the same loop whose body is
//     [ray setup, rare]  ->  [one HDDA step, if the lane is looking for a majorant segment]
//                        ->  [one free-flight sample, if the lane has a segment]
// so lanes on camera rays, bounce rays and shadow rays, in any pixel of any tile, execute the two hot blocks
// together. The arithmetic inside the blocks is the same statement-for-statement restatement as vpt_device.cuh.
//
// Phases of a lane (`ph`): MARCH (in the two hot blocks) | NEXT_PIXEL | BEGIN_SHADOW | AFTER_SHADOW.
#pragma once
#include "vpt_device.cuh"
#include "vpt_kernels.hpp"

namespace vptd {

// 128^3-cell code with a one-entry per-thread cache: successive lookups of a marching ray almost always fall in
// the same cell (a dense 512^3 grid has only 4^3 of them), which removes two of the three dependent loads of an
// HDDA step (getDim probe + majorant descent) and one of the two of a trilinear fetch.
struct CellCache {
    int x, y, z;   // any voxel of the cached cell. Cells are 128-aligned in index space (cell_org is a multiple of 4096,
    uint32_t code; // bake.cpp), so "same cell" is "all bits above the low 7 equal": one xor/or chain instead of 3 sub+shift+compare
    // The sentinel voxel INT_MIN lies outside every grid bake.cpp accepts (|coordinate| <= 2^30): background, dim 4096.
    __device__ __forceinline__ void reset() { x = y = z = INT_MIN; code = CELL_KIND_TILE | (2u << CELL_DIM_SHIFT); }
};
__device__ __forceinline__ uint32_t cell_code_cached(const GridDev& G, CellCache& cc, int x, int y, int z)
{
    if (unsigned((x ^ cc.x) | (y ^ cc.y) | (z ^ cc.z)) < 128u) return cc.code;
    const int cx = (x - G.cell_org[0]) >> 7, cy = (y - G.cell_org[1]) >> 7, cz = (z - G.cell_org[2]) >> 7;
    uint32_t code;
    if (unsigned(cx) >= unsigned(G.cell_dim[0]) || unsigned(cy) >= unsigned(G.cell_dim[1]) || unsigned(cz) >= unsigned(G.cell_dim[2]))
        code = CELL_KIND_TILE | (2u << CELL_DIM_SHIFT);
    else
        code = __ldg(&G.cells[(cx * G.cell_dim[1] + cy) * G.cell_dim[2] + cz]); // < 2^27 cells (bake.cpp): 32-bit index
    cc.x = x; cc.y = y; cc.z = z; cc.code = code;
    return code;
}
__device__ __forceinline__ int dim_of_code(uint32_t c)
{
    c &= CELL_DIM_MASK;
    return c == 0u ? 8 : (c == (1u << CELL_DIM_SHIFT) ? 128 : 4096);
}
__device__ __forceinline__ float majorant_cached(const GridDev& G, CellCache& cc, int x, int y, int z)
{
    const uint32_t c = cell_code_cached(G, cc, x, y, z);
    if (c & CELL_KIND_TILE) return __ldg(&G.tile_majorant[c & CELL_INDEX_MASK]);
#if VPT_SPLIT_MAJ
    return __ldg(&G.slot_majorant[(c & CELL_INDEX_MASK) * 4096u + maj_index(uint32_t(x & 127) >> 3, uint32_t(y & 127) >> 3, uint32_t(z & 127) >> 3)]);
#else
    return load_slot(G, slot_index(c, x, y, z)).majorant;
#endif
}
// Base voxel in a tile / background (its corners may still reach into leaves): 8 full getValue descents. Rare, and
// deliberately NOT inlined — the render loop has to stay small for the instruction cache.
__device__ VPT_COLD_INLINE float trilinear_general(const GridDev& G, int x, int y, int z, float u, float v, float w)
{
    const float v000 = value_at(G, x, y, z),         v001 = value_at(G, x, y, z + 1);
    const float v010 = value_at(G, x, y + 1, z),     v011 = value_at(G, x, y + 1, z + 1);
    const float v100 = value_at(G, x + 1, y, z),     v101 = value_at(G, x + 1, y, z + 1);
    const float v110 = value_at(G, x + 1, y + 1, z), v111 = value_at(G, x + 1, y + 1, z + 1);
    const float a00 = v000 + w * (v001 - v000), a01 = v010 + w * (v011 - v010);
    const float a10 = v100 + w * (v101 - v100), a11 = v110 + w * (v111 - v110);
    const float b0 = a00 + v * (a01 - a00), b1 = a10 + v * (a11 - a10);
    return b0 + u * (b1 - b0);
}
// trilinear() of vpt_device.cuh with the cached cell code for the base voxel.
__device__ __forceinline__ float trilinear_cached(const GridDev& G, CellCache& cc, float px, float py, float pz)
{
    const float fx = floorf(px), fy = floorf(py), fz = floorf(pz);
    const float u = px - fx, v = py - fy, w = pz - fz;
    const int x = int(fx), y = int(fy), z = int(fz);
    float v000, v001, v010, v011, v100, v101, v110, v111;
    bool fast = false;
    const uint32_t c = cell_code_cached(G, cc, x, y, z);
    if (!(c & CELL_KIND_TILE)) {
        const Slot s = load_slot(G, slot_index(c, x, y, z));
        if (!(s.code & SLOT_KIND_TILE)) {
            brick_corners(G, s.code, x, y, z, v000, v001, v010, v011, v100, v101, v110, v111);
            fast = true;
        }
    }
    if (!fast) return trilinear_general(G, x, y, z, u, v, w);
    const float a00 = v000 + w * (v001 - v000), a01 = v010 + w * (v011 - v010);
    const float a10 = v100 + w * (v101 - v100), a11 = v110 + w * (v111 - v110);
    const float b0 = a00 + v * (a01 - a00), b1 = a10 + v * (a11 - a10);
    return b0 + u * (b1 - b0);
}

enum : int { PH_MARCH = 0, PH_NEXT_PIXEL = 1, PH_BEGIN_SHADOW = 2, PH_AFTER_SHADOW = 3, PH_EXIT = 4 };

template <bool COUNT>
__device__ __forceinline__ bool flat_first_segment(March& m, const GridDev& G, CellCache& cc, bool& ray_done, Stats& st)
{
    ray_done = false;
    // RayMajorantIterator::next, first call: `if (time >= maxTime) return nullopt` comes BEFORE the first fetch
    if (m.t0 >= m.t1) { ray_done = true; return true; }
    m.majorant = majorant_cached(G, cc, m.vx, m.vy, m.vz);
    if (COUNT) ++st.majorant_fetches;
    m.seg_t0 = m.t0; m.seg_dmaj = m.majorant;
    return true;
}
// march_begin of vpt_device.cuh + the iterator's first majorant fetch + the first segment start.
// Returns false when the ray misses the index bbox (Volume::intersect -> nullopt) — no majorant is fetched then.
template <bool COUNT>
__device__ __forceinline__ bool flat_begin_ray(March& m, const GridDev& G, CellCache& cc, V3 o, V3 d, bool& ray_done, Stats& st)
{
    if (!march_begin(m, G, o, d)) return false;
    return flat_first_segment<COUNT>(m, G, cc, ray_done, st);
}
// Same with a precomputed direction (shadow rays) and the cached getDim lookup.
template <bool COUNT>
__device__ __forceinline__ bool flat_begin_ray_dir(March& m, const GridDev& G, CellCache& cc, V3 o, const RayDir& r, bool& ray_done, Stats& st)
{
    if (!march_begin_dir(m, G, o, r)) return false;
    const V3 s = m.at(m.t0);
    hdda_init(m, dim_of_code(cell_code_cached(G, cc, int(floorf(s.x)), int(floorf(s.y)), int(floorf(s.z)))));
    return flat_first_segment<COUNT>(m, G, cc, ray_done, st);
}

template <bool TRACE, bool COUNT>
__global__ void __launch_bounds__(VPT_BLOCK_THREADS) vpt_render_flat_kernel(const RenderArgs A)
{
    __shared__ float bb[1504];
    __shared__ double logtab[32];
    for (int i = threadIdx.x; i < 1504; i += blockDim.x) bb[i] = i < 1500 ? A.bb_table[i] : 0.0f;
    if (threadIdx.x < 32) logtab[threadIdx.x] = c_logf_table[threadIdx.x];
    __syncthreads();

    const SceneDev& S = A.scene;
    const GridDev& G = A.density;
    Stats st = {0, 0, 0, 0, 0, 0};
    unsigned long long draws_total = 0;
    Rng rng; rng.state = 0; rng.draws = 0;
    CellCache cc, ct;
    cc.reset(); ct.reset();
    March m;
    m.has_seg = false;
    V3 pos = mk(0, 0, 0), rd = mk(0, 0, 1), L = mk(0, 0, 0);
    float T_ray = 1.0f;
    unsigned int depth = 0;
    bool shadow = false, ray_done = false, pixel_open = false;
    int ph = PH_NEXT_PIXEL;
    unsigned long long j = 0;
    int x0 = 0, y0 = 0, sx = 0, sy = 0, x = 0, y = 0; // tile rect and the pixel inside it (x == sx: job exhausted)

    for (;;) {
        // ------------------------------------------------------------ ray setup (rare: ~once per 20 hot steps)
        if (ph != PH_MARCH) {
            if (ph == PH_NEXT_PIXEL) {
                if (pixel_open) { // film(y,x) += (imaging_ratio * L, 1)   worker.cpp:203-204
                    if (COUNT) ++st.pixel_samples;
                    if (TRACE) {
                        const size_t slot = size_t(j) * size_t(S.tw * S.th) + size_t(y) * size_t(S.tw) + size_t(x);
                        A.trace_xyz[slot * 3 + 0] = L.x; A.trace_xyz[slot * 3 + 1] = L.y; A.trace_xyz[slot * 3 + 2] = L.z;
                        A.trace_ndraws[slot] = rng.draws;
                    } else {
                        film_add(A.film + (size_t(y0 + y) * size_t(S.W) + size_t(x0 + x)) * 4, S.imaging_ratio * L.x, S.imaging_ratio * L.y, S.imaging_ratio * L.z);
                    }
                    pixel_open = false;
                }
                for (;;) { // next pixel of the job (row-major, worker.cpp:109-110), next job when exhausted
                    if (++x >= sx) { x = 0; ++y; }
                    if (y >= sy) {
                        if (COUNT) draws_total += rng.draws;
                        j = atomicAdd(A.queue, 1ULL); // TileProvider::next(): m_job_idx.fetch_add(1)
                        if (j >= A.n_jobs) { ph = PH_EXIT; break; }
                        uint64_t jid;
                        if (TRACE) jid = A.trace_jids[j];
                        else {
                            const uint64_t k = j / S.ntiles, tile_k = j - k * S.ntiles;
                            jid = (uint64_t(A.first_wave - 1) + k * A.wave_stride) * S.ntiles + tile_k;
                        }
                        const uint64_t tile = jid % S.ntiles;
                        const int64_t tx0 = int64_t(tile % uint64_t(S.ntx)) * S.tw, ty0 = int64_t(tile / uint64_t(S.ntx)) * S.th;
                        x0 = int(tx0); y0 = int(ty0);
                        sx = int(min(S.W - tx0, S.tw)); sy = int(min(S.H - ty0, S.th));
                        rng.begin_job(S.seed, jid);
                        x = 0; y = 0;
                    }
                    if (S.single_pixel_enabled && (int64_t(x0 + x) != S.single_pixel_x || int64_t(y0 + y) != S.single_pixel_y)) continue;
                    break;
                }
                if (ph == PH_EXIT) break;
                // camera path start   worker.cpp:121-128, camera.hpp:14-23
                const float j0 = rng.uniform(), j1 = rng.uniform();
                const float jm = S.use_jitter ? 0.5f : 0.0f;
                const float rx = (float(x0 + x) + 0.5f) + j0 * jm, ry = (float(y0 + y) + 0.5f) + j1 * jm;
                rd = normalized(mk(sum3(S.cam_m[0] * rx, S.cam_m[1] * ry, S.cam_m[2] * 0.0f) + S.cam_t[0],
                                   sum3(S.cam_m[3] * rx, S.cam_m[4] * ry, S.cam_m[5] * 0.0f) + S.cam_t[1],
                                   sum3(S.cam_m[6] * rx, S.cam_m[7] * ry, S.cam_m[8] * 0.0f) + S.cam_t[2]));
                pos = mk(S.cam_pos[0], S.cam_pos[1], S.cam_pos[2]);
                L = mk(0.0f, 0.0f, 0.0f);
                depth = 0; shadow = false; pixel_open = true;
            } else if (ph == PH_BEGIN_SHADOW) { // sample_Ld   worker.cpp:59-66
                shadow = true; T_ray = 1.0f;
                if (COUNT) ++st.shadow_rays;
            } else { // PH_AFTER_SHADOW: worker.cpp:88-89 then :175-183
                const V3 wi = mk(S.dist_wi[0], S.dist_wi[1], S.dist_wi[2]);
                const float k = henyey_greenstein(dot3(rd, wi), S.g) * T_ray;
                L.x += k * S.dist_Li[0]; L.y += k * S.dist_Li[1]; L.z += k * S.dist_Li[2];
                const float u0 = rng.uniform(), u1 = rng.uniform();
                rd = sample_henyey_greenstein(rd, u0, u1, S.g);
                shadow = false;
                depth += 2; // `depth++` at the scatter (worker.cpp:169) and `++depth` of the bounce loop (:130)
            }
            m.has_seg = false;
            if (!shadow && depth >= S.max_depth) { // bounce loop ran out of depth: not terminated -> infinite light (:198-200)
                L.x += S.inf_light[0]; L.y += S.inf_light[1]; L.z += S.inf_light[2];
                ph = PH_NEXT_PIXEL;
            } else {
                const V3 dir = shadow ? mk(S.dist_wi[0], S.dist_wi[1], S.dist_wi[2]) : rd;
                if (flat_begin_ray<COUNT>(m, G, cc, pos, dir, ray_done, st)) ph = PH_MARCH;
                else if (shadow) ph = PH_AFTER_SHADOW; // the shadow ray misses the volume: T stays 1
                else { L.x += S.inf_light[0]; L.y += S.inf_light[1]; L.z += S.inf_light[2]; ph = PH_NEXT_PIXEL; }
            }
        }
        if (ph != PH_MARCH) continue;

        // ------------------------------------------------------------ hot block 1: one HDDA step   volume.cpp:53-71
        if (!ray_done && !m.has_seg) {
            bool closed;
            if (!hdda_step(m)) { m.seg_t1 = m.t1; closed = true; }
            else {
                const V3 p = m.at(m.t0 + 1.0001f);
                hdda_update(m, dim_of_code(cell_code_cached(G, cc, int(floorf(p.x)), int(floorf(p.y)), int(floorf(p.z)))));
                m.majorant = majorant_cached(G, cc, m.vx, m.vy, m.vz);
                if (COUNT) ++st.majorant_fetches;
                closed = m.majorant != m.seg_dmaj;
                if (closed) m.seg_t1 = m.t0;
            }
            if (closed) {
                if (m.seg_dmaj > 0.0f) m.has_seg = true;
                else if (m.t0 >= m.t1) ray_done = true;          // sampler drops the empty segment, iterator.next() -> nullopt
                else { m.seg_t0 = m.t0; m.seg_dmaj = m.majorant; } // ... or starts the next segment
            }
        }
        // ------------------------------------------------------------ hot block 2: one free-flight sample   majorant_transmittance_sampler.cpp:41-77
        if (m.has_seg) {
            const float sigma_maj = m.seg_dmaj * S.sigma_t;
            const float dt_m = -vptm::logf_exact(1.0f - rng.uniform(), logtab) / sigma_maj;
            const float t = m.seg_t0 + dt_m / m.scale;
            if (t < m.seg_t1) {
                m.seg_t0 = t;
                const V3 p = m.at(t);
                if (COUNT) ++st.density_evals;
                const float density = trilinear_cached(G, cc, p.x, p.y, p.z);
                if (density > 0.0f) {
                    if (shadow) { // ratio tracking + Russian roulette   worker.cpp:68-84
                        const float sigma_n = fmaxf(0.0f, sigma_maj - S.sigma_t * density);
                        T_ray *= sigma_n / sigma_maj;
                        if (T_ray <= 0.05f) {
                            if (rng.uniform() < 0.75f) T_ray = 0.0f;
                            else T_ray /= 0.25f;
                        }
                        if (T_ray <= 0.0f) { T_ray = 0.0f; ph = PH_AFTER_SHADOW; }
                    } else { // worker.cpp:148-188
                        const V3 point = index_to_world(G, p);
                        const float p_a = (S.sigma_a * density) / sigma_maj;
                        const float p_s = (S.sigma_s * density) / sigma_maj;
                        const float p_n = fmaxf(1.0f - p_a - p_s, 0.0f);
                        if (S.has_temperature) {
                            const V3 tc = world_to_index(A.temperature, point);
                            if (COUNT) ++st.temperature_evals;
                            const float temp_adim = trilinear_cached(A.temperature, ct, tc.x, tc.y, tc.z);
                            const float temp_K = temp_adim * S.temperature_scale + S.temperature_offset;
                            const V3 e = blackbody_xyz(bb, temp_K);
                            const float k = p_a * S.le_scale;
                            L.x += k * e.x; L.y += k * e.y; L.z += k * e.z;
                        }
                        float u = rng.uniform() * (((0.0f + p_n) + p_a) + p_s); // sample_discrete  random.hpp:31-47
                        u -= p_n;
                        if (u > 0.0f) {
                            u -= p_a;
                            if (u <= 0.0f) ph = PH_NEXT_PIXEL; // absorption: terminated, no infinite light
                            else {                               // scatter (also the fallback choice)
                                if (COUNT) ++st.scatter_events;
                                pos = point;
                                if (S.dist_light_on) ph = PH_BEGIN_SHADOW;
                                else { T_ray = 0.0f; ph = PH_AFTER_SHADOW; } // sample_Ld returns 0 before any draw (:57)
                            }
                        }
                    }
                }
            } else { // overshoot: drop the segment, iterator.next()
                m.has_seg = false;
                if (m.t0 >= m.t1) ray_done = true;
                else { m.seg_t0 = m.t0; m.seg_dmaj = m.majorant; }
            }
        }
        if (ray_done && ph == PH_MARCH) { // the ray left the volume
            if (shadow) ph = PH_AFTER_SHADOW;
            else { L.x += S.inf_light[0]; L.y += S.inf_light[1]; L.z += S.inf_light[2]; ph = PH_NEXT_PIXEL; }
        }
    }
    if (COUNT) {
        atomicAdd(&A.counters->pixel_samples, (unsigned long long)st.pixel_samples);
        atomicAdd(&A.counters->density_evals, (unsigned long long)st.density_evals);
        atomicAdd(&A.counters->temperature_evals, (unsigned long long)st.temperature_evals);
        atomicAdd(&A.counters->majorant_fetches, (unsigned long long)st.majorant_fetches);
        atomicAdd(&A.counters->rng_draws, draws_total);
        atomicAdd(&A.counters->scatter_events, (unsigned long long)st.scatter_events);
        atomicAdd(&A.counters->shadow_rays, (unsigned long long)st.shadow_rays);
    }
}

} // namespace vptd
