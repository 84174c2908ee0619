// vpt_kernels.cu — the path-integrator megakernel and its small companions (sm_100a).
//
// Replaces vpt::run (src/worker.cpp:92-208) + sample_Ld (:52-90) + TileProvider::next
// (src/tile_provider.cpp:27-67). One thread = one (tile, wave) job, so the job's pcg32_fast stream is consumed
// in the reference's order (pixels row-major inside the tile, worker.cpp:107-110) and per-path radiance is
// comparable with the CPU oracle draw for draw. Jobs are handed out by a device-side atomic counter
// (the reference's m_job_idx.fetch_add, tile_provider.cpp:28); the film is accumulated with float4-wide
// reductions in L2, which replaces the per-tile wave handshake (tile_provider.cpp:40-60).
#include <algorithm>
#include <cstdio>

#include "vpt_device.cuh"
#include "vpt_kernels.hpp"

namespace vptd {

__constant__ float c_cie[471 * 3];
__constant__ float c_lambda5[471];                                   // float(pow(double(lambda_m), 5)) per wavelength, from the HOST's libm (spectral.cpp:15)
__constant__ double c_logf_table[32] = VPT_LOGF_TABLE_INIT;
__constant__ unsigned long long c_exp2f_table[32] = VPT_EXP2F_TABLE_INIT;   // glibc __exp2f_data.tab (libm_exact.cuh)
__constant__ double c_powf_log2_table[32] = VPT_POWF_LOG2_TABLE_INIT;       // glibc __powf_log2_data.tab

// ---------------------------------------------------------------- blackbody  src/precompute_blackbody.cpp:24-52
// Bit-exact with the host: lambda^5 is the host libm's own value per wavelength (471 of them, uploaded at vptc_create), expf is glibc's
// scheme (libm_exact.cuh, exhaustively checked), everything else is single IEEE fp32 operations in the reference's order.
__device__ __noinline__ V3 spectrum_to_xyz_direct(float T) // T >= 49900 K: direct 471-term integration (spectral.hpp:62-75, spectral.cpp:7-20)
{
    float acc[3];
    for (int ch = 0; ch < 3; ++ch) {
        float integral = 0.0f;
        for (int i = 0; i < 471; ++i) {
            const float lambda_m = float(360 + i) * 1e-9f;
            const float c = 299792458.f, h = 6.62606957e-34f, kb = 1.3806488e-23f;
            const float num = 2 * h * c * c;
            const float lambda5 = c_lambda5[i];
            const float e = vptm::expf_exact((h * c) / (lambda_m * kb * T), reinterpret_cast<const uint64_t*>(c_exp2f_table)); // argument in (0, 0.81] for T >= 49900
            const float den = lambda5 * (e - 1);
            integral += c_cie[i * 3 + ch] * (num / den);
        }
        acc[ch] = integral;
    }
    const float Y = 106.856895f;
    return mk(acc[0] / Y, acc[1] / Y, acc[2] / Y);
}
__device__ VPT_COLD_INLINE V3 blackbody_xyz(const float* __restrict__ bb, float T)
{
    if (!isfinite(T)) { const float n = __int_as_float(0x7fc00000); return mk(n, n, n); }
    if (T <= 0.0f) return mk(0.0f, 0.0f, 0.0f);
    if (T >= 49900.0f) return spectrum_to_xyz_direct(T);
    int dn = int(T / 100.0f);
    while (T <= float(dn - 2) * 100.0f) --dn;  // idx_to_temp(dn - 1)
    while (T >= float(dn) * 100.0f) ++dn;      // idx_to_temp(dn + 1)
    const float dn_temp = float(dn - 1) * 100.0f;
    const float* a = bb + 3 * dn;
    if (T == dn_temp) return mk(a[0], a[1], a[2]);
    const float t = (T - dn_temp) / 100.0f;
    return mk(a[0] + (a[3] - a[0]) * t, a[1] + (a[4] - a[1]) * t, a[2] + (a[5] - a[2]) * t);
}

// ---------------------------------------------------------------- sample_Ld  src/worker.cpp:52-90
template <bool COUNT>
__device__ __forceinline__ V3 sample_Ld(const RenderArgs& A, const double* logtab, Rng& rng, V3 pos, V3 w, Stats& st)
{
    const SceneDev& S = A.scene;
    if (!S.dist_light_on) return mk(0.0f, 0.0f, 0.0f);
    const V3 wi = mk(S.dist_wi[0], S.dist_wi[1], S.dist_wi[2]);
    float T_ray = 1.0f;
    if (COUNT) ++st.shadow_rays;
    March m;
    if (march_begin(m, A.density, pos, wi)) {
        Collision c;
        while (sampler_next<COUNT>(m, A.density, rng, S.sigma_t, logtab, c, st)) {
            const float sigma_n = fmaxf(0.0f, c.sigma_maj - S.sigma_t * c.density);
            T_ray *= sigma_n / c.sigma_maj;
            if (T_ray <= 0.05f) {
                if (rng.uniform() < 0.75f) T_ray = 0.0f;
                else T_ray /= 0.25f; // 1 - q
            }
            if (T_ray <= 0.0f) return mk(0.0f, 0.0f, 0.0f);
        }
    }
    const float p = henyey_greenstein(dot3(w, wi), S.g);
    const float k = p * T_ray;
    return mk(k * S.dist_Li[0], k * S.dist_Li[1], k * S.dist_Li[2]);
}

// ---------------------------------------------------------------- one camera path  src/worker.cpp:121-200
template <bool COUNT>
__device__ __forceinline__ V3 trace_pixel(const RenderArgs& A, const float* __restrict__ bb, const double* logtab, Rng& rng, int64_t px, int64_t py, Stats& st)
{
    const SceneDev& S = A.scene;
    const float j0 = rng.uniform(), j1 = rng.uniform();
    const float jm = S.use_jitter ? 0.5f : 0.0f;
    // Camera::generate_ray  camera.hpp:14-23
    const float rx = (float(px) + 0.5f) + j0 * jm, ry = (float(py) + 0.5f) + j1 * jm;
    V3 rd = normalized(mk(sum3(S.cam_m[0] * rx, S.cam_m[1] * ry, S.cam_m[2] * 0.0f) + S.cam_t[0],
                          sum3(S.cam_m[3] * rx, S.cam_m[4] * ry, S.cam_m[5] * 0.0f) + S.cam_t[1],
                          sum3(S.cam_m[6] * rx, S.cam_m[7] * ry, S.cam_m[8] * 0.0f) + S.cam_t[2]));
    V3 ro = mk(S.cam_pos[0], S.cam_pos[1], S.cam_pos[2]);

    V3 L = mk(0.0f, 0.0f, 0.0f);
    bool terminated = false;
    for (unsigned int depth = 0; depth < S.max_depth; ++depth) {
        bool scattered = false;
        March m;
        if (!march_begin(m, A.density, ro, rd)) break;
        Collision c;
        while (sampler_next<COUNT>(m, A.density, rng, S.sigma_t, logtab, c, st)) {
            const float p_a = (S.sigma_a * c.density) / c.sigma_maj;
            const float p_s = (S.sigma_s * c.density) / c.sigma_maj;
            const float p_n = fmaxf(1.0f - p_a - p_s, 0.0f);
            if (S.has_temperature) {
                const V3 tc = world_to_index(A.temperature, c.point);
                if (COUNT) ++st.temperature_evals;
                const float temp_adim = trilinear(A.temperature, tc.x, tc.y, tc.z);
                const float temp_K = temp_adim * S.temperature_scale + S.temperature_offset;
                const V3 e = blackbody_xyz(bb, temp_K);
                const float k = p_a * S.le_scale;
                L.x += k * e.x; L.y += k * e.y; L.z += k * e.z;
            }
            // sample_discrete({Null p_n, Absorption p_a, Scatter p_s}, u)  random.hpp:31-47
            float u = rng.uniform() * (((0.0f + p_n) + p_a) + p_s);
            int event; // 0 null, 1 absorption, 2 scatter
            u -= p_n;
            if (u <= 0.0f) event = 0;
            else { u -= p_a; if (u <= 0.0f) event = 1; else event = 2; } // third choice is also the fallback
            if (event == 0) continue;
            if (event == 2) {
                if (depth++ >= S.max_depth) { terminated = true; break; }
                if (COUNT) ++st.scatter_events;
                const V3 ld = sample_Ld<COUNT>(A, logtab, rng, c.point, rd, st);
                L.x += ld.x; L.y += ld.y; L.z += ld.z;
                const float u0 = rng.uniform(), u1 = rng.uniform();
                rd = sample_henyey_greenstein(rd, u0, u1, S.g);
                ro = c.point;
                scattered = true;
                break;
            }
            terminated = true; // absorption
            break;
        }
        if (!scattered) break;
    }
    if (!terminated) { L.x += S.inf_light[0]; L.y += S.inf_light[1]; L.z += S.inf_light[2]; }
    if (COUNT) ++st.pixel_samples;
    return L;
}

__device__ __forceinline__ void film_add(float* film, float x, float y, float z)
{
    // one 16-byte reduction in L2: red.global.add.v4.f32 (sm_90+)
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(film), "f"(x), "f"(y), "f"(z), "f"(1.0f) : "memory");
}

} // namespace vptd
#include "vpt_render_flat.cuh"  // variant 1: the flat state-machine megakernel
#include "vpt_render_sched.cuh" // variant 2: flat + warp-level stage scheduling
#include "vpt_render_pool.cuh"  // variant 3: per-warp job pool in shared memory
namespace vptd {

// ---------------------------------------------------------------- variant 0: the first, nested megakernel (kept for the ncu comparison)
template <bool TRACE, bool COUNT>
__global__ void __launch_bounds__(VPT_BLOCK_THREADS) vpt_render_kernel(const RenderArgs A)
{
    __shared__ float bb[1504]; // precompute_blackbody table: 500 x XYZ (+1 padding entry, see oracle note)
    __shared__ double logtab[32]; // glibc logf's {1/c, log c} table (libm_exact.cuh)
    for (int i = threadIdx.x; i < 1504; i += blockDim.x) bb[i] = i < 1500 ? A.bb_table[i] : 0.0f;
    if (threadIdx.x < 32) logtab[threadIdx.x] = c_logf_table[threadIdx.x];
    __syncthreads();

    const SceneDev& S = A.scene;
    Stats st = {0, 0, 0, 0, 0, 0};
    unsigned long long draws_total = 0;
    Rng rng;
    for (;;) {
        // TileProvider::next(): job = fetch_add(1); wave = 1 + job / ntiles; tile = job % ntiles
        const unsigned long long j = atomicAdd(A.queue, 1ULL);
        if (j >= A.n_jobs) break;
        uint64_t jid;
        if (TRACE) jid = A.trace_jids[j];
        else {
            const uint64_t k = j / S.ntiles, tile = j - k * S.ntiles;
            jid = (uint64_t(A.first_wave - 1) + k * A.wave_stride) * S.ntiles + tile;
        }
        const uint64_t tile = jid % S.ntiles;
        // compute_tile_rect  tile_provider.cpp:95-105
        const int64_t x0 = int64_t(tile % uint64_t(S.ntx)) * S.tw, y0 = int64_t(tile / uint64_t(S.ntx)) * S.th;
        const int sx = int(min(S.W - x0, S.tw)), sy = int(min(S.H - y0, S.th));
        rng.begin_job(S.seed, jid);
        for (int y = 0; y < sy; ++y) {
            for (int x = 0; x < sx; ++x) {
                const int64_t px = x0 + x, py = y0 + y;
                if (S.single_pixel_enabled && (px != S.single_pixel_x || py != S.single_pixel_y)) continue;
                const V3 L = trace_pixel<COUNT>(A, bb, logtab, rng, px, py, st);
                if (TRACE) {
                    const size_t slot = size_t(j) * size_t(S.tw * S.th) + size_t(y) * size_t(S.tw) + size_t(x);
                    A.trace_xyz[slot * 3 + 0] = L.x; A.trace_xyz[slot * 3 + 1] = L.y; A.trace_xyz[slot * 3 + 2] = L.z;
                    A.trace_ndraws[slot] = rng.draws;
                } else {
                    film_add(A.film + (size_t(py) * size_t(S.W) + size_t(px)) * 4, S.imaging_ratio * L.x, S.imaging_ratio * L.y, S.imaging_ratio * L.z);
                }
            }
        }
        if (COUNT) draws_total += rng.draws;
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

// ---------------------------------------------------------------- parity hooks
__global__ void vpt_segments_kernel(GridDev G, const float* origins, const float* dirs, size_t n, uint32_t max_segs, float* segs, uint32_t* n_segs)
{
    const size_t r = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (r >= n) return;
    Stats st = {0, 0, 0, 0, 0, 0};
    March m;
    uint32_t cnt = 0;
    if (march_begin(m, G, mk(origins[3 * r], origins[3 * r + 1], origins[3 * r + 2]), mk(dirs[3 * r], dirs[3 * r + 1], dirs[3 * r + 2]))) {
        while (iterator_next<false>(m, G, st)) {
            if (cnt < max_segs) { float* o = segs + (r * max_segs + cnt) * 3; o[0] = m.seg_t0; o[1] = m.seg_t1; o[2] = m.seg_dmaj; }
            ++cnt;
        }
    }
    n_segs[r] = cnt;
}

__global__ void vpt_sample_kernel(GridDev G, const float* xyz, size_t n, float* values, uint32_t* dims, float* majorants)
{
    const size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    const float px = xyz[3 * i], py = xyz[3 * i + 1], pz = xyz[3 * i + 2];
    if (values) values[i] = trilinear(G, px, py, pz);
    const int x = int(floorf(px)), y = int(floorf(py)), z = int(floorf(pz));
    if (dims) dims[i] = uint32_t(hdda_dim(G, x, y, z));
    if (majorants) majorants[i] = majorant_at(G, x, y, z);
}

// ---------------------------------------------------------------- film_to_image  src/main.cpp:12-24, color.hpp:8-30
// Byte-exact with the host: the sRGB OETF's powf(x, 1/2.4f) is glibc's scheme (libm_exact.cuh: zero mismatches over every float of
// (0.0031308, FLT_MAX]); +inf stays +inf (clamps to 255), NaN (a w = 0 pixel: 0/0) has an undefined cast in the reference and is written as 0.
__global__ void vpt_resolve_kernel(const float4* film, size_t n, uint8_t* rgb)
{
    const size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    const float4 f = film[i];
    const float X = f.x / f.w, Y = f.y / f.w, Z = f.z / f.w;
    const float lin[3] = {sum3(3.240479f * X, -1.537150f * Y, -0.498535f * Z), sum3(-0.969256f * X, 1.875991f * Y, 0.041556f * Z),
                          sum3(0.055648f * X, -0.204043f * Y, 1.057311f * Z)};
    for (int c = 0; c < 3; ++c) {
        const float x = lin[c];
        float s;
        if (x <= 0.0031308f) s = 12.92f * x;
        else {
            const float p = x <= FLT_MAX ? vptm::powf_exact(x, 1.0f / 2.4f, reinterpret_cast<const uint64_t*>(c_exp2f_table), c_powf_log2_table) : x; // inf, NaN pass through
            s = 1.055f * p - 0.055f;
        }
        s = fminf(fmaxf(s, 0.0f), 1.0f) * 255.0f; // NaN -> fmaxf gives 0
        rgb[i * 3 + c] = (unsigned char)(s);
    }
}

// Parity hook: the device's blackbody_radiation_xyz (table in global memory here; the render kernels keep it in shared memory).
__global__ void vpt_blackbody_kernel(const float* __restrict__ bb, const float* temps, size_t n, float* xyz)
{
    const size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x;
    if (i >= n) return;
    const V3 v = blackbody_xyz(bb, temps[i]);
    xyz[3 * i] = v.x; xyz[3 * i + 1] = v.y; xyz[3 * i + 2] = v.z;
}

// ---------------------------------------------------------------- multi-GPU film combine (vptc_group_wait, wave sharding)
// dst += src[0] + ... + src[k-1], float4 per pixel. `dst` lives on this device; the sources are the OTHER devices' films, read in place through
// peer mappings (NVLink / NVSwitch): the sum and the transfer are one pass, nothing is staged. HBM/NVLink bound, k + 2 float4 per pixel.
__global__ void vpt_film_reduce_kernel(float4* __restrict__ dst, FilmSources src, int k, size_t n)
{
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
        float4 a = dst[i];
        for (int j = 0; j < k; ++j) {
            const float4 b = __ldcs(src.p[j] + i); // streamed once: no reuse
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        dst[i] = a;
    }
}

// ---------------------------------------------------------------- launchers
cudaError_t upload_cie_table(const float* cie471x3, const float* lambda5_471)
{
    const cudaError_t e = cudaMemcpyToSymbol(c_cie, cie471x3, sizeof(float) * 471 * 3);
    return e != cudaSuccess ? e : cudaMemcpyToSymbol(c_lambda5, lambda5_471, sizeof(float) * 471);
}

template <bool TRACE, bool COUNT>
static void launch_variant(const RenderArgs& args, int variant, int grid_blocks, cudaStream_t stream)
{
    if (variant == 0) vpt_render_kernel<TRACE, COUNT><<<grid_blocks, VPT_BLOCK_THREADS, 0, stream>>>(args);
    else if (variant == 1) vpt_render_flat_kernel<TRACE, COUNT><<<grid_blocks, VPT_BLOCK_THREADS, 0, stream>>>(args);
    else if (variant == 2) vpt_render_sched_kernel<TRACE, COUNT><<<grid_blocks, VPT_BLOCK_THREADS, 0, stream>>>(args);
    else vpt_render_pool_kernel<TRACE, COUNT><<<grid_blocks, POOL_THREADS, pool_smem_bytes<TRACE, COUNT>(), stream>>>(args);
}
cudaError_t launch_render_stats(const RenderArgs& args, int grid_blocks, cudaStream_t stream) // the STATS instantiation of the default kernel
{
    vpt_render_sched_kernel<false, false, true><<<grid_blocks, VPT_BLOCK_THREADS, 0, stream>>>(args);
    return cudaGetLastError();
}
cudaError_t launch_render_ordered(const RenderArgs& args, int ord, int grid_blocks, cudaStream_t stream) // tile order: 1 = learning, 2 = ordered
{
    if (ord == 1) vpt_render_sched_kernel<false, false, false, 1><<<grid_blocks, VPT_BLOCK_THREADS, 0, stream>>>(args);
    else vpt_render_sched_kernel<false, false, false, 2><<<grid_blocks, VPT_BLOCK_THREADS, 0, stream>>>(args);
    return cudaGetLastError();
}
int render_ordered_blocks_per_sm(int ord)
{
    int n = 0;
    const cudaError_t e = ord == 1 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vpt_render_sched_kernel<false, false, false, 1>, VPT_BLOCK_THREADS, 0)
                                   : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vpt_render_sched_kernel<false, false, false, 2>, VPT_BLOCK_THREADS, 0);
    return e == cudaSuccess ? n : 0;
}
int render_stats_blocks_per_sm()
{
    int n = 0;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vpt_render_sched_kernel<false, false, true>, VPT_BLOCK_THREADS, 0) == cudaSuccess ? n : 0;
}
// Adaptive rendering: per-tile relative standard error of the mean luminance from the film (sum y in .y, count in .w) and the second-moment
// plane: err[t] = sqrt(mean over the tile's pixels of s^2 / n) / (mean over the pixels of y-bar), s^2 the unbiased sample variance.
// One block per tile; tiles clipped at the frame edge use their own pixel count.
__global__ void vpt_tile_error_kernel(const float4* film, const float* sum_y2, long long W, long long H, long long tw, long long th, long long ntx, unsigned long long ntiles, float* err)
{
    const unsigned long long t = blockIdx.x;
    if (t >= ntiles) return;
    const long long x0 = (long long)(t % (unsigned long long)ntx) * tw, y0 = (long long)(t / (unsigned long long)ntx) * th;
    const long long sx = min(W - x0, tw), sy = min(H - y0, th);
    double se2 = 0.0, mean = 0.0;
    for (long long i = threadIdx.x; i < sx * sy; i += blockDim.x) {
        const size_t p = size_t(y0 + i / sx) * size_t(W) + size_t(x0 + i % sx);
        const float4 f = film[p];
        const double n = f.w, m = n > 0.0 ? f.y / n : 0.0;
        const double var = n > 1.0 ? fmax(0.0, (double(sum_y2[p]) - n * m * m) / (n - 1.0)) : 1e30; // fewer than 2 samples: not converged
        se2 += n > 0.0 ? var / n : 1e30;
        mean += m;
    }
    __shared__ double s_a[128], s_b[128];
    s_a[threadIdx.x] = se2; s_b[threadIdx.x] = mean;
    __syncthreads();
    for (int k = blockDim.x / 2; k > 0; k >>= 1) {
        if (int(threadIdx.x) < k) { s_a[threadIdx.x] += s_a[threadIdx.x + k]; s_b[threadIdx.x] += s_b[threadIdx.x + k]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double npx = double(sx * sy), rms_se = sqrt(s_a[0] / npx), mu = s_b[0] / npx;
        err[t] = mu > 0.0 ? float(fmin(rms_se / mu, 1e30)) : (rms_se > 0.0 ? 1e30f : 0.0f); // a black, noise-free tile is converged
    }
}
cudaError_t launch_tile_error(const float* film, const float* sum_y2, long long W, long long H, long long tw, long long th, long long ntx, unsigned long long ntiles, float* err, cudaStream_t s)
{
    if (ntiles == 0) return cudaSuccess;
    vpt_tile_error_kernel<<<unsigned(ntiles), 128, 0, s>>>(reinterpret_cast<const float4*>(film), sum_y2, W, H, tw, th, ntx, ntiles, err);
    return cudaGetLastError();
}
cudaError_t launch_render(const RenderArgs& args, int variant, bool trace, bool count, int grid_blocks, cudaStream_t stream)
{
    if (trace) { if (count) launch_variant<true, true>(args, variant, grid_blocks, stream); else launch_variant<true, false>(args, variant, grid_blocks, stream); }
    else { if (count) launch_variant<false, true>(args, variant, grid_blocks, stream); else launch_variant<false, false>(args, variant, grid_blocks, stream); }
    return cudaGetLastError();
}
template <bool TRACE, bool COUNT>
static int occupancy_variant(int variant)
{
    int n = 0;
    if (variant == 3) { // the pool needs more than the default 48 KB of dynamic shared memory: opt in (idempotent)
        if (cudaFuncSetAttribute(vpt_render_pool_kernel<TRACE, COUNT>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(pool_smem_bytes<TRACE, COUNT>())) != cudaSuccess) return 0;
        return cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vpt_render_pool_kernel<TRACE, COUNT>, POOL_THREADS, pool_smem_bytes<TRACE, COUNT>()) == cudaSuccess ? n : 0;
    }
    const cudaError_t e = variant == 0   ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vpt_render_kernel<TRACE, COUNT>, VPT_BLOCK_THREADS, 0)
                          : variant == 1 ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vpt_render_flat_kernel<TRACE, COUNT>, VPT_BLOCK_THREADS, 0)
                                         : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, vpt_render_sched_kernel<TRACE, COUNT>, VPT_BLOCK_THREADS, 0);
    return e == cudaSuccess ? n : 0;
}
int render_jobs_per_block(int variant) { return variant == 3 ? POOL_JOBS_PER_BLOCK : VPT_BLOCK_THREADS; }
int render_blocks_per_sm(int variant, bool trace, bool count)
{
    if (trace) return count ? occupancy_variant<true, true>(variant) : occupancy_variant<true, false>(variant);
    return count ? occupancy_variant<false, true>(variant) : occupancy_variant<false, false>(variant);
}
cudaError_t launch_segments(const GridDev& G, const float* origins, const float* dirs, size_t n, uint32_t max_segs, float* segs, uint32_t* n_segs, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    vpt_segments_kernel<<<unsigned((n + 127) / 128), 128, 0, s>>>(G, origins, dirs, n, max_segs, segs, n_segs);
    return cudaGetLastError();
}
cudaError_t launch_sample(const GridDev& G, const float* xyz, size_t n, float* values, uint32_t* dims, float* majorants, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    vpt_sample_kernel<<<unsigned((n + 127) / 128), 128, 0, s>>>(G, xyz, n, values, dims, majorants);
    return cudaGetLastError();
}
cudaError_t launch_film_reduce(float* dst, const float* const* srcs, int k, size_t n_pixels, cudaStream_t s)
{
    if (n_pixels == 0 || k <= 0) return cudaSuccess;
    FilmSources fs{};
    for (int j = 0; j < k && j < FILM_REDUCE_MAX_SRC; ++j) fs.p[j] = reinterpret_cast<const float4*>(srcs[j]);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const unsigned blocks = unsigned(std::min<size_t>((n_pixels + 255) / 256, size_t(sms) * 8));
    vpt_film_reduce_kernel<<<blocks, 256, 0, s>>>(reinterpret_cast<float4*>(dst), fs, std::min(k, FILM_REDUCE_MAX_SRC), n_pixels);
    return cudaGetLastError();
}
cudaError_t launch_blackbody(const float* bb_table, const float* temps, size_t n, float* xyz, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    vpt_blackbody_kernel<<<unsigned((n + 127) / 128), 128, 0, s>>>(bb_table, temps, n, xyz);
    return cudaGetLastError();
}
cudaError_t launch_resolve(const float* film, size_t n_pixels, uint8_t* rgb, cudaStream_t s)
{
    if (n_pixels == 0) return cudaSuccess;
    vpt_resolve_kernel<<<unsigned((n_pixels + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4*>(film), n_pixels, rgb);
    return cudaGetLastError();
}

} // namespace vptd
