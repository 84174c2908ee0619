// vpt_render_pool.cuh — variant 3: the megakernel over a per-warp JOB POOL in shared memory.
//
// Evidence (profiles/r01_dense512_16waves_sched72p_full.summary.txt, profiles/r01_bench_launches.csv): the sched
// kernel (variant 2) is instruction-issue bound and issues its warp instructions with 12-13 of 32 lanes active.
// With exactly 32 jobs per warp, every lane that waits for a rare stage (ray setup, exact evaluation) is a lane the
// hot stage runs without: HOT averages ~17 lanes, EVAL 16, SETUP 12. The only way to fill the lanes while keeping
// one serial RNG stream per (tile, wave) job (the reference's bit-exact contract, worker.cpp:107-110) is to give a
// warp MORE JOBS THAN LANES. Here a warp owns VPT_POOL_SLOTS (64) job slots whose complete path state lives in
// shared memory as structure-of-arrays (`pool[word][slot]`, 40 words = 160 B per job). Each iteration the warp
//   1. takes a census of the slots' stages (2 LDS + 6 ballots),
//   2. picks the stage that fills most lanes,
//   3. compacts the slots queued for it onto lanes (ballot + popc prefix ranks scattered through a 64-byte list),
//   4. loads the words that stage needs, runs it (HOT: a few passes), stores what changed.
// This is the wavefront design (stage queues, SoA state, ballot/prefix-sum compaction) at warp scope, with the
// state streaming through shared memory instead of HBM — tools/wavefront_bound.cu shows the HBM form is capped
// at 93 M pixel-samples/s before any arithmetic. Per-job arithmetic and its order are those of variants 1/2, so the
// film stays bit-identical (tests/test_gpu_parity.py::test_kernel_variants_agree).
#pragma once
#include "vpt_render_sched.cuh"

namespace vptd {

#ifndef VPT_POOL_SLOTS
#define VPT_POOL_SLOTS 64 // job slots per warp (33..64)
#endif
#ifndef VPT_POOL_WARPS
#define VPT_POOL_WARPS 10 // warps per block; 10 x 64 slots x 160 B + tables = 107 KB -> 2 blocks per SM
#endif
#ifndef VPT_POOL_MIN_BLOCKS
#define VPT_POOL_MIN_BLOCKS 2
#endif
constexpr int POOL_THREADS = VPT_POOL_WARPS * 32;
constexpr int POOL_JOBS_PER_BLOCK = VPT_POOL_WARPS * VPT_POOL_SLOTS;

// slot words. Mutable marching state first (what HOT stores back), then per-ray constants, then the cold path state.
enum : int {
    W_FLAGS = 0, W_RLO, W_RHI, W_T0, W_VX, W_VY, W_VZ, W_NX, W_NY, W_NZ, W_MAJ, W_ST0, W_ST1, W_SDM, W_PENDU,
    W_EX, W_EY, W_EZ, W_DX, W_DY, W_DZ, W_IX, W_IY, W_IZ, W_SCALE, W_T1,
    W_POS, W_POS1, W_POS2, W_RD, W_RD1, W_RD2, W_L, W_L1, W_L2, W_TRAY, W_DEPTH, W_XY0, W_SXY, W_XY,
    W_WORDS = 40, W_J = 40, W_DRAWS = 41, W_WORDS_DIAG = 42 // TRACE / COUNT builds carry the job index and the draw count
};
// W_FLAGS: bits 0-1 stage, 2-3 sub-phase, 4 shadow, 5 pixel_open, 6-7 marching mode, 8.. HDDA dim
constexpr uint32_t FL_SHADOW = 16u, FL_PIXEL_OPEN = 32u;

constexpr size_t POOL_SMEM_TABLES = 256 /*logf table*/ + 1504 * 4 /*blackbody*/ + 32 /*shadow RayDir*/ + VPT_POOL_WARPS * 64 /*assign lists*/;
template <bool TRACE, bool COUNT>
constexpr size_t pool_smem_bytes() { return POOL_SMEM_TABLES + size_t(VPT_POOL_WARPS) * VPT_POOL_SLOTS * ((TRACE || COUNT) ? W_WORDS_DIAG : W_WORDS) * 4; }

template <bool TRACE, bool COUNT>
__global__ void __launch_bounds__(POOL_THREADS, VPT_POOL_MIN_BLOCKS) vpt_render_pool_kernel(const RenderArgs A)
{
    constexpr int P = VPT_POOL_SLOTS;
    constexpr int NW = (TRACE || COUNT) ? W_WORDS_DIAG : W_WORDS;
    static_assert(P > 32 && P <= 64, "a warp's pool holds 33..64 job slots");
    extern __shared__ __align__(16) unsigned char pool_smem[];
    double* logtab = reinterpret_cast<double*>(pool_smem);
    float* bb = reinterpret_cast<float*>(pool_smem + 256);
    RayDir* shadow_dir_p = reinterpret_cast<RayDir*>(pool_smem + 256 + 1504 * 4);
    uint8_t* assign = pool_smem + 256 + 1504 * 4 + 32 + (threadIdx.x >> 5) * 64;
    uint32_t* pool = reinterpret_cast<uint32_t*>(pool_smem + POOL_SMEM_TABLES) + size_t(threadIdx.x >> 5) * NW * P;

    for (int i = threadIdx.x; i < 1504; i += blockDim.x) bb[i] = i < 1500 ? A.bb_table[i] : 0.0f;
    if (threadIdx.x < 32) logtab[threadIdx.x] = c_logf_table[threadIdx.x];
    if (threadIdx.x == 32) *shadow_dir_p = make_ray_dir(A.density, mk(A.scene.dist_wi[0], A.scene.dist_wi[1], A.scene.dist_wi[2]));
    const int lane = threadIdx.x & 31;
    for (int s = lane; s < P; s += 32) { // every slot starts in SETUP with an exhausted (empty) job: its first SETUP fetches one
        for (int k = 0; k < NW; ++k) pool[k * P + s] = 0u;
        pool[W_FLAGS * P + s] = uint32_t(ST_SETUP) | uint32_t(PH_NEXT_PIXEL) << 2;
    }
    __syncthreads();

    const SceneDev& S = A.scene;
    const GridDev& G = A.density;
    const unsigned FULL = 0xffffffffu;
    const unsigned lt = (1u << lane) - 1u;
    const int hot_reps = A.pool_params[0], hot_min = 1, thr_hot = A.pool_params[1], thr_eval = A.pool_params[2], thr_setup = A.pool_params[3];
    Stats st = {0, 0, 0, 0, 0, 0};
    unsigned long long draws_total = 0, sd_iters = 0, sd_runs[3] = {0, 0, 0}, sd_lanes[3] = {0, 0, 0}, sd_wait = 0, sd_exit = 0;
    enum : int { M_STEP = 0, M_PEND = 1, M_DRAW = 2 };
    unsigned flip = 0;

#define PLD(k) (sp[(k) * P])
#define PLDF(k) __uint_as_float(sp[(k) * P])
#define PSTF(k, v) (sp[(k) * P] = __float_as_uint(v))
#define PSTI(k, v) (sp[(k) * P] = uint32_t(v))
#define PLD3(k) mk(PLDF(k), PLDF((k) + 1), PLDF((k) + 2))
#define PST3(k, v) do { PSTF(k, (v).x); PSTF((k) + 1, (v).y); PSTF((k) + 2, (v).z); } while (0)
#define PADD_INF_LIGHT() do { PSTF(W_L, PLDF(W_L) + S.inf_light[0]); PSTF(W_L1, PLDF(W_L1) + S.inf_light[1]); PSTF(W_L2, PLDF(W_L2) + S.inf_light[2]); } while (0)
    // the ray left the volume (iterator.next() -> nullopt): back to SETUP
#define POOL_RAY_FINISHED()                                            \
    do {                                                               \
        if (shadow) sub = PH_AFTER_SHADOW;                             \
        else { PADD_INF_LIGHT(); sub = PH_NEXT_PIXEL; }                \
        stage = ST_SETUP;                                              \
    } while (0)

    for (;;) {
        __syncwarp(); // the slots' flags written by the lanes that ran them are visible to the census
        // ---------------------------------------------------------------- 1. census
        const int s0 = int(pool[W_FLAGS * P + lane] & 3u);
        const int s1 = lane + 32 < P ? int(pool[W_FLAGS * P + lane + 32] & 3u) : int(ST_EXIT);
        const unsigned a0 = __ballot_sync(FULL, s0 == ST_SETUP), a1 = __ballot_sync(FULL, s1 == ST_SETUP);
        const unsigned h0 = __ballot_sync(FULL, s0 == ST_HOT), h1 = __ballot_sync(FULL, s1 == ST_HOT);
        const unsigned e0 = __ballot_sync(FULL, s0 == ST_EVAL), e1 = __ballot_sync(FULL, s1 == ST_EVAL);
        const int n_setup = __popc(a0) + __popc(a1), n_hot = __popc(h0) + __popc(h1), n_eval = __popc(e0) + __popc(e1);
        if ((n_setup | n_hot | n_eval) == 0) break; // every job slot has exited
        // ---------------------------------------------------------------- 2. pick a stage. The expensive, rare stages (EVAL ~350, SETUP ~800
        // instructions per lane) run only once a (nearly) full warp of slots is queued for them; HOT, cheap per pass, takes whatever
        // is left — with 64 slots that is still ~32 lanes, because the two waiting queues hold ~16 slots each on average. (Running
        // "the fullest queue" instead equalises the three queues at P/3 = 21 lanes each: measured, profiles/r01_pool_kernel.txt.)
        int pick, best;
        if (n_eval >= thr_eval) { pick = ST_EVAL; best = n_eval; }
        else if (n_setup >= thr_setup) { pick = ST_SETUP; best = n_setup; }
        else if (n_hot >= thr_hot) { pick = ST_HOT; best = n_hot; }
        else { // progress guarantee: the fullest queue (ties: EVAL, SETUP, HOT — they feed HOT)
            pick = ST_EVAL; best = n_eval;
            if (n_setup > best) { pick = ST_SETUP; best = n_setup; }
            if (n_hot > best) { pick = ST_HOT; best = n_hot; }
        }
        best = min(best, 32);
        const unsigned M0 = pick == ST_SETUP ? a0 : (pick == ST_HOT ? h0 : e0), M1 = pick == ST_SETUP ? a1 : (pick == ST_HOT ? h1 : e1);
        // ---------------------------------------------------------------- 3. compaction: the first 32 queued slots, upper / lower half first in turn
        const int c0 = __popc(M0), c1 = __popc(M1);
        if (s0 == pick) { const int r = __popc(M0 & lt) + (flip ? c1 : 0); if (r < 32) assign[r] = uint8_t(lane); }
        if (s1 == pick) { const int r = __popc(M1 & lt) + (flip ? 0 : c0); if (r < 32) assign[r] = uint8_t(lane + 32); }
        flip ^= 1u;
        __syncwarp();
        const bool active = lane < best;
        uint32_t* sp = pool + (active ? int(assign[lane]) : 0);
        if (COUNT && lane == 0) {
            ++sd_iters; ++sd_runs[pick]; sd_lanes[pick] += best;
            sd_wait += n_setup + n_hot + n_eval - best; sd_exit += P - n_setup - n_hot - n_eval;
        }
        const unsigned amask = __ballot_sync(FULL, active);

        // ---------------------------------------------------------------- SETUP: pixel / job bookkeeping, ray begin
        if (pick == ST_SETUP) {
            if (active) {
                uint32_t fl = PLD(W_FLAGS);
                int sub = int(fl >> 2) & 3, stage = ST_SETUP;
                bool shadow = (fl & FL_SHADOW) != 0, pixel_open = (fl & FL_PIXEL_OPEN) != 0;
                Rng rng; rng.state = uint64_t(PLD(W_RLO)) | uint64_t(PLD(W_RHI)) << 32; rng.draws = (TRACE || COUNT) ? PLD(W_DRAWS) : 0u;
                V3 rd = mk(0.0f, 0.0f, 1.0f); // direction of the ray to begin (unused by a shadow ray: shadow_dir)
                unsigned int depth = 0;
                if (sub == PH_NEXT_PIXEL) {
                    const uint32_t xy0 = PLD(W_XY0), sxy = PLD(W_SXY), xy = PLD(W_XY);
                    int x0 = int(xy0 & 0xffffu), y0 = int(xy0 >> 16), sx = int(sxy & 0xffffu), sy = int(sxy >> 16), x = int(xy & 0xffffu), y = int(xy >> 16);
                    if (pixel_open) { // film(y,x) += (imaging_ratio * L, 1)   worker.cpp:203-204
                        const V3 L = PLD3(W_L);
                        if (COUNT) ++st.pixel_samples;
                        if (TRACE) {
                            const size_t slot = size_t(PLD(W_J)) * size_t(S.tw * S.th) + size_t(y) * size_t(S.tw) + size_t(x);
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
                            const unsigned long long j = atomicAdd(A.queue, 1ULL); // TileProvider::next(): m_job_idx.fetch_add(1)
                            if (j >= A.n_jobs) { stage = ST_EXIT; break; }
                            uint64_t jid;
                            if (TRACE) { jid = A.trace_jids[j]; PSTI(W_J, uint32_t(j)); }
                            else {
                                const uint64_t k = j / S.ntiles, tile_k = j - k * S.ntiles;
                                jid = (uint64_t(A.first_wave - 1) + k * A.wave_stride) * S.ntiles + tile_k;
                            }
                            const uint64_t tile = jid % S.ntiles;
                            const int64_t tx0 = int64_t(tile % uint64_t(S.ntx)) * S.tw, ty0 = int64_t(tile / uint64_t(S.ntx)) * S.th;
                            x0 = int(tx0); y0 = int(ty0);
                            sx = int(min(S.W - tx0, S.tw)); sy = int(min(S.H - ty0, S.th));
                            PSTI(W_XY0, uint32_t(x0) | uint32_t(y0) << 16); PSTI(W_SXY, uint32_t(sx) | uint32_t(sy) << 16);
                            rng.begin_job(S.seed, jid);
                            x = 0; y = 0;
                        }
                        if (S.single_pixel_enabled && (int64_t(x0 + x) != S.single_pixel_x || int64_t(y0 + y) != S.single_pixel_y)) continue;
                        break;
                    }
                    PSTI(W_XY, uint32_t(x) | uint32_t(y) << 16);
                    if (stage != ST_EXIT) { // camera path start   worker.cpp:121-128, camera.hpp:14-23
                        const float j0 = rng.uniform(), j1 = rng.uniform();
                        const float jm = S.use_jitter ? 0.5f : 0.0f;
                        const float rx = (float(x0 + x) + 0.5f) + j0 * jm, ry = (float(y0 + y) + 0.5f) + j1 * jm;
                        rd = normalized(mk(sum3(S.cam_m[0] * rx, S.cam_m[1] * ry, S.cam_m[2] * 0.0f) + S.cam_t[0],
                                           sum3(S.cam_m[3] * rx, S.cam_m[4] * ry, S.cam_m[5] * 0.0f) + S.cam_t[1],
                                           sum3(S.cam_m[6] * rx, S.cam_m[7] * ry, S.cam_m[8] * 0.0f) + S.cam_t[2]));
                        PST3(W_RD, rd);
                        PSTF(W_POS, S.cam_pos[0]); PSTF(W_POS1, S.cam_pos[1]); PSTF(W_POS2, S.cam_pos[2]);
                        PSTF(W_L, 0.0f); PSTF(W_L1, 0.0f); PSTF(W_L2, 0.0f);
                        PSTI(W_DEPTH, 0);
                        shadow = false; pixel_open = true;
                    }
                } else if (sub == PH_BEGIN_SHADOW) { // sample_Ld   worker.cpp:59-66
                    shadow = true; PSTF(W_TRAY, 1.0f);
                    if (COUNT) ++st.shadow_rays;
                } else { // PH_AFTER_SHADOW: worker.cpp:88-89 then :175-183
                    const V3 wi = mk(S.dist_wi[0], S.dist_wi[1], S.dist_wi[2]);
                    rd = PLD3(W_RD);
                    const float k = henyey_greenstein(dot3(rd, wi), S.g) * PLDF(W_TRAY);
                    PSTF(W_L, PLDF(W_L) + k * S.dist_Li[0]); PSTF(W_L1, PLDF(W_L1) + k * S.dist_Li[1]); PSTF(W_L2, PLDF(W_L2) + k * S.dist_Li[2]);
                    const float u0 = rng.uniform(), u1 = rng.uniform();
                    rd = sample_henyey_greenstein(rd, u0, u1, S.g);
                    PST3(W_RD, rd);
                    shadow = false;
                    depth = PLD(W_DEPTH) + 2u; // `depth++` at the scatter (worker.cpp:169) and `++depth` of the bounce loop (:130)
                    PSTI(W_DEPTH, depth);
                }
                int dim = 8, mode = M_STEP;
                if (stage != ST_EXIT) {
                    if (!shadow && depth >= S.max_depth) { // bounce loop out of depth: not terminated -> infinite light (:198-200)
                        PADD_INF_LIGHT();
                        sub = PH_NEXT_PIXEL;
                    } else {
                        bool ray_done;
                        March m;
                        CellCache cc; cc.reset();
                        const RayDir rdir = shadow ? *shadow_dir_p : make_ray_dir(G, rd);
                        if (flat_begin_ray_dir<COUNT>(m, G, cc, PLD3(W_POS), rdir, ray_done, st)) {
                            if (ray_done) POOL_RAY_FINISHED();
                            else {
                                stage = ST_HOT; dim = m.dim;
                                PSTF(W_EX, m.ex); PSTF(W_EY, m.ey); PSTF(W_EZ, m.ez); PSTF(W_DX, m.dx); PSTF(W_DY, m.dy); PSTF(W_DZ, m.dz);
                                PSTF(W_IX, m.ix); PSTF(W_IY, m.iy); PSTF(W_IZ, m.iz); PSTF(W_SCALE, m.scale); PSTF(W_T0, m.t0); PSTF(W_T1, m.t1);
                                PSTI(W_VX, m.vx); PSTI(W_VY, m.vy); PSTI(W_VZ, m.vz); PSTF(W_NX, m.nx); PSTF(W_NY, m.ny); PSTF(W_NZ, m.nz);
                                PSTF(W_MAJ, m.majorant); PSTF(W_ST0, m.seg_t0); PSTF(W_SDM, m.seg_dmaj);
                            }
                        } else if (shadow) sub = PH_AFTER_SHADOW; // the shadow ray misses the volume: T stays 1
                        else { PADD_INF_LIGHT(); sub = PH_NEXT_PIXEL; }
                    }
                }
                PSTI(W_RLO, uint32_t(rng.state)); PSTI(W_RHI, uint32_t(rng.state >> 32));
                if (TRACE || COUNT) PSTI(W_DRAWS, rng.draws);
                PSTI(W_FLAGS, uint32_t(stage) | uint32_t(sub) << 2 | (shadow ? FL_SHADOW : 0u) | (pixel_open ? FL_PIXEL_OPEN : 0u) | uint32_t(mode) << 6 | uint32_t(dim) << 8);
            }
        }
        // ---------------------------------------------------------------- HOT: HDDA steps and free-flight draws (software-pipelined as in variant 2)
        else if (pick == ST_HOT) {
            if (active) {
                const uint32_t fl = PLD(W_FLAGS);
                int sub = int(fl >> 2) & 3, stage = ST_HOT, mode = int(fl >> 6) & 3;
                const bool shadow = (fl & FL_SHADOW) != 0;
                Rng rng; rng.state = uint64_t(PLD(W_RLO)) | uint64_t(PLD(W_RHI)) << 32; rng.draws = (TRACE || COUNT) ? PLD(W_DRAWS) : 0u;
                March m;
                m.ex = PLDF(W_EX); m.ey = PLDF(W_EY); m.ez = PLDF(W_EZ); m.dx = PLDF(W_DX); m.dy = PLDF(W_DY); m.dz = PLDF(W_DZ);
                m.ix = PLDF(W_IX); m.iy = PLDF(W_IY); m.iz = PLDF(W_IZ); m.scale = PLDF(W_SCALE); m.t0 = PLDF(W_T0); m.t1 = PLDF(W_T1);
                m.vx = int(PLD(W_VX)); m.vy = int(PLD(W_VY)); m.vz = int(PLD(W_VZ)); m.nx = PLDF(W_NX); m.ny = PLDF(W_NY); m.nz = PLDF(W_NZ);
                m.majorant = PLDF(W_MAJ); m.seg_t0 = PLDF(W_ST0); m.seg_t1 = PLDF(W_ST1); m.seg_dmaj = PLDF(W_SDM);
                m.dim = int(fl >> 8);
                float pend_u = 0.0f;
                CellCache cc; cc.reset();
#pragma unroll 1
                for (int rep = 0; rep < hot_reps; ++rep) {
                    if (stage == ST_HOT) {
                        if (mode == M_PEND) { // RayMajorantIterator::next after the step: volume.cpp:60-70
                            mode = M_STEP;
                            if (m.majorant != m.seg_dmaj) {                 // NaN after a failed step: the segment closes at t1
                                m.seg_t1 = fminf(m.t0, m.t1);               // t0 on a successful step (t0 <= t1), t1 when the step left the ray
                                if (m.seg_dmaj > 0.0f) mode = M_DRAW;
                                else if (m.t0 >= m.t1) POOL_RAY_FINISHED();        // empty segment dropped, iterator.next() -> nullopt
                                else { m.seg_t0 = m.t0; m.seg_dmaj = m.majorant; } // ... or the next segment starts
                            }
                        }
                        if (mode == M_DRAW && stage == ST_HOT) { // MajorantTransmittanceSampler::next, the draw   majorant_transmittance_sampler.cpp:41-47
                            pend_u = rng.uniform();
                            // conservative SFU prefilter of `t < seg.t1` (see variant 2): clear overshoots are settled here, the rest goes to EVAL
                            float l2;
                            asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(1.0f - pend_u));
                            const float lp = l2 * (-0.69314718f * 0.9999f) - 1.0e-6f;
                            const float gap = (m.seg_t1 - m.seg_t0) + (1.0e-5f * fabsf(m.seg_t1) + 1.0e-5f);
                            if (lp > gap * (m.seg_dmaj * S.sigma_t * m.scale)) { // certainly t >= seg.t1: segment dropped
                                mode = M_STEP;
                                if (m.t0 >= m.t1) POOL_RAY_FINISHED();
                                else { m.seg_t0 = m.t0; m.seg_dmaj = m.majorant; }
                            } else stage = ST_EVAL;
                        }
                        if (mode == M_STEP && stage == ST_HOT) { // one HDDA step, majorant fetch issued   volume.cpp:53-59
                            if (!hdda_step_bf(m)) m.majorant = __int_as_float(0x7fc00000);
                            else {
                                const V3 p = m.at(m.t0 + 1.0001f);
                                hdda_update(m, dim_of_code(cell_code_cached(G, cc, int(floorf(p.x)), int(floorf(p.y)), int(floorf(p.z)))));
                                m.majorant = majorant_cached(G, cc, m.vx, m.vy, m.vz);
                                if (COUNT) ++st.majorant_fetches;
                            }
                            mode = M_PEND;
                        }
                    }
                    if (__popc(__ballot_sync(amask, stage == ST_HOT)) < hot_min) break; // too few lanes left: reschedule
                }
                PSTF(W_T0, m.t0); PSTI(W_VX, m.vx); PSTI(W_VY, m.vy); PSTI(W_VZ, m.vz); PSTF(W_NX, m.nx); PSTF(W_NY, m.ny); PSTF(W_NZ, m.nz);
                PSTF(W_MAJ, m.majorant); PSTF(W_ST0, m.seg_t0); PSTF(W_ST1, m.seg_t1); PSTF(W_SDM, m.seg_dmaj); PSTF(W_PENDU, pend_u);
                PSTI(W_RLO, uint32_t(rng.state)); PSTI(W_RHI, uint32_t(rng.state >> 32));
                if (TRACE || COUNT) PSTI(W_DRAWS, rng.draws);
                PSTI(W_FLAGS, uint32_t(stage) | uint32_t(sub) << 2 | (fl & (FL_SHADOW | FL_PIXEL_OPEN)) | uint32_t(mode) << 6 | uint32_t(m.dim) << 8);
            }
        }
        // ---------------------------------------------------------------- EVAL: exact free flight, trilinear density, collision   worker.cpp:68-84,148-188
        else {
            if (active) {
                const uint32_t fl = PLD(W_FLAGS);
                int sub = int(fl >> 2) & 3, stage = ST_HOT, mode = int(fl >> 6) & 3;
                const bool shadow = (fl & FL_SHADOW) != 0;
                Rng rng; rng.state = uint64_t(PLD(W_RLO)) | uint64_t(PLD(W_RHI)) << 32; rng.draws = (TRACE || COUNT) ? PLD(W_DRAWS) : 0u;
                March m;
                m.ex = PLDF(W_EX); m.ey = PLDF(W_EY); m.ez = PLDF(W_EZ); m.dx = PLDF(W_DX); m.dy = PLDF(W_DY); m.dz = PLDF(W_DZ);
                m.scale = PLDF(W_SCALE); m.t0 = PLDF(W_T0); m.t1 = PLDF(W_T1);
                m.majorant = PLDF(W_MAJ); m.seg_t0 = PLDF(W_ST0); m.seg_t1 = PLDF(W_ST1); m.seg_dmaj = PLDF(W_SDM);
                const float pend_u = PLDF(W_PENDU);
                const float sigma_maj = m.seg_dmaj * S.sigma_t;
                const float dt_m = -vptm::logf_exact(1.0f - pend_u, logtab) / sigma_maj; // sample_exponential  random.hpp:19-22
                const float t = m.seg_t0 + dt_m / m.scale;
                float density = 0.0f;
                V3 pos_idx = mk(0.0f, 0.0f, 0.0f);
                if (t < m.seg_t1) {
                    m.seg_t0 = t;
                    const V3 p = m.at(t);
                    if (COUNT) ++st.density_evals;
                    CellCache cc; cc.reset();
                    density = trilinear_cached(G, cc, p.x, p.y, p.z);
                    if (density > 0.0f) pos_idx = p;
                } else { // overshoot after all: segment dropped, iterator.next()
                    mode = M_STEP;
                    if (m.t0 >= m.t1) POOL_RAY_FINISHED();
                    else { m.seg_t0 = m.t0; m.seg_dmaj = m.majorant; }
                }
                if (density > 0.0f) {
                    if (shadow) { // ratio tracking + Russian roulette
                        const float sigma_n = fmaxf(0.0f, sigma_maj - S.sigma_t * density);
                        float T_ray = PLDF(W_TRAY) * (sigma_n / sigma_maj);
                        if (T_ray <= 0.05f) {
                            if (rng.uniform() < 0.75f) T_ray = 0.0f;
                            else T_ray /= 0.25f;
                        }
                        if (T_ray <= 0.0f) { T_ray = 0.0f; sub = PH_AFTER_SHADOW; stage = ST_SETUP; }
                        PSTF(W_TRAY, T_ray);
                    } else {
                        const V3 point = index_to_world(G, pos_idx);
                        const float p_a = (S.sigma_a * density) / sigma_maj;
                        const float p_s = (S.sigma_s * density) / sigma_maj;
                        const float p_n = fmaxf(1.0f - p_a - p_s, 0.0f);
                        if (S.has_temperature) {
                            const V3 tc = world_to_index(A.temperature, point);
                            if (COUNT) ++st.temperature_evals;
                            CellCache ct; ct.reset();
                            const float temp_adim = trilinear_cached(A.temperature, ct, tc.x, tc.y, tc.z);
                            const float temp_K = temp_adim * S.temperature_scale + S.temperature_offset;
                            const V3 e = blackbody_xyz(bb, temp_K);
                            const float k = p_a * S.le_scale;
                            PSTF(W_L, PLDF(W_L) + k * e.x); PSTF(W_L1, PLDF(W_L1) + k * e.y); PSTF(W_L2, PLDF(W_L2) + k * e.z);
                        }
                        float u = rng.uniform() * (((0.0f + p_n) + p_a) + p_s); // sample_discrete  random.hpp:31-47
                        u -= p_n;
                        if (u > 0.0f) {
                            u -= p_a;
                            stage = ST_SETUP;
                            if (u <= 0.0f) sub = PH_NEXT_PIXEL; // absorption: terminated, no infinite light
                            else {                               // scatter (also the fallback choice)
                                if (COUNT) ++st.scatter_events;
                                PST3(W_POS, point);
                                if (S.dist_light_on) sub = PH_BEGIN_SHADOW;
                                else { PSTF(W_TRAY, 0.0f); sub = PH_AFTER_SHADOW; } // sample_Ld returns 0 before any draw (:57)
                            }
                        }
                    }
                }
                PSTF(W_ST0, m.seg_t0); PSTF(W_SDM, m.seg_dmaj);
                PSTI(W_RLO, uint32_t(rng.state)); PSTI(W_RHI, uint32_t(rng.state >> 32));
                if (TRACE || COUNT) PSTI(W_DRAWS, rng.draws);
                PSTI(W_FLAGS, uint32_t(stage) | uint32_t(sub) << 2 | (fl & (FL_SHADOW | FL_PIXEL_OPEN)) | uint32_t(mode) << 6 | (fl & ~0xffu));
            }
        }
    }
#undef POOL_RAY_FINISHED
#undef PADD_INF_LIGHT
#undef PLD
#undef PLDF
#undef PSTF
#undef PSTI
#undef PLD3
#undef PST3
    if (COUNT) {
        atomicAdd(&A.counters->pixel_samples, (unsigned long long)st.pixel_samples);
        atomicAdd(&A.counters->density_evals, (unsigned long long)st.density_evals);
        atomicAdd(&A.counters->temperature_evals, (unsigned long long)st.temperature_evals);
        atomicAdd(&A.counters->majorant_fetches, (unsigned long long)st.majorant_fetches);
        atomicAdd(&A.counters->rng_draws, draws_total);
        atomicAdd(&A.counters->scatter_events, (unsigned long long)st.scatter_events);
        atomicAdd(&A.counters->shadow_rays, (unsigned long long)st.shadow_rays);
        if (lane == 0) {
            atomicAdd(&A.counters->sched_iters, sd_iters);
            for (int i = 0; i < 3; ++i) { atomicAdd(&A.counters->sched_runs[i], sd_runs[i]); atomicAdd(&A.counters->sched_lanes[i], sd_lanes[i]); }
            atomicAdd(&A.counters->sched_waiting, sd_wait);
            atomicAdd(&A.counters->sched_exited, sd_exit);
        }
    }
}

} // namespace vptd
