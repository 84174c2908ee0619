// vpt_render_sched.cuh — variant 2: the flat megakernel with WARP-LEVEL STAGE SCHEDULING.
//
// Evidence (profiles/r01_dense512_8waves_flat.summary.txt + the per-line source view): in the flat kernel the
// two hot blocks run with 13-21 active lanes, but the rare-and-long blocks (ray setup: Volume::intersect with
// its divisions and square roots, HG sampling, camera ray; the trilinear fetch + collision logic) run with
// 1.6-3 lanes and account for ~45 % of all issued warp instructions — almost every iteration finds at least
// ONE of the 32 lanes needing them. Here a lane carries an explicit stage
//     SETUP -> HOT (HDDA steps and free-flight draws) -> EVAL (-> HOT | SETUP)
// and each loop iteration runs a stage only if enough lanes are queued for it (ballot + popc against a per-stage
// threshold) or it is the most populated one (progress guarantee); HOT runs VPT_HOT_REPS passes per decision. Lanes
// queued for a stage that is skipped simply wait; the arithmetic and its order per lane are unchanged, so results
// stay bit-identical to variant 1. This is the in-register, in-warp form of wavefront regrouping: the hot path
// state never goes to memory (the cold part sits in shared memory, see below).
#pragma once
#include "vpt_render_flat.cuh"

namespace vptd {

#ifndef VPT_SCHED_MIN_BLOCKS
#define VPT_SCHED_MIN_BLOCKS 7 // resident 128-thread blocks per SM the register allocation must allow (72 registers)
#endif

// Explicit reconvergence point (A/B switch). The kernel's speed hangs on where ptxas lets the lanes of a stage
// rejoin: an edit as small as one conditional load in the job fetch made SETUP's ray begin run at 4.5 instead of 7.7
// lanes and the kernel 25 % slower (profiles/r01_codegen_cliff.txt). -DVPT_SYNC_SETUP=1 pins the rejoin with a
// __syncwarp over the lanes that entered the stage; it costs ~1 % when the compiler already does the right thing
// (the present code), so it is off — turn it on first when an edit to SETUP slows the kernel unexpectedly.
#ifndef VPT_HOT_REPS
#define VPT_HOT_REPS 3 // HOT passes per scheduling decision: cuts the ballot/threshold overhead per HDDA step (+4..7 %, profiles/r01_codegen_cliff.txt)
#endif
#ifndef VPT_PREFILTER_MUL
#define VPT_PREFILTER_MUL 1 // overshoot prefilter without the reciprocal: +1 % (profiles/r01_codegen_cliff.txt); 0 = the MUFU.RCP form
#endif
#ifndef VPT_HOT_SINGLE_FINISH
#define VPT_HOT_SINGLE_FINISH 1 // HOT's two ray-finished sites only mark the lane; the bookkeeping runs once per pass: +1..3 % (profiles/r01_codegen_cliff.txt)
#endif
#ifndef VPT_SYNC_SETUP
#define VPT_SYNC_SETUP 0
#endif
#ifndef VPT_SHADOW_IN_EVAL
#define VPT_SHADOW_IN_EVAL 0 // a scatter begins its shadow ray inside EVAL instead of queueing for SETUP: one stage transition less per scatter (profiles/r02_kernel_ab.txt)
#endif
#ifndef VPT_DEFER_INF
#define VPT_DEFER_INF 0 // a ray that leaves the volume only marks the pixel; the infinite light is added where the pixel is closed (no cold-state traffic in HOT)
#endif
#ifndef VPT_FILM_IN_COLD
#define VPT_FILM_IN_COLD 0 // the film pointer of the lane's current range is kept in two cold words (written at the job fetch) instead of being re-read from the
                           // ring descriptor (an L2 round trip) at every pixel: matters where a pixel-sample is cheap (fire: ~10 HDDA steps)
#endif
#ifndef VPT_FETCH_DIV32
#define VPT_FETCH_DIV32 1 // job fetch: 32-bit divisions where the operands fit, tile taken from the decomposition instead of jid % ntiles (profiles/r02_kernel_ab.txt)
#endif
#ifndef VPT_LOOKAHEAD
#define VPT_LOOKAHEAD 0 // the HDDA runs ONE STEP AHEAD of the segment logic: a pass first takes the pair (t0, majorant) whose load the previous pass issued, steps again and
                        // issues the next load, and only then consumes the pair — a whole pass between a slot load and its use (profiles/r02_kernel_ab.txt)
#endif
enum : int { ST_SETUP = 0, ST_HOT = 1, ST_EVAL = 2, ST_EXIT = 3 };
enum : int { PH_NEXT_PIXEL_INF = 5 }; // NEXT_PIXEL for a path that escaped to the infinite light (VPT_DEFER_INF)

// STATS: the instantiation adaptive rendering launches — an explicit tile list per range and a second-moment plane beside the film. A template
// parameter, not a run-time test, so that the production instantiation's SASS does not change (the job fetch is where one extra conditional
// load once cost 25 %: profiles/r01_codegen_cliff.txt).
// ORD: the tile-order instantiations (launch_render_ordered) — 1 LEARNS (natural order, each finished job adds its duration to the range's
// tile_cost), 2 walks the range's tile_list tile-major. Template parameters for the same reason as STATS: ORD = 0 is the kernel as it was.
template <bool TRACE, bool COUNT, bool STATS = false, int ORD = 0>
__global__ void __launch_bounds__(VPT_BLOCK_THREADS, VPT_SCHED_MIN_BLOCKS) vpt_render_sched_kernel(const RenderArgs A)
{
    __shared__ float bb[1504];
    __shared__ double logtab[32];
    for (int i = threadIdx.x; i < 1504; i += blockDim.x) bb[i] = i < 1500 ? A.bb_table[i] : 0.0f;
    if (threadIdx.x < 32) logtab[threadIdx.x] = c_logf_table[threadIdx.x];
    __shared__ RayDir shadow_dir; // every shadow ray points at the one distant light: direction terms once per block
    if (threadIdx.x == 32) shadow_dir = make_ray_dir(A.density, mk(A.scene.dist_wi[0], A.scene.dist_wi[1], A.scene.dist_wi[2]));
    __syncthreads();

    const SceneDev& S = A.scene;
    const GridDev& G = A.density;
    const unsigned FULL = 0xffffffffu;
    Stats st = {0, 0, 0, 0, 0, 0};
    unsigned long long draws_total = 0, sd_iters = 0, sd_runs[3] = {0, 0, 0}, sd_lanes[3] = {0, 0, 0}, sd_wait = 0, sd_exit = 0;
    Rng rng; rng.state = 0; rng.draws = 0;
    CellCache cc;
    cc.reset();
    March m;
    enum : int { M_STEP = 0, M_PEND = 1, M_DRAW = 2 }; // marching: step next / majorant fetch in flight; drawing in a closed segment
    int mode = M_STEP;
    float pend_u = 0.0f;
#if VPT_LOOKAHEAD
    float c_t0 = 0.0f, c_maj = 0.0f; // the pair the segment logic is looking at; (m.t0, m.majorant) is the HDDA's, one step further
#define VPT_CUR_T0 c_t0
#define VPT_CUR_MAJ c_maj
#else
#define VPT_CUR_T0 m.t0
#define VPT_CUR_MAJ m.majorant
#endif
    // one HDDA step with the next node's majorant fetch issued   volume.cpp:53-59
#define VPT_HDDA_ADVANCE()                                                                                                              \
    do {                                                                                                                                \
        if (!hdda_step_bf(m)) m.majorant = __int_as_float(0x7fc00000);                                                                  \
        else {                                                                                                                          \
            const V3 p_ = m.at(m.t0 + 1.0001f);                                                                                         \
            hdda_update(m, dim_of_code(cell_code_cached(G, cc, int(floorf(p_.x)), int(floorf(p_.y)), int(floorf(p_.z)))));              \
            m.majorant = majorant_cached(G, cc, m.vx, m.vy, m.vz);                                                                      \
            if (COUNT && !VPT_LOOKAHEAD) ++st.majorant_fetches;                                                                         \
        }                                                                                                                               \
    } while (0)
    bool shadow = false, pixel_open = false;
    int stage = ST_SETUP, sub = PH_NEXT_PIXEL;
    unsigned long long j = 0;
    // COLD per-lane state lives in shared memory, one word column per lane (conflict-free): everything HOT never
    // touches — path origin/direction/radiance, shadow transmittance, depth, the job's pixel bookkeeping, the
    // temperature grid's cell cache. That is 21 registers the hot loop does not have to carry, which is what lets
    // the kernel fit VPT_SCHED_MIN_BLOCKS resident blocks per SM without spilling.
    enum { C_POS = 0, C_RD = 3, C_L = 6, C_TRAY = 9, C_DEPTH = 10, C_X0 = 11, C_Y0 = 12, C_SX = 13, C_SY = 14, C_X = 15, C_Y = 16, C_CT = 17, C_RANGE = 21, C_FILM = 22, C_TILE = 22 + 2 * VPT_FILM_IN_COLD, C_T0 = C_TILE + 1, C_WORDS = 22 + 2 * VPT_FILM_IN_COLD + 2 * (ORD == 1) };
    __shared__ uint32_t cold[C_WORDS][VPT_BLOCK_THREADS];
    const int lane_col = threadIdx.x;
#define CLDF(i) __uint_as_float(cold[i][lane_col])
#define CLDI(i) int(cold[i][lane_col])
#define CSTF(i, v) (cold[i][lane_col] = __float_as_uint(v))
#define CSTI(i, v) (cold[i][lane_col] = uint32_t(v))
#define CLD3(i) mk(CLDF(i), CLDF((i) + 1), CLDF((i) + 2))
#define CST3(i, v) do { CSTF(i, (v).x); CSTF((i) + 1, (v).y); CSTF((i) + 2, (v).z); } while (0)
    for (int i = 0; i < C_WORDS; ++i) cold[i][lane_col] = 0u; // sx = sy = x = y = 0: the first SETUP fetches a job
    cold[C_RANGE][lane_col] = A.range_begin;                  // the range (render call) this lane takes jobs from: starts at this launch's own
    { CellCache r; r.reset(); CSTI(C_CT, r.x); CSTI(C_CT + 1, r.y); CSTI(C_CT + 2, r.z); cold[C_CT + 3][lane_col] = r.code; }

    // Logger<true> (worker.cpp:15-49), trace launches only: {type, 6 arguments, 0} appended in program order (callers trace ONE job)
#define VPT_EVENT(type, a, b, c, d, e, f)                                                          \
    do {                                                                                           \
        if (TRACE && A.trace_events) {                                                             \
            const unsigned long long i_ = atomicAdd(A.trace_event_count, 1ULL);                    \
            if (i_ < A.trace_event_cap) {                                                          \
                float* e_ = A.trace_events + i_ * 8;                                               \
                e_[0] = float(type); e_[1] = (a); e_[2] = (b); e_[3] = (c); e_[4] = (d); e_[5] = (e); e_[6] = (f); e_[7] = 0.0f; \
            }                                                                                      \
        }                                                                                          \
    } while (0)
    // the ray left the volume (iterator.next() -> nullopt): back to SETUP
#if VPT_DEFER_INF
#define VPT_ADD_INF_LIGHT() (sub = PH_NEXT_PIXEL_INF) /* L += infinite light happens where the pixel is closed: same single fp32 add per channel */
#else
#define VPT_ADD_INF_LIGHT() do { CSTF(C_L, CLDF(C_L) + S.inf_light[0]); CSTF(C_L + 1, CLDF(C_L + 1) + S.inf_light[1]); CSTF(C_L + 2, CLDF(C_L + 2) + S.inf_light[2]); sub = PH_NEXT_PIXEL; } while (0)
#endif
#define VPT_RAY_FINISHED()                                                                            \
    do {                                                                                              \
        if (shadow) sub = PH_AFTER_SHADOW;                                                            \
        else VPT_ADD_INF_LIGHT();                                                                     \
        stage = ST_SETUP;                                                                             \
    } while (0)

    for (;;) {
        const int n_setup = __popc(__ballot_sync(FULL, stage == ST_SETUP)), n_hot = __popc(__ballot_sync(FULL, stage == ST_HOT));
        const int n_eval = __popc(__ballot_sync(FULL, stage == ST_EVAL));
        const int n_max = max(n_setup, max(n_hot, n_eval));
        if (n_max == 0) break; // every lane has exited
        const bool run_setup = n_setup >= A.sched_thresh[0] || n_setup == n_max;
        const bool run_hot = n_hot >= A.sched_thresh[1] || n_hot == n_max;
        const bool run_eval = n_eval >= A.sched_thresh[3] || n_eval == n_max;
        if (COUNT && (threadIdx.x & 31) == 0) {
            ++sd_iters;
            sd_runs[0] += run_setup && n_setup > 0; sd_runs[1] += run_hot && n_hot > 0; sd_runs[2] += run_eval && n_eval > 0;
            sd_lanes[0] += run_setup ? n_setup : 0; sd_lanes[1] += run_hot ? n_hot : 0; sd_lanes[2] += run_eval ? n_eval : 0;
            sd_wait += (run_setup ? 0 : n_setup) + (run_hot ? 0 : n_hot) + (run_eval ? 0 : n_eval);
            sd_exit += 32 - n_setup - n_hot - n_eval;
        }

        // ------------------------------------------------------------ SETUP: pixel / job bookkeeping, ray begin
#if VPT_SYNC_SETUP
        const unsigned setup_mask = __ballot_sync(FULL, run_setup && stage == ST_SETUP);
#endif
        if (run_setup && stage == ST_SETUP) {
            V3 rd = mk(0.0f, 0.0f, 1.0f); // direction of the ray to begin (unused by a shadow ray: shadow_dir)
            unsigned int depth = 0;
            if (sub == PH_NEXT_PIXEL || sub == PH_NEXT_PIXEL_INF) {
                int x0 = CLDI(C_X0), y0 = CLDI(C_Y0), sx = CLDI(C_SX), sy = CLDI(C_SY), x = CLDI(C_X), y = CLDI(C_Y);
                if (pixel_open) { // film(y,x) += (imaging_ratio * L, 1)   worker.cpp:203-204
                    V3 L = CLD3(C_L);
                    if (VPT_DEFER_INF && sub == PH_NEXT_PIXEL_INF) { L.x += S.inf_light[0]; L.y += S.inf_light[1]; L.z += S.inf_light[2]; } // worker.cpp:198-200
                    if (COUNT) ++st.pixel_samples;
                    if (TRACE) {
                        const size_t slot = size_t(j) * size_t(S.tw * S.th) + size_t(y) * size_t(S.tw) + size_t(x);
                        A.trace_xyz[slot * 3 + 0] = L.x; A.trace_xyz[slot * 3 + 1] = L.y; A.trace_xyz[slot * 3 + 2] = L.z;
                        A.trace_ndraws[slot] = rng.draws;
                    } else {
#if VPT_FILM_IN_COLD
                        float* film = reinterpret_cast<float*>((unsigned long long)(cold[C_FILM][lane_col]) | ((unsigned long long)(cold[C_FILM + 1][lane_col]) << 32));
#else
                        float* film = reinterpret_cast<float*>(__ldcg(reinterpret_cast<const unsigned long long*>(&A.ring[cold[C_RANGE][lane_col] % RING_SLOTS].film))); // the film of the call this job belongs to
#endif
                        film_add(film + (size_t(y0 + y) * size_t(S.W) + size_t(x0 + x)) * 4, S.imaging_ratio * L.x, S.imaging_ratio * L.y, S.imaging_ratio * L.z);
                        if (STATS) {
                            float* sq = reinterpret_cast<float*>(__ldcg(reinterpret_cast<const unsigned long long*>(&A.ring[cold[C_RANGE][lane_col] % RING_SLOTS].sum_y2)));
                            const float yl = S.imaging_ratio * L.y;
                            if (sq) atomicAdd(sq + (size_t(y0 + y) * size_t(S.W) + size_t(x0 + x)), yl * yl);
                        }
                    }
                    pixel_open = false;
                }
                for (;;) { // next pixel of the job (row-major, worker.cpp:109-110), next job when exhausted
                    if (++x >= sx) { x = 0; ++y; }
                    if (y >= sy) {
                        if (COUNT) draws_total += rng.draws;
                        // TileProvider::next(): job = m_job_idx.fetch_add(1) (tile_provider.cpp:28) — on the current range of the ring; an exhausted
                        // range hands over to the next published one (a later render call) instead of letting the lane drain.
                        unsigned int rc = cold[C_RANGE][lane_col];
                        RangeDesc* R = A.ring + (rc % RING_SLOTS);
                        if (sy > 0) { // a job of range rc just finished: count it; the last one raises the host's completion flag
                            if (ORD == 1) { // how long the job took, per tile: what the host sorts the next calls' tiles by
                                uint32_t* cost = reinterpret_cast<uint32_t*>(__ldcg(reinterpret_cast<const unsigned long long*>(&R->tile_cost)));
                                if (cost) atomicAdd(cost + cold[C_TILE][lane_col], max(1u, (uint32_t(clock()) - cold[C_T0][lane_col]) >> 10));
                            }
                            __threadfence(); // this lane's film reductions before the count
                            if (atomicAdd(&R->done, 1ULL) + 1ULL == __ldcg(&R->n_jobs)) {
                                __threadfence_system();
                                if (A.host_done) *reinterpret_cast<volatile unsigned int*>(A.host_done + (rc % RING_SLOTS)) = rc + 1u;
                            }
                            sy = 0;
                        }
                        for (;;) {
                            j = atomicAdd(&R->counter, 1ULL);
                            if (j < __ldcg(&R->n_jobs)) break;
                            // ring dry — or this launch has thinned out (lanes leave whenever they find the ring momentarily dry; the survivors
                            // must not nurse later ranges along while a full grid waits in the stream behind them)
                            if (rc + 1u >= *reinterpret_cast<const volatile unsigned int*>(A.ring_tail) ||
                                *reinterpret_cast<const volatile unsigned int*>(A.exited) > A.exit_limit) { stage = ST_EXIT; break; }
                            ++rc;
                            R = A.ring + (rc % RING_SLOTS);
                        }
                        cold[C_RANGE][lane_col] = rc;
#if VPT_FILM_IN_COLD
                        if (!TRACE && stage != ST_EXIT) {
                            const unsigned long long fp = __ldcg(reinterpret_cast<const unsigned long long*>(&R->film));
                            cold[C_FILM][lane_col] = uint32_t(fp); cold[C_FILM + 1][lane_col] = uint32_t(fp >> 32);
                        }
#endif
                        if (stage == ST_EXIT) { CSTI(C_SY, 0); atomicAdd(A.exited, 1u); break; }
                        uint64_t jid, tile;
                        if (TRACE) { jid = A.trace_jids[j]; tile = jid % S.ntiles; }
                        else {
                            const uint64_t tcount = __ldcg(&R->tile_count);
#if VPT_FETCH_DIV32
                            // the 64-bit divisions are library calls of ~70 instructions each, run by ONE lane while 31 wait: where every operand fits
                            // 32 bits (any frame below 2^32 jobs per call) the inline 32-bit form does the same arithmetic
                            uint64_t k, ti, trow;
                            if (((j | tcount) >> 32) == 0) {
                                const uint32_t k32 = uint32_t(j) / uint32_t(tcount);
                                k = k32; ti = uint32_t(j) - k32 * uint32_t(tcount); trow = uint32_t(ti) / uint32_t(S.ntx);
                            } else { k = j / tcount; ti = j - k * tcount; trow = ti / uint64_t(S.ntx); }
#else
                            uint64_t k = j / tcount;
                            const uint64_t ti = j - k * tcount, trow = ti / uint64_t(S.ntx); // interleaved stripes: every tile_row_stride-th row of tiles
#endif
                            uint64_t tile_k = __ldcg(&R->tile_begin) + trow * (uint64_t(__ldcg(&R->tile_row_stride)) * uint64_t(S.ntx)) + (ti - trow * uint64_t(S.ntx));
                            if (STATS || ORD == 2) { // explicit tile list: the range is walked tile-major (layout.hpp RangeDesc)
                                const uint32_t* list = reinterpret_cast<const uint32_t*>(__ldcg(reinterpret_cast<const unsigned long long*>(&R->tile_list)));
                                if (list) {
                                    const uint32_t nw = __ldcg(&R->n_waves);
                                    const uint64_t slot = ((j >> 32) == 0) ? uint64_t(uint32_t(j) / nw) : j / nw;
                                    k = j - slot * nw;
                                    tile_k = __ldg(list + slot);
                                }
                            }
                            jid = (uint64_t(__ldcg(&R->first_wave) - 1u) + k * __ldcg(&R->wave_stride)) * S.ntiles + tile_k;
#if VPT_FETCH_DIV32
                            tile = tile_k; // = jid % ntiles by construction (tile_k < ntiles)
#else
                            tile = jid % S.ntiles;
#endif
                        }
#if VPT_FETCH_DIV32
                        uint64_t tyq;
                        if ((tile >> 32) == 0) tyq = uint32_t(tile) / uint32_t(S.ntx); else tyq = tile / uint64_t(S.ntx);
                        const int64_t tx0 = int64_t(tile - tyq * uint64_t(S.ntx)) * S.tw, ty0 = int64_t(tyq) * S.th;
#else
                        const int64_t tx0 = int64_t(tile % uint64_t(S.ntx)) * S.tw, ty0 = int64_t(tile / uint64_t(S.ntx)) * S.th;
#endif
                        x0 = int(tx0); y0 = int(ty0);
                        sx = int(min(S.W - tx0, S.tw)); sy = int(min(S.H - ty0, S.th));
                        CSTI(C_X0, x0); CSTI(C_Y0, y0); CSTI(C_SX, sx); CSTI(C_SY, sy);
                        if (ORD == 1) { cold[C_TILE][lane_col] = uint32_t(tile); cold[C_T0][lane_col] = uint32_t(clock()); }
                        rng.begin_job(S.seed, jid);
                        x = 0; y = 0;
                    }
                    if (S.single_pixel_enabled && (int64_t(x0 + x) != S.single_pixel_x || int64_t(y0 + y) != S.single_pixel_y)) continue;
                    break;
                }
                CSTI(C_X, x); CSTI(C_Y, y);
                if (stage != ST_EXIT) { // camera path start   worker.cpp:121-128, camera.hpp:14-23
                    const float j0 = rng.uniform(), j1 = rng.uniform();
                    const float jm = S.use_jitter ? 0.5f : 0.0f;
                    const float rx = (float(x0 + x) + 0.5f) + j0 * jm, ry = (float(y0 + y) + 0.5f) + j1 * jm;
                    rd = normalized(mk(sum3(S.cam_m[0] * rx, S.cam_m[1] * ry, S.cam_m[2] * 0.0f) + S.cam_t[0],
                                       sum3(S.cam_m[3] * rx, S.cam_m[4] * ry, S.cam_m[5] * 0.0f) + S.cam_t[1],
                                       sum3(S.cam_m[6] * rx, S.cam_m[7] * ry, S.cam_m[8] * 0.0f) + S.cam_t[2]));
                    CST3(C_RD, rd);
                    VPT_EVENT(0, S.cam_pos[0], S.cam_pos[1], S.cam_pos[2], rd.x, rd.y, rd.z); // logger.new_ray   worker.cpp:125
                    CSTF(C_POS, S.cam_pos[0]); CSTF(C_POS + 1, S.cam_pos[1]); CSTF(C_POS + 2, S.cam_pos[2]);
                    CSTF(C_L, 0.0f); CSTF(C_L + 1, 0.0f); CSTF(C_L + 2, 0.0f);
                    CSTI(C_DEPTH, 0);
                    shadow = false; pixel_open = true;
                }
            } else if (sub == PH_BEGIN_SHADOW) { // sample_Ld   worker.cpp:59-66
                shadow = true; CSTF(C_TRAY, 1.0f);
                if (COUNT) ++st.shadow_rays;
            } else { // PH_AFTER_SHADOW: worker.cpp:88-89 then :175-183
                const V3 wi = mk(S.dist_wi[0], S.dist_wi[1], S.dist_wi[2]);
                rd = CLD3(C_RD);
                const float k = henyey_greenstein(dot3(rd, wi), S.g) * CLDF(C_TRAY);
                CSTF(C_L, CLDF(C_L) + k * S.dist_Li[0]); CSTF(C_L + 1, CLDF(C_L + 1) + k * S.dist_Li[1]); CSTF(C_L + 2, CLDF(C_L + 2) + k * S.dist_Li[2]);
                const float u0 = rng.uniform(), u1 = rng.uniform();
                rd = sample_henyey_greenstein(rd, u0, u1, S.g);
                CST3(C_RD, rd);
                VPT_EVENT(3, CLDF(C_POS), CLDF(C_POS + 1), CLDF(C_POS + 2), rd.x, rd.y, rd.z); // logger.scatter(r)   worker.cpp:181
                shadow = false;
                depth = unsigned(CLDI(C_DEPTH)) + 2u; // `depth++` at the scatter (worker.cpp:169) and `++depth` of the bounce loop (:130)
                CSTI(C_DEPTH, depth);
            }
#if VPT_SYNC_SETUP
            __syncwarp(setup_mask); // the three sub-phases rejoin here: ray begin runs with every SETUP lane
#endif
            if (stage != ST_EXIT) {
                if (!shadow && depth >= S.max_depth) { // bounce loop out of depth: not terminated -> infinite light (:198-200)
                    VPT_ADD_INF_LIGHT();
                } else {
                    bool ray_done;
                    const RayDir rdir = shadow ? shadow_dir : make_ray_dir(G, rd);
                    if (flat_begin_ray_dir<COUNT>(m, G, cc, CLD3(C_POS), rdir, ray_done, st)) {
                        if (ray_done) VPT_RAY_FINISHED();
                        else {
                            stage = ST_HOT; mode = M_STEP;
#if VPT_LOOKAHEAD
                            c_t0 = m.t0; c_maj = m.majorant; // the first pair is current; the HDDA goes one step ahead right here
                            VPT_HDDA_ADVANCE();
#endif
                        }
                    } else if (shadow) sub = PH_AFTER_SHADOW; // the shadow ray misses the volume: T stays 1
                    else VPT_ADD_INF_LIGHT(); // the camera / bounce ray misses the volume
                }
            }
        }

        // ------------------------------------------------------------ HOT: one HDDA step and/or one free-flight draw
#if VPT_HOT_REPS > 1
#pragma unroll 1
        for (int rep = 0; rep < VPT_HOT_REPS; ++rep) // several HOT passes per scheduling decision (A/B switch)
#endif
        // Software-pipelined: a pass CONSUMES the majorant whose fetch the previous pass issued, draws if a segment closed,
        // then takes the next HDDA step and ISSUES that voxel's fetch — the L2 latency of the slot load (the kernel's top
        // stall: 10 % of all stall samples at `majorant != seg_dmaj`) overlaps the draw, the prefilter and the loop
        // overhead instead of sitting between the step and its use. Same operations per lane in the same order.
        if (run_hot && stage == ST_HOT) {
#if VPT_HOT_SINGLE_FINISH
            bool fin = false;
#define HOT_FINISH() do { fin = true; stage = ST_SETUP; } while (0)
#else
#define HOT_FINISH() VPT_RAY_FINISHED()
#endif
#if VPT_LOOKAHEAD
            if (mode == M_STEP) { // the pair fetched a pass ago becomes current (its load is waited for HERE), and the HDDA moves on at once
                c_t0 = m.t0; c_maj = m.majorant;
                if (COUNT && c_t0 <= m.t1) ++st.majorant_fetches; // counted when consumed: the reference never fetches the look-ahead node of a ray that ends first
                if (m.t0 <= m.t1) VPT_HDDA_ADVANCE();             // a step that left the ray is the last one
                mode = M_PEND;
            }
#endif
            if (mode == M_PEND) { // RayMajorantIterator::next after the step: volume.cpp:60-70
                mode = M_STEP;
                if (VPT_CUR_MAJ != m.seg_dmaj) {                 // NaN after a failed step: the segment closes at t1
                    m.seg_t1 = fminf(VPT_CUR_T0, m.t1);               // t0 on a successful step (t0 <= t1), t1 when the step left the ray
                    if (m.seg_dmaj > 0.0f) mode = M_DRAW;
                    else if (VPT_CUR_T0 >= m.t1) HOT_FINISH();               // empty segment dropped, iterator.next() -> nullopt
                    else { m.seg_t0 = VPT_CUR_T0; m.seg_dmaj = VPT_CUR_MAJ; } // ... or the next segment starts
                }
            }
            if (mode == M_DRAW && stage == ST_HOT) { // MajorantTransmittanceSampler::next, the draw   majorant_transmittance_sampler.cpp:41-47
                pend_u = rng.uniform();
                // ~85 % of the draws overshoot the segment. Only the DECISION `t < seg.t1` matters then, not t itself, so
                // a conservative test with the SFU approximations (MUFU.LG2 / RCP, error margins far above their
                // documented bounds) settles the clear cases; everything else — the hits and the ~0.1 % near the
                // boundary — goes to EVAL, which evaluates t exactly (glibc-exact logf, IEEE divisions) and decides.
#if VPT_PREFILTER_MUL
                // the same test multiplied through by k = d_maj * sigma_t * scale > 0 (no reciprocal, no denormal rescaling: 1 - u >= 2^-24):
                //   0.9999 * (-ln(1 - u)) - 1e-6  >  ((t1 - t0) + margin) * k      (k = 0 / denormal -> true overshoot; k = inf -> EVAL)
                float l2;
                asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(1.0f - pend_u));
                const float lp = l2 * (-0.69314718f * 0.9999f) - 1.0e-6f;
                const float gap = (m.seg_t1 - m.seg_t0) + (1.0e-5f * fabsf(m.seg_t1) + 1.0e-5f);
                if (lp > gap * (m.seg_dmaj * S.sigma_t * m.scale)) { // certainly t >= seg.t1: segment dropped
#else
                const float lp = -__logf(1.0f - pend_u) - 1.0e-6f;
                const float dtp = __fdividef(lp, m.seg_dmaj * S.sigma_t * m.scale) * 0.9999f;
                if (m.seg_t0 + dtp > m.seg_t1 + (1.0e-5f * fabsf(m.seg_t1) + 1.0e-5f)) { // certainly t >= seg.t1: segment dropped
#endif
                    mode = M_STEP;
                    if (VPT_CUR_T0 >= m.t1) HOT_FINISH();
                    else { m.seg_t0 = VPT_CUR_T0; m.seg_dmaj = VPT_CUR_MAJ; }
                } else stage = ST_EVAL;
            }
#if VPT_HOT_SINGLE_FINISH
            if (fin) VPT_RAY_FINISHED();
#endif
#undef HOT_FINISH
#if !VPT_LOOKAHEAD
            if (mode == M_STEP && stage == ST_HOT) { // one HDDA step, majorant fetch issued   volume.cpp:53-59
                VPT_HDDA_ADVANCE();
                mode = M_PEND;
            }
#endif
        }
        // ------------------------------------------------------------ EVAL: trilinear density + collision   worker.cpp:68-84,148-188
        if (run_eval && stage == ST_EVAL) {
            const float sigma_maj = m.seg_dmaj * S.sigma_t;
            const float dt_m = -vptm::logf_exact(1.0f - pend_u, logtab) / sigma_maj; // sample_exponential  random.hpp:19-22
            const float t = m.seg_t0 + dt_m / m.scale;
            stage = ST_HOT;
            float density = 0.0f;
            V3 pos_idx = mk(0.0f, 0.0f, 0.0f);
            if (t < m.seg_t1) {
                m.seg_t0 = t;
                const V3 p = m.at(t);
                if (COUNT) ++st.density_evals;
                density = trilinear_cached(G, cc, p.x, p.y, p.z);
                if (density > 0.0f) pos_idx = p;
            } else { // overshoot after all: segment dropped, iterator.next()
                mode = M_STEP;
                if (VPT_CUR_T0 >= m.t1) VPT_RAY_FINISHED();
                else { m.seg_t0 = VPT_CUR_T0; m.seg_dmaj = VPT_CUR_MAJ; }
            }
            if (density > 0.0f) {
                if (shadow) { // ratio tracking + Russian roulette
                    const float sigma_n = fmaxf(0.0f, sigma_maj - S.sigma_t * density);
                    float T_ray = CLDF(C_TRAY) * (sigma_n / sigma_maj);
                    if (T_ray <= 0.05f) {
                        if (rng.uniform() < 0.75f) T_ray = 0.0f;
                        else T_ray /= 0.25f;
                    }
                    if (T_ray <= 0.0f) { T_ray = 0.0f; sub = PH_AFTER_SHADOW; stage = ST_SETUP; }
                    CSTF(C_TRAY, T_ray);
                } else {
                    const V3 point = index_to_world(G, pos_idx);
                    VPT_EVENT(1, point.x, point.y, point.z, density, 0.0f, 0.0f); // logger.sampled_point   worker.cpp:146
                    const float p_a = (S.sigma_a * density) / sigma_maj;
                    const float p_s = (S.sigma_s * density) / sigma_maj;
                    const float p_n = fmaxf(1.0f - p_a - p_s, 0.0f);
                    if (S.has_temperature) {
                        const V3 tc = world_to_index(A.temperature, point);
                        if (COUNT) ++st.temperature_evals;
                        CellCache ct;
                        ct.x = CLDI(C_CT); ct.y = CLDI(C_CT + 1); ct.z = CLDI(C_CT + 2); ct.code = cold[C_CT + 3][lane_col];
                        const float temp_adim = trilinear_cached(A.temperature, ct, tc.x, tc.y, tc.z);
                        CSTI(C_CT, ct.x); CSTI(C_CT + 1, ct.y); CSTI(C_CT + 2, ct.z); cold[C_CT + 3][lane_col] = ct.code;
                        const float temp_K = temp_adim * S.temperature_scale + S.temperature_offset;
                        const V3 e = blackbody_xyz(bb, temp_K);
                        const float k = p_a * S.le_scale;
                        CSTF(C_L, CLDF(C_L) + k * e.x); CSTF(C_L + 1, CLDF(C_L + 1) + k * e.y); CSTF(C_L + 2, CLDF(C_L + 2) + k * e.z);
                    }
                    float u = rng.uniform() * (((0.0f + p_n) + p_a) + p_s); // sample_discrete  random.hpp:31-47
                    u -= p_n;
                    if (TRACE && !(u > 0.0f)) VPT_EVENT(2, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f); // logger.null()   worker.cpp:166
                    if (u > 0.0f) {
                        u -= p_a;
                        stage = ST_SETUP;
                        if (u <= 0.0f) { sub = PH_NEXT_PIXEL; VPT_EVENT(4, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f); } // absorption: terminated, no infinite light (logger.absorbed, :185)
                        else {                               // scatter (also the fallback choice)
                            if (COUNT) ++st.scatter_events;
                            CST3(C_POS, point);
#if VPT_SHADOW_IN_EVAL
                            if (S.dist_light_on) { // sample_Ld (worker.cpp:59-66): the shadow ray begins right here; the lane goes straight back to HOT
                                shadow = true; CSTF(C_TRAY, 1.0f);
                                if (COUNT) ++st.shadow_rays;
                                bool ray_done;
                                sub = PH_AFTER_SHADOW; // where the lane goes if the shadow ray misses the volume (T stays 1) or is over at once
                                if (flat_begin_ray_dir<COUNT>(m, G, cc, point, shadow_dir, ray_done, st) && !ray_done) { stage = ST_HOT; mode = M_STEP; }
                            } else { CSTF(C_TRAY, 0.0f); sub = PH_AFTER_SHADOW; } // sample_Ld returns 0 before any draw (:57)
#else
                            if (S.dist_light_on) sub = PH_BEGIN_SHADOW;
                            else { CSTF(C_TRAY, 0.0f); sub = PH_AFTER_SHADOW; } // sample_Ld returns 0 before any draw (:57)
#endif
                        }
                    }
                }
            }
        }
    }
#undef VPT_RAY_FINISHED
#undef VPT_HDDA_ADVANCE
#undef VPT_CUR_T0
#undef VPT_CUR_MAJ
#undef VPT_ADD_INF_LIGHT
#undef VPT_EVENT
#undef CLDF
#undef CLDI
#undef CSTF
#undef CSTI
#undef CLD3
#undef CST3
    if (COUNT) {
        atomicAdd(&A.counters->pixel_samples, (unsigned long long)st.pixel_samples);
        atomicAdd(&A.counters->density_evals, (unsigned long long)st.density_evals);
        atomicAdd(&A.counters->temperature_evals, (unsigned long long)st.temperature_evals);
        atomicAdd(&A.counters->majorant_fetches, (unsigned long long)st.majorant_fetches);
        atomicAdd(&A.counters->rng_draws, draws_total);
        atomicAdd(&A.counters->scatter_events, (unsigned long long)st.scatter_events);
        atomicAdd(&A.counters->shadow_rays, (unsigned long long)st.shadow_rays);
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&A.counters->sched_iters, sd_iters);
            for (int i = 0; i < 3; ++i) { atomicAdd(&A.counters->sched_runs[i], sd_runs[i]); atomicAdd(&A.counters->sched_lanes[i], sd_lanes[i]); }
            atomicAdd(&A.counters->sched_waiting, sd_wait);
            atomicAdd(&A.counters->sched_exited, sd_exit);
        }
    }
}

} // namespace vptd
