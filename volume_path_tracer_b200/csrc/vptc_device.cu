// vptc_device.cu — the device-facing entry points of include/vptc.h: context, volume staging, render,
// parity hooks. Plain CUDA runtime; no torch types cross this boundary.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <memory>
#include <mutex>
#include <chrono>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "device/vpt_kernels.hpp"
#include "host/bake.hpp"
#include "host/scene_prep.hpp"
#include "vptc.h"
#include "vptc_internal.hpp"

using namespace vptd;

struct DeviceGrid {
    GridDev dev{};
    void *cells = nullptr, *slots = nullptr, *tile_value = nullptr, *tile_majorant = nullptr, *bricks = nullptr, *slot_majorant = nullptr;
    size_t bytes = 0;
    void release()
    {
        cudaFree(cells); cudaFree(slots); cudaFree(tile_value); cudaFree(tile_majorant); cudaFree(bricks); cudaFree(slot_majorant);
        cells = slots = tile_value = tile_majorant = bricks = slot_majorant = nullptr; bytes = 0;
    }
};

struct vptc_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    std::string error;
    DeviceGrid density, temperature;
    bool has_volume = false, has_temperature = false;
    float* bb_table = nullptr;
    unsigned long long* queue = nullptr;
    CountersDev* counters = nullptr;
    bool count = false;
    int variant = 2; // render kernel variant (vptc_set_kernel_variant)
    int sched_thresh[4] = {12, 8, 8, 16};
    int pool_params[4] = {3, 8, 28, 28}; // variant 3: HOT passes per scheduling decision; queued slots needed to run HOT / EVAL / SETUP
    float last_ms = 0.0f;
    vptc_counters last_counters{};
    int num_sms = 0;
    // ---- range ring (device/layout.hpp RangeDesc): every render call of the default kernel is published here ----
    cudaStream_t ctl = nullptr;      // non-blocking stream: ring publication, film zeroing and staging copies run BESIDE the render kernel
    cudaEvent_t ev_ctl = nullptr;
    RangeDesc* d_ring = nullptr; unsigned int* d_tail = nullptr; unsigned int* d_exited = nullptr;
    RangeDesc* h_ring = nullptr; unsigned int* h_tail = nullptr; // pinned staging, one slot per ring slot
    unsigned int* h_done = nullptr; unsigned int* d_done = nullptr; // mapped pinned completion flags (host pointer / device alias)
    unsigned int n_ranges = 0;       // ranges published so far (absolute range number of the next one)
    // ---- tile order: full-frame ranges are walked longest-tiles-first once the per-tile job times of this scene are known (prepare_tile_order) ----
    uint32_t* d_tile_cost = nullptr; uint32_t* d_tile_order = nullptr; size_t order_tiles = 0, order_capacity = 0;
    SceneDev order_scene{}; bool have_order = false; unsigned long long cost_jobs = 0;
    std::vector<uint32_t> h_order;   // host copy of the sorted tile list: tile STRIPES take their own tiles out of it, in the same order
    uint32_t* d_stripe_order = nullptr; size_t stripe_capacity = 0;
    unsigned long long stripe_begin = ~0ull, stripe_count = 0; uint32_t stripe_stride = 0; // which stripe d_stripe_order holds
    bool order_enabled = true;       // VPT_TILE_ORDER=0 in the environment: natural order (row-major tiles, wave-major), as TileProvider hands them out
    // open series of asynchronous launches (vptc_render_device_async / vptc_render_async ... vptc_wait)
    bool series_open = false; int series_launches = 0;
    bool series_has_plain = false;   // the open series holds a launch that does not read tile lists (see enqueue_render)
    SceneDev series_scene{};         // a call joins the running series only when it renders the same scene
    cudaEvent_t ev_series = nullptr; // recorded before the first launch of the series
    // host-film path: two device films + pinned staging; at most one call's film is still on the device
    float* film[2] = {nullptr, nullptr}; size_t film_floats = 0;
    float* pinned = nullptr; size_t pinned_floats = 0;
    float* pinned_zero = nullptr;    // as large as a film, all zero: films are cleared by an H2D COPY (copy engine), because a memset is a kernel
                                     // and would wait behind the persistent render kernel that fills every SM
    struct Pending { bool active = false; unsigned int rid = 0; float* host_film = nullptr; size_t n = 0; int buf = 0; } pending;
    int next_buf = 0;
    cudaEvent_t ev_launch[2] = {nullptr, nullptr}; // after the kernel launched for the call using film[b]
};

namespace {

int fail(vptc_ctx* c, int code, const std::string& msg)
{
    if (c) c->error = msg;
    vptc_internal::set_global_error(msg);
    return code;
}
#define CU(call)                                                                                                   \
    do {                                                                                                           \
        cudaError_t e_ = (call);                                                                                   \
        if (e_ != cudaSuccess) return fail(ctx, VPTC_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// Stages a baked grid into a FRESH DeviceGrid; on any failure the partial allocations are released and `g` is left empty.
int upload(vptc_ctx* ctx, const vpt_host::BakedGrid& b, DeviceGrid& g)
{
    g.release();
    g.dev = b.header;
    auto up = [&](void** dst, const void* src, size_t bytes) -> cudaError_t {
        cudaError_t e = cudaMalloc(dst, bytes ? bytes : 16);
        if (e != cudaSuccess) return e;
        g.bytes += bytes;
        return bytes ? cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream) : cudaSuccess;
    };
    // The brick stage on the device (device/vpt_bake.cu): the caller's serialized grid is staged as it is, one block per leaf fills its apron
    // brick and reduces the patched maximum, a second pass folds the maxima into the slot table uploaded just before. Temporaries are freed here.
    void *d_raw = nullptr, *d_max = nullptr, *d_order = nullptr;
    auto device_brick_stage = [&]() -> cudaError_t {
        const size_t nb = b.header.n_bricks, brick_bytes = nb * size_t(BRICK_STRIDE) * 4;
        cudaError_t e = cudaMalloc(&g.bricks, brick_bytes ? brick_bytes : 16);
        if (e != cudaSuccess) return e;
        g.bytes += brick_bytes;
        if (nb == 0) return cudaSuccess;
        if ((e = cudaMemsetAsync(g.bricks, 0, brick_bytes, ctx->stream)) != cudaSuccess) return e; // the padding floats of a brick are zero, as in the host baker
        if ((e = cudaMalloc(&d_raw, b.raw_bytes)) != cudaSuccess || (e = cudaMalloc(&d_max, nb * 4)) != cudaSuccess) return e;
        if ((e = cudaMemcpyAsync(d_raw, b.raw, b.raw_bytes, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return e;
        if (!b.brick_of.empty()) {
            if ((e = cudaMalloc(&d_order, nb * 4)) != cudaSuccess) return e;
            if ((e = cudaMemcpyAsync(d_order, b.brick_of.data(), nb * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) return e;
        }
        return launch_bake_bricks(d_raw, nb, static_cast<const uint32_t*>(d_order), b.patch_majorants, static_cast<float*>(g.bricks), static_cast<float*>(d_max),
                                  g.slots, b.slots.size() / 2, ctx->stream);
    };
    struct Temps { void *&a, *&b, *&c; ~Temps() { cudaFree(a); cudaFree(b); cudaFree(c); } } temps{d_raw, d_max, d_order}; // after the stream synchronisation below
    cudaError_t e;
    if ((e = up(&g.cells, b.cells.data(), b.cells.size() * 4)) != cudaSuccess || (e = up(&g.slots, b.slots.data(), b.slots.size() * 4)) != cudaSuccess ||
        (e = up(&g.tile_value, b.tile_value.data(), b.tile_value.size() * 4)) != cudaSuccess ||
        (e = up(&g.tile_majorant, b.tile_majorant.data(), b.tile_majorant.size() * 4)) != cudaSuccess ||
        (e = b.device_bricks ? device_brick_stage() : up(&g.bricks, b.bricks.data(), b.bricks.size() * 4)) != cudaSuccess ||
        (!b.slot_majorant.empty() && (e = up(&g.slot_majorant, b.slot_majorant.data(), b.slot_majorant.size() * 4)) != cudaSuccess) ||
        (e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) {
        cudaStreamSynchronize(ctx->stream); // the host vectors must outlive any copy still in flight
        g.release();
        cudaGetLastError();
        return fail(ctx, VPTC_ERR_CUDA, std::string("staging the baked volume to the device: ") + cudaGetErrorString(e));
    }
    g.dev.cells = static_cast<const uint32_t*>(g.cells);
    g.dev.slots = static_cast<const Slot*>(g.slots);
    g.dev.tile_value = static_cast<const float*>(g.tile_value);
    g.dev.tile_majorant = static_cast<const float*>(g.tile_majorant);
    g.dev.bricks = static_cast<const float*>(g.bricks);
    g.dev.slot_majorant = static_cast<const float*>(g.slot_majorant);
    return VPTC_OK;
}

int check_scene(vptc_ctx* ctx, const vptc_scene* s)
{
    if (!s) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "scene is null");
    if (s->output_size[0] <= 0 || s->output_size[1] <= 0 || s->tile_size[0] <= 0 || s->tile_size[1] <= 0)
        return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "output_size and tile_size must be positive");
    // The kernels hold pixel coordinates in 32-bit ints and the film is W*H*16 bytes: bound every factor BEFORE multiplying, so that no
    // product below can wrap (a [2^31, 2^31] frame used to wrap the film size to 0 and launch on a null film).
    constexpr int64_t kMaxSide = int64_t(1) << 20, kMaxPixels = int64_t(1) << 32; // 64 Gi B of film at most
    if (s->output_size[0] > kMaxSide || s->output_size[1] > kMaxSide || s->output_size[0] * s->output_size[1] > kMaxPixels)
        return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "output_size too large (each side <= 2^20, at most 2^32 pixels)");
    if (s->tile_size[0] > kMaxSide || s->tile_size[1] > kMaxSide || s->tile_size[0] * s->tile_size[1] > (int64_t(1) << 20))
        return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "tile too large (at most 2^20 pixels per tile)");
    return VPTC_OK;
}

// A tile slice must stay inside the frame: contiguous [begin, begin + count), or — with a row stride — whole rows of tiles begin/ntx, +stride, ...
bool range_inside(const SceneDev& sd, unsigned long long begin, unsigned long long count, uint32_t row_stride)
{
    if (row_stride <= 1) return begin <= sd.ntiles && count <= sd.ntiles - begin;
    const unsigned long long ntx = (unsigned long long)sd.ntx, nty = (unsigned long long)sd.nty;
    if (begin % ntx != 0 || count % ntx != 0) return false; // interleaved stripes are made of whole rows
    const unsigned long long row0 = begin / ntx, rows = count / ntx;
    return rows == 0 || (row0 < nty && row0 + (rows - 1) * row_stride < nty);
}

// What one render call covers. tile_count == 0 means every tile of the frame.
struct RangeSpec {
    float* film = nullptr;
    unsigned long long n_jobs = 0, tile_begin = 0, tile_count = 0;
    uint32_t first_wave = 1, wave_stride = 1, tile_row_stride = 1;
    // statistics launches (adaptive rendering): explicit tile list, second-moment plane; run the STATS instantiation of the default kernel
    const uint32_t* tile_list = nullptr; float* sum_y2 = nullptr; bool stats = false;
    uint32_t* tile_cost = nullptr; // where the kernel adds each finished job's duration (per tile), or null
};

// Publish a range into the device ring (descriptor first, then the tail, both on the control stream, which runs beside any render kernel
// already in flight: that kernel's lanes carry on into the new range as soon as they see the tail move). -> absolute range number in *rid.
int publish_range(vptc_ctx* ctx, const RangeSpec& r, unsigned int* rid)
{
    const unsigned int id = ctx->n_ranges, slot = id % RING_SLOTS;
    RangeDesc& d = ctx->h_ring[slot];
    d.counter = 0; d.done = 0; d.n_jobs = r.n_jobs; d.film = r.film;
    d.tile_begin = r.tile_begin; d.tile_count = r.tile_count;
    d.first_wave = r.first_wave; d.wave_stride = r.wave_stride; d.rid = id; d.tile_row_stride = std::max(1u, r.tile_row_stride);
    d.tile_list = r.tile_list; d.sum_y2 = r.sum_y2;
    d.tile_cost = r.tile_cost; d.n_waves = r.tile_count ? uint32_t(r.n_jobs / r.tile_count) : 0u; d.pad_ = 0;
    ctx->h_tail[slot] = id + 1;
    ctx->h_done[slot] = 0;
    CU(cudaMemcpyAsync(ctx->d_ring + slot, &d, sizeof(RangeDesc), cudaMemcpyHostToDevice, ctx->ctl));
    CU(cudaMemcpyAsync(ctx->d_tail, ctx->h_tail + slot, sizeof(unsigned int), cudaMemcpyHostToDevice, ctx->ctl));
    CU(cudaEventRecord(ctx->ev_ctl, ctx->ctl));
    CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_ctl, 0)); // this call's own launch starts after its range is visible
    ctx->n_ranges = id + 1;
    *rid = id;
    return VPTC_OK;
}

// Longest tiles first. A (tile, wave) job is rendered serially by one lane, so a launch ends with a tail as long as its last jobs; jobs of
// the expensive tiles (through the core of a cloud) take 100x the time of sky tiles. The kernel records how long each tile's jobs take; as soon
// as one whole wave of a scene has been timed (and nothing is in flight), the tiles are sorted by that time and every later full-frame range
// of the scene is walked tile-major in that order: all waves of the expensive tiles start at once, the cheap ones fill the tail. Job ids —
// and with them every sample — are unchanged; only which lane renders which job when.
int sort_tiles_if_timed(vptc_ctx* ctx);
int prepare_tile_order(vptc_ctx* ctx, const SceneDev& sd)
{
    const size_t nt = size_t(sd.ntiles);
    if (ctx->order_tiles != nt || std::memcmp(&ctx->order_scene, &sd, sizeof(SceneDev)) != 0) { // another frame / camera / medium: start over
        CU(cudaStreamSynchronize(ctx->stream));
        ctx->have_order = false; ctx->cost_jobs = 0;
        if (ctx->order_capacity < nt) { // a moving camera keeps the frame size: the tables are reallocated only when the tile count grows
            cudaFree(ctx->d_tile_cost); cudaFree(ctx->d_tile_order);
            ctx->d_tile_cost = ctx->d_tile_order = nullptr; ctx->order_tiles = 0; ctx->order_capacity = 0;
            CU(cudaMalloc(&ctx->d_tile_cost, nt * 4));
            CU(cudaMalloc(&ctx->d_tile_order, nt * 4));
            ctx->order_capacity = nt;
        }
        CU(cudaMemset(ctx->d_tile_cost, 0, nt * 4));
        ctx->order_tiles = nt;
        std::memcpy(&ctx->order_scene, &sd, sizeof(SceneDev));
    }
    return sort_tiles_if_timed(ctx);
}
// One whole wave of the scene has been timed and nothing is in flight: sort the tiles by the time their jobs took.
int sort_tiles_if_timed(vptc_ctx* ctx)
{
    const size_t nt = ctx->order_tiles;
    if (!ctx->have_order && nt && ctx->cost_jobs >= nt && !ctx->series_open) {
        CU(cudaStreamSynchronize(ctx->stream));
        std::vector<uint32_t> cost(nt), order(nt);
        CU(cudaMemcpy(cost.data(), ctx->d_tile_cost, nt * 4, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < nt; ++i) order[i] = uint32_t(i);
        std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return cost[a] > cost[b]; });
        CU(cudaMemcpy(ctx->d_tile_order, order.data(), nt * 4, cudaMemcpyHostToDevice));
        ctx->h_order.swap(order);
        ctx->stripe_begin = ~0ull;
        ctx->have_order = true;
    }
    return VPTC_OK;
}

// A stripe of tiles (multi-GPU tile sharding: contiguous, or every row_stride-th row of tiles) of a scene whose full-frame order is known:
// the stripe's own tiles, in that order. One list is cached per context (a rank renders the same stripe call after call); another stripe while
// a series is in flight keeps the natural order. -> true when ctx->d_stripe_order holds the list for this range.
bool prepare_stripe_order(vptc_ctx* ctx, const SceneDev& sd, unsigned long long begin, unsigned long long count, uint32_t row_stride)
{
    if (ctx->order_tiles != size_t(sd.ntiles) || std::memcmp(&ctx->order_scene, &sd, sizeof(SceneDev)) != 0 || count == 0) return false;
    if (!ctx->have_order && sort_tiles_if_timed(ctx) != VPTC_OK) return false; // a rank that renders only its stripe after the first full-frame waves
    if (!ctx->have_order) return false;
    if (ctx->stripe_begin == begin && ctx->stripe_count == count && ctx->stripe_stride == row_stride) return true;
    if (ctx->series_open) return false; // the cached list may still be read by ranges in flight
    const unsigned long long ntx = (unsigned long long)sd.ntx, row0 = begin / ntx, rows = count / ntx;
    std::vector<uint32_t> list;
    list.reserve(size_t(count));
    for (uint32_t t : ctx->h_order) {
        bool in;
        if (row_stride <= 1) in = t >= begin && t - begin < count;
        else { const unsigned long long row = t / ntx; in = row >= row0 && (row - row0) % row_stride == 0 && (row - row0) / row_stride < rows; }
        if (in) list.push_back(t);
    }
    if (list.size() != size_t(count)) return false;
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) return false;
    if (ctx->stripe_capacity < list.size()) {
        cudaFree(ctx->d_stripe_order); ctx->d_stripe_order = nullptr; ctx->stripe_capacity = 0; ctx->stripe_begin = ~0ull;
        if (cudaMalloc(&ctx->d_stripe_order, list.size() * 4) != cudaSuccess) { cudaGetLastError(); return false; }
        ctx->stripe_capacity = list.size();
    }
    if (cudaMemcpy(ctx->d_stripe_order, list.data(), list.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess) { cudaGetLastError(); ctx->stripe_begin = ~0ull; return false; }
    ctx->stripe_begin = begin; ctx->stripe_count = count; ctx->stripe_stride = row_stride;
    return true;
}

// Enqueue the megakernel for one range on the render stream; does not wait. `full_grid`: a launch that later calls may join needs every SM.
int enqueue_render(vptc_ctx* ctx, const vptc_scene* scene, RenderArgs& A, RangeSpec r, bool trace, bool full_grid, unsigned int* rid_out = nullptr)
{
    vpt_host::prepare_scene(*scene, ctx->has_temperature, A.scene);
    A.density = ctx->density.dev;
    A.temperature = ctx->has_temperature ? ctx->temperature.dev : ctx->density.dev;
    A.bb_table = ctx->bb_table;
    A.queue = ctx->queue;
    A.counters = ctx->counters;
    for (int i = 0; i < 4; ++i) A.sched_thresh[i] = std::max(1, ctx->sched_thresh[i]); // the kernel relies on >= 1
    for (int i = 0; i < 4; ++i) A.pool_params[i] = std::max(1, ctx->pool_params[i]);
    if (r.tile_count == 0) { r.tile_begin = 0; r.tile_count = A.scene.ntiles; }
    if (!trace && !range_inside(A.scene, r.tile_begin, r.tile_count, r.tile_row_stride)) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "tile range outside the frame's tiles");
    // variant 3 packs pixel coordinates into 16-bit halves of its slot words: larger frames / tiles take variant 2
    int variant = (ctx->variant == 3 && (scene->output_size[0] > 65535 || scene->output_size[1] > 65535 || scene->tile_size[0] > 65535 || scene->tile_size[1] > 65535)) ? 2 : ctx->variant;
    if (variant != 2 && !trace && (r.tile_count != A.scene.ntiles || r.tile_row_stride > 1 || r.stats)) variant = 2; // tile stripes / lists exist in the default kernel only
    const bool stats = r.stats && !trace && !ctx->count;
    int ord = 0; // tile-order instantiation of the default kernel: 0 none, 1 learning the tiles' costs, 2 ordered (decided below)
    const bool orderable = variant == 2 && !trace && !r.stats && !ctx->count && ctx->order_enabled && r.tile_begin == 0 && r.tile_count == A.scene.ntiles &&
                           r.tile_row_stride <= 1 && A.scene.ntiles >= 2 && A.scene.ntiles <= 0xffffffffull && r.n_jobs % r.tile_count == 0;
    if (orderable) {
        if (int rc = prepare_tile_order(ctx, A.scene)) return rc;
        ord = ctx->have_order ? 2 : 1;
        // A range with a tile list is numbered tile-major, and only the ORD = 2 instantiation reads it that way: lanes of a plain launch of the
        // same series could carry on into it and number its jobs wave-major. Once a series holds a plain launch, the rest of it stays plain.
        if (ord == 2 && ctx->series_open && ctx->series_has_plain) ord = 0;
    }
    // a stripe of a scene whose order is known: its own tiles, longest first (no learning on stripes: the costs come from full-frame calls)
    const bool stripe = variant == 2 && !trace && !r.stats && !ctx->count && ctx->order_enabled && !orderable && r.tile_count > 1 && r.tile_count < A.scene.ntiles &&
                        A.scene.ntiles <= 0xffffffffull && r.n_jobs % r.tile_count == 0 && !(ctx->series_open && ctx->series_has_plain);
    const uint32_t* stripe_list = nullptr;
    if (stripe && prepare_stripe_order(ctx, A.scene, r.tile_begin, r.tile_count, std::max(1u, r.tile_row_stride))) { ord = 2; stripe_list = ctx->d_stripe_order; }
    if (variant == 2 && !trace && ord != 2) ctx->series_has_plain = true; // reset when the series closes (wait_all) / a blocking call finishes
    int per_sm = stats ? render_stats_blocks_per_sm() : ord ? render_ordered_blocks_per_sm(ord) : render_blocks_per_sm(variant, trace, ctx->count);
    if (const char* v = std::getenv("VPT_MAX_BLOCKS_PER_SM")) per_sm = std::min(per_sm, std::max(1, std::atoi(v))); // occupancy experiments
    if (per_sm <= 0) return fail(ctx, VPTC_ERR_CUDA, "render kernel cannot be resident on this device (occupancy query returned 0)");
    if (ctx->count) CU(cudaMemsetAsync(ctx->counters, 0, sizeof(CountersDev), ctx->stream));
    if (variant == 2) {
        if (ord == 2) r.tile_list = stripe_list ? stripe_list : ctx->d_tile_order;
        else if (ord == 1) { r.tile_cost = ctx->d_tile_cost; ctx->cost_jobs += r.n_jobs; }
        unsigned int rid = 0;
        if (int rc = publish_range(ctx, r, &rid)) return rc;
        A.ring = ctx->d_ring; A.ring_tail = ctx->d_tail; A.range_begin = rid; A.host_done = ctx->d_done;
        if (rid_out) *rid_out = rid;
    } else {
        A.film = r.film; A.n_jobs = r.n_jobs; A.first_wave = r.first_wave; A.wave_stride = r.wave_stride;
        CU(cudaMemsetAsync(ctx->queue, 0, sizeof(unsigned long long), ctx->stream));
    }
    // persistent grid: every SM filled to its occupancy limit, never more blocks than the jobs of a stand-alone call need
    const unsigned long long per_block = (unsigned long long)(render_jobs_per_block(variant));
    const unsigned long long want = (r.n_jobs + per_block - 1) / per_block, full = (unsigned long long)(ctx->num_sms) * per_sm;
    const int blocks = int(full_grid ? full : std::min<unsigned long long>(want ? want : 1ULL, full));
    A.exited = ctx->d_exited; A.exit_limit = unsigned(blocks) * unsigned(per_block) / 2u;
    CU(cudaMemsetAsync(ctx->d_exited, 0, sizeof(unsigned int), ctx->stream)); // same stream: runs between the previous launch and this one
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    CU(stats ? launch_render_stats(A, blocks, ctx->stream) : ord ? launch_render_ordered(A, ord, blocks, ctx->stream) : launch_render(A, variant, trace, ctx->count, blocks, ctx->stream));
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    return VPTC_OK;
}

int read_counters(vptc_ctx* ctx);
// Wait for the launch enqueue_render made; fills last_ms / counters.
int finish_render(vptc_ctx* ctx)
{
    CU(cudaEventSynchronize(ctx->ev1));
    CU(cudaEventElapsedTime(&ctx->last_ms, ctx->ev0, ctx->ev1));
    return read_counters(ctx);
}
int read_counters(vptc_ctx* ctx)
{
    if (ctx->count) {
        CountersDev h;
        CU(cudaMemcpy(&h, ctx->counters, sizeof(h), cudaMemcpyDeviceToHost));
        ctx->last_counters = {h.pixel_samples, h.density_evals, h.temperature_evals, h.majorant_fetches, h.rng_draws, h.scatter_events, h.shadow_rays};
        if (std::getenv("VPT_SCHED_STATS") && h.sched_iters) { // development aid: where the lanes of a warp spend their iterations
            const double it = double(h.sched_iters);
            std::fprintf(stderr, "[sched] warp-iterations %.4g; per iteration: SETUP run %.3f (%.1f lanes when run), HOT run %.3f (%.1f), EVAL run %.3f (%.1f); waiting %.2f lanes, exited %.2f lanes\n",
                         it, h.sched_runs[0] / it, double(h.sched_lanes[0]) / double(h.sched_runs[0] ? h.sched_runs[0] : 1), h.sched_runs[1] / it,
                         double(h.sched_lanes[1]) / double(h.sched_runs[1] ? h.sched_runs[1] : 1), h.sched_runs[2] / it,
                         double(h.sched_lanes[2]) / double(h.sched_runs[2] ? h.sched_runs[2] : 1), h.sched_waiting / it, h.sched_exited / it);
        }
    }
    return VPTC_OK;
}

int run_render(vptc_ctx* ctx, const vptc_scene* scene, RenderArgs& A, const RangeSpec& r, bool trace)
{
    if (int rc = enqueue_render(ctx, scene, A, r, trace, /*full_grid=*/false)) return rc;
    return finish_render(ctx);
}

// Add `src` into `dst` (n floats) on a few host threads: the reference's film accumulates (worker.cpp:203-204), so a call's device film is
// ADDED into the caller's host film. 8.3 M floats at 1080p, 33 M at 4K: memory bound, ~1 ms per 10 M floats on 8 threads.
void host_add(float* dst, const float* src, size_t n)
{
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const size_t nt = std::min<size_t>({size_t(8), size_t(hw), n / (size_t(1) << 18) + 1});
    auto work = [=](size_t a, size_t b) { for (size_t i = a; i < b; ++i) dst[i] += src[i]; };
    if (nt <= 1) { work(0, n); return; }
    std::vector<std::thread> th;
    const size_t chunk = (n + nt - 1) / nt;
    for (size_t t = 1; t < nt; ++t) th.emplace_back(work, std::min(n, t * chunk), std::min(n, (t + 1) * chunk));
    work(0, std::min(n, chunk));
    for (auto& t : th) t.join();
}

// The film of the pending host-film call is complete on the device (its range's completion flag is up, or the launch made for it has
// finished): bring it to the host and add it into that call's film. The next call's kernel keeps running meanwhile.
int collect_pending(vptc_ctx* ctx)
{
    if (!ctx->pending.active) return VPTC_OK;
    vptc_ctx::Pending p = ctx->pending;
    ctx->pending.active = false;
    const unsigned int slot = p.rid % RING_SLOTS;
    for (unsigned spin = 0;; ++spin) {
        if (p.rid != ~0u && *reinterpret_cast<volatile unsigned int*>(ctx->h_done + slot) == p.rid + 1u) break;
        const cudaError_t q = cudaEventQuery(ctx->ev_launch[p.buf]);
        if (q == cudaSuccess) break; // the kernel launched for this range has exited: the range is complete
        if (q != cudaErrorNotReady) return fail(ctx, VPTC_ERR_CUDA, std::string("render kernel failed: ") + cudaGetErrorString(q));
        if (spin > 64) std::this_thread::sleep_for(std::chrono::microseconds(50));
    }
    CU(cudaMemcpyAsync(ctx->pinned, ctx->film[p.buf], p.n * sizeof(float), cudaMemcpyDeviceToHost, ctx->ctl));
    CU(cudaStreamSynchronize(ctx->ctl));
    host_add(p.host_film, ctx->pinned, p.n);
    return VPTC_OK;
}

// Close the open series of asynchronous calls: collect the pending host film, wait for every launch, total device time.
int wait_all(vptc_ctx* ctx)
{
    int rc = collect_pending(ctx);
    if (ctx->series_open) {
        ctx->series_open = false; ctx->series_launches = 0; ctx->series_has_plain = false;
        cudaError_t e = cudaEventSynchronize(ctx->ev1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ctx->last_ms, ctx->ev_series, ctx->ev1);
        if (e != cudaSuccess && rc == VPTC_OK) rc = fail(ctx, VPTC_ERR_CUDA, std::string("vptc_wait: ") + cudaGetErrorString(e));
        if (rc == VPTC_OK) rc = read_counters(ctx);
    }
    return rc;
}

// Common front of the asynchronous render calls: argument checks, joins (or opens) the series, enqueues one launch for the range.
int enqueue_async(vptc_ctx* ctx, const vptc_scene* scene, const vptc_job_range* jr, float* d_film, unsigned int* rid, bool* launched = nullptr)
{
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    if (int rc = check_scene(ctx, scene)) return rc;
    if (!jr || !d_film || jr->first_wave < 1 || jr->wave_stride < 1) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "render: film or range null, first_wave < 1 or wave_stride < 1");
    CU(cudaSetDevice(ctx->device));
    SceneDev sd; vpt_host::prepare_scene(*scene, ctx->has_temperature, sd);
    const bool joinable = ctx->variant == 2 && !ctx->count;
    // a call may ride on the running kernel only if that kernel renders the same scene; and the ring must never wrap onto a live range
    if (ctx->series_open && (!joinable || std::memcmp(&sd, &ctx->series_scene, sizeof(SceneDev)) != 0 || ctx->series_launches >= int(RING_SLOTS) / 2))
        if (int rc = wait_all(ctx)) return rc;
    RangeSpec r;
    r.film = d_film; r.first_wave = jr->first_wave; r.wave_stride = jr->wave_stride;
    r.tile_begin = jr->tile_begin; r.tile_count = jr->tile_count ? jr->tile_count : sd.ntiles;
    r.tile_row_stride = std::max(1u, jr->tile_row_stride);
    if (!range_inside(sd, r.tile_begin, r.tile_count, r.tile_row_stride)) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "render: tile range outside the frame's tiles (a row stride takes whole rows of tiles)");
    r.n_jobs = r.tile_count * (unsigned long long)jr->n_waves;
    if (launched) *launched = false;
    if (rid) *rid = ~0u; // stays ~0 for the kernels that do not read the ring (variants 0, 1, 3)
    if (r.n_jobs == 0) return VPTC_OK;
    if (!ctx->series_open) {
        ctx->series_has_plain = false;
        CU(cudaEventRecord(ctx->ev_series, ctx->stream));
        std::memset(&ctx->series_scene, 0, sizeof(SceneDev)); // padding bytes take part in the memcmp above
        ctx->series_scene = sd;
    }
    RenderArgs A{};
    unsigned int id = ~0u;
    if (int rc = enqueue_render(ctx, scene, A, r, /*trace=*/false, /*full_grid=*/joinable, &id)) return rc;
    ctx->series_open = true; ++ctx->series_launches;
    if (rid) *rid = id;
    if (launched) *launched = true;
    return VPTC_OK;
}
// fix_majorants_for_interpolation + the brick layout, on the host (csrc/host/bake.cpp). A corrupt buffer is VPTC_ERR_BAD_GRID.
int bake_volume(vptc_ctx* ctx, const void* density_grid, size_t density_bytes, const void* temperature_grid, size_t temperature_bytes,
                vpt_host::BakedGrid& b, vpt_host::BakedGrid& t, bool& has_temperature)
{
    has_temperature = false;
    try {
        // the brick stage (97 % of the bake) runs on the device unless asked otherwise; the split-majorant experiment builds its table on the host
        const bool device_bricks = !std::getenv("VPT_HOST_BAKE") && !VPT_SPLIT_MAJ;
        vpt_host::bake_grid(density_grid, density_bytes, /*patch_majorants=*/true, b, device_bricks);
        if (temperature_grid && temperature_bytes) {
            vpt_host::bake_grid(temperature_grid, temperature_bytes, /*patch_majorants=*/false, t, device_bricks);
            has_temperature = true;
        }
    } catch (const std::exception& e) {
        return fail(ctx, VPTC_ERR_BAD_GRID, std::string("vptc_set_volume: ") + e.what());
    }
    return VPTC_OK;
}

// Stage baked grids into FRESH device grids and swap only when everything succeeded: a failed call (out of device memory on a larger
// second volume) leaves the context rendering the previous volume, never a half-released one.
int stage_volume(vptc_ctx* ctx, const vpt_host::BakedGrid& b, const vpt_host::BakedGrid* t)
{
    CU(cudaSetDevice(ctx->device));
    if (int rc = wait_all(ctx)) return rc; // nothing may still be rendering the volume about to be replaced
    CU(cudaStreamSynchronize(ctx->stream));
    DeviceGrid new_density, new_temperature;
    if (int rc = upload(ctx, b, new_density)) return rc;
    if (t) if (int rc = upload(ctx, *t, new_temperature)) { new_density.release(); return rc; }
    ctx->density.release(); ctx->temperature.release();
    ctx->density = new_density; ctx->temperature = new_temperature;
    ctx->has_temperature = t != nullptr;
    ctx->has_volume = true;
    ctx->order_tiles = 0; ctx->have_order = false; // another volume: the tiles' costs are learnt afresh
    return VPTC_OK;
}

} // namespace

extern "C" {

const char* vptc_last_error(const vptc_ctx* ctx) { return ctx ? ctx->error.c_str() : vptc_internal::global_error(); }

int vptc_create(vptc_ctx** out, int device)
{
    vptc_ctx* ctx = nullptr;
    if (!out) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_create: out is null");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(nullptr, VPTC_ERR_CUDA, std::string("no CUDA device available (") + cudaGetErrorString(e) + "); this library has no CPU fallback");
    if (device < 0 || device >= n) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_create: device index out of range");
    std::unique_ptr<vptc_ctx> c(new vptc_ctx);
    ctx = c.get();
    ctx->device = device;
    CU(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    ctx->num_sms = prop.multiProcessorCount;
    if (const char* v = std::getenv("VPT_L2_FETCH_GRANULARITY")) // 32 / 64 / 128 bytes (profiles/r02_fetch_layout.txt); a hint, device-wide
        CU(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, size_t(std::max(32, std::atoi(v)))));
    CU(cudaStreamCreate(&ctx->stream)); // blocking stream: ordered after work on the legacy default stream (torch)
    CU(cudaEventCreate(&ctx->ev0));
    CU(cudaEventCreate(&ctx->ev1));
    CU(cudaMalloc(&ctx->bb_table, 1504 * sizeof(float))); // +1 zero entry: the reference reads bb_xyz[500] for 49 800 K < T < 49 900 K
    CU(cudaMemset(ctx->bb_table, 0, 1504 * sizeof(float)));
    CU(cudaMalloc(&ctx->queue, sizeof(unsigned long long)));
    CU(cudaMalloc(&ctx->counters, sizeof(CountersDev)));
    CU(cudaStreamCreateWithFlags(&ctx->ctl, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&ctx->ev_ctl, cudaEventDisableTiming));
    CU(cudaEventCreate(&ctx->ev_series));
    for (int b = 0; b < 2; ++b) CU(cudaEventCreateWithFlags(&ctx->ev_launch[b], cudaEventDisableTiming));
    CU(cudaMalloc(&ctx->d_ring, sizeof(RangeDesc) * RING_SLOTS));
    CU(cudaMemset(ctx->d_ring, 0, sizeof(RangeDesc) * RING_SLOTS));
    CU(cudaMalloc(&ctx->d_exited, sizeof(unsigned int)));
    CU(cudaMalloc(&ctx->d_tail, sizeof(unsigned int)));
    CU(cudaMemset(ctx->d_tail, 0, sizeof(unsigned int)));
    CU(cudaMallocHost(&ctx->h_ring, sizeof(RangeDesc) * RING_SLOTS));
    CU(cudaMallocHost(&ctx->h_tail, sizeof(unsigned int) * RING_SLOTS));
    CU(cudaHostAlloc(&ctx->h_done, sizeof(unsigned int) * RING_SLOTS, cudaHostAllocMapped));
    std::memset(ctx->h_done, 0, sizeof(unsigned int) * RING_SLOTS);
    CU(cudaHostGetDevicePointer(&ctx->d_done, ctx->h_done, 0));
    std::vector<float> bb(1500);
    vpt_host::compute_blackbody_table(bb.data()); // init_blackbody_radiation_xyz (main.cpp:32)
    CU(cudaMemcpy(ctx->bb_table, bb.data(), 1500 * sizeof(float), cudaMemcpyHostToDevice));
    float lambda5[471];
    vpt_host::lambda5_table(lambda5);
    CU(upload_cie_table(vpt_host::cie_table(), lambda5));
    if (const char* v = std::getenv("VPT_TILE_ORDER")) ctx->order_enabled = std::atoi(v) != 0;
    if (const char* v = std::getenv("VPT_KERNEL_VARIANT")) ctx->variant = std::min(3, std::max(0, std::atoi(v)));
    if (const char* v = std::getenv("VPT_POOL_PARAMS")) // "HOT passes,HOT slots,EVAL slots,SETUP slots" (tuning aid)
        std::sscanf(v, "%d,%d,%d,%d", &ctx->pool_params[0], &ctx->pool_params[1], &ctx->pool_params[2], &ctx->pool_params[3]);
    if (const char* v = std::getenv("VPT_SCHED_THRESH")) // "setup,dda,sample,eval" lanes (tuning aid)
        std::sscanf(v, "%d,%d,%d,%d", &ctx->sched_thresh[0], &ctx->sched_thresh[1], &ctx->sched_thresh[2], &ctx->sched_thresh[3]);
    *out = c.release();
    return VPTC_OK;
}

void vptc_destroy(vptc_ctx* ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    ctx->density.release(); ctx->temperature.release();
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->ctl) cudaStreamSynchronize(ctx->ctl);
    cudaFree(ctx->bb_table); cudaFree(ctx->queue); cudaFree(ctx->counters); cudaFree(ctx->film[0]); cudaFree(ctx->film[1]);
    cudaFree(ctx->d_ring); cudaFree(ctx->d_tail); cudaFree(ctx->d_exited);
    cudaFree(ctx->d_tile_cost); cudaFree(ctx->d_tile_order); cudaFree(ctx->d_stripe_order);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    if (ctx->pinned_zero) cudaFreeHost(ctx->pinned_zero);
    if (ctx->h_ring) cudaFreeHost(ctx->h_ring);
    if (ctx->h_tail) cudaFreeHost(ctx->h_tail);
    if (ctx->h_done) cudaFreeHost(ctx->h_done);
    if (ctx->ev_ctl) cudaEventDestroy(ctx->ev_ctl);
    if (ctx->ev_series) cudaEventDestroy(ctx->ev_series);
    for (int b = 0; b < 2; ++b) if (ctx->ev_launch[b]) cudaEventDestroy(ctx->ev_launch[b]);
    if (ctx->ctl) cudaStreamDestroy(ctx->ctl);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int vptc_set_volume(vptc_ctx* ctx, const void* density_grid, size_t density_bytes, const void* temperature_grid, size_t temperature_bytes)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_set_volume: ctx is null");
    if (!density_grid) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_set_volume: the density grid is required (volume_grids.cpp:48-56)");
    vpt_host::BakedGrid b, t;
    bool has_t = false;
    if (int rc = bake_volume(ctx, density_grid, density_bytes, temperature_grid, temperature_bytes, b, t, has_t)) return rc;
    return stage_volume(ctx, b, has_t ? &t : nullptr);
}

int vptc_wait(vptc_ctx* ctx)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_wait: ctx is null");
    CU(cudaSetDevice(ctx->device));
    return wait_all(ctx);
}

int vptc_render_device_async(vptc_ctx* ctx, const vptc_scene* scene, const vptc_job_range* range, void* d_film)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_device_async: ctx is null");
    return enqueue_async(ctx, scene, range, static_cast<float*>(d_film), nullptr);
}

int vptc_render_device_range(vptc_ctx* ctx, const vptc_scene* scene, const vptc_job_range* range, void* d_film)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_device_range: ctx is null");
    if (int rc = wait_all(ctx)) return rc;
    ctx->last_ms = 0.0f;
    if (int rc = enqueue_async(ctx, scene, range, static_cast<float*>(d_film), nullptr)) return rc;
    if (!ctx->series_open) return VPTC_OK; // empty range
    ctx->series_open = false; ctx->series_launches = 0; ctx->series_has_plain = false;
    return finish_render(ctx);
}

int vptc_render_device(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, uint32_t wave_stride, void* d_film)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_device: ctx is null");
    const vptc_job_range jr{first_wave, n_waves, wave_stride, 1, 0, 0};
    return vptc_render_device_range(ctx, scene, &jr, d_film);
}

int vptc_render_async(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, float* film_xyzw)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_async: ctx is null");
    if (int rc = check_scene(ctx, scene)) return rc;
    if (!film_xyzw) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_async: film is null");
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    CU(cudaSetDevice(ctx->device));
    const size_t n = size_t(scene->output_size[0]) * size_t(scene->output_size[1]) * 4;
    if (ctx->film_floats < n || ctx->pinned_floats < n) { // (re)size the two device films and the pinned staging buffer: nothing may be in flight
        if (int rc = wait_all(ctx)) return rc;
        CU(cudaStreamSynchronize(ctx->stream));
        for (int b = 0; b < 2; ++b) { cudaFree(ctx->film[b]); ctx->film[b] = nullptr; }
        ctx->film_floats = 0;
        if (ctx->pinned) cudaFreeHost(ctx->pinned);
        if (ctx->pinned_zero) cudaFreeHost(ctx->pinned_zero);
        ctx->pinned = ctx->pinned_zero = nullptr; ctx->pinned_floats = 0;
        for (int b = 0; b < 2; ++b) CU(cudaMalloc(&ctx->film[b], n * sizeof(float)));
        ctx->film_floats = n;
        CU(cudaMallocHost(&ctx->pinned, n * sizeof(float)));
        CU(cudaMallocHost(&ctx->pinned_zero, n * sizeof(float)));
        std::memset(ctx->pinned_zero, 0, n * sizeof(float));
        ctx->pinned_floats = n;
    }
    // The device renders this call's waves into a zeroed film of its own; the host later ADDS it into the caller's film, which keeps the
    // reference's "film accumulates" contract without uploading the old film. Buffer b was last used two calls ago and has been collected.
    const int b = ctx->next_buf;
    CU(cudaMemcpyAsync(ctx->film[b], ctx->pinned_zero, n * sizeof(float), cudaMemcpyHostToDevice, ctx->ctl)); // cleared before the range is published (same stream)
    CU(cudaEventRecord(ctx->ev_ctl, ctx->ctl));
    CU(cudaStreamWaitEvent(ctx->stream, ctx->ev_ctl, 0)); // kernels that do not read the ring (variants 0, 1, 3) are ordered after the clear by this
    const vptc_job_range jr{first_wave, n_waves, 1, 1, 0, 0};
    unsigned int rid = ~0u;
    bool launched = false;
    if (int rc = enqueue_async(ctx, scene, &jr, ctx->film[b], &rid, &launched)) return rc;
    if (!launched) return VPTC_OK; // nothing to render
    CU(cudaEventRecord(ctx->ev_launch[b], ctx->stream));
    // now that the GPU has this call's work, bring home the previous call's film (its range completes while this one is already running)
    const int rc = collect_pending(ctx);
    ctx->pending.active = true; ctx->pending.rid = rid; ctx->pending.host_film = film_xyzw; ctx->pending.n = n; ctx->pending.buf = b;
    ctx->next_buf ^= 1;
    return rc;
}

int vptc_render(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, float* film_xyzw)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_render: ctx is null");
    if (int rc = wait_all(ctx)) return rc;
    ctx->last_ms = 0.0f;
    if (int rc = vptc_render_async(ctx, scene, first_wave, n_waves, film_xyzw)) return rc;
    return wait_all(ctx);
}

int vptc_render_adaptive(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, const vptc_adaptive_params* p, float* film_xyzw, float* sum_y2, uint32_t* tile_waves)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_adaptive: ctx is null");
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    if (int rc = check_scene(ctx, scene)) return rc;
    if (!p || !film_xyzw || first_wave < 1 || p->batch_waves < 1 || !(p->rel_error >= 0.0f)) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_adaptive: null argument, first_wave < 1, batch_waves < 1 or rel_error < 0");
    CU(cudaSetDevice(ctx->device));
    if (int rc = wait_all(ctx)) return rc;
    SceneDev sd; vpt_host::prepare_scene(*scene, ctx->has_temperature, sd);
    if (sd.ntiles > 0xffffffffull) return fail(ctx, VPTC_ERR_UNSUPPORTED, "vptc_render_adaptive: more than 2^32 - 1 tiles");
    const size_t npx = size_t(sd.W) * size_t(sd.H), nt = size_t(sd.ntiles);
    float *d_film = nullptr, *d_sq = nullptr, *d_err = nullptr; uint32_t* d_list = nullptr;
    auto cleanup = [&]() { cudaFree(d_film); cudaFree(d_sq); cudaFree(d_err); cudaFree(d_list); };
    cudaError_t e;
    if ((e = cudaMalloc(&d_film, npx * 16)) != cudaSuccess || (e = cudaMalloc(&d_sq, npx * 4)) != cudaSuccess || (e = cudaMalloc(&d_err, nt * 4)) != cudaSuccess ||
        (e = cudaMalloc(&d_list, nt * 4)) != cudaSuccess || (e = cudaMemsetAsync(d_film, 0, npx * 16, ctx->stream)) != cudaSuccess ||
        (e = cudaMemsetAsync(d_sq, 0, npx * 4, ctx->stream)) != cudaSuccess) {
        cleanup(); return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_render_adaptive: ") + cudaGetErrorString(e));
    }
    std::vector<uint32_t> active(nt), waves(nt, 0u);
    for (size_t t = 0; t < nt; ++t) active[t] = uint32_t(t);
    std::vector<float> err(nt);
    uint32_t wave = first_wave, remaining = p->max_waves;
    float total_ms = 0.0f;
    int rc = VPTC_OK;
    bool list_dirty = false; // the first batch covers every tile: a contiguous range, no list
    while (rc == VPTC_OK && remaining > 0 && !active.empty()) {
        const uint32_t b = std::min(p->batch_waves, remaining);
        if (list_dirty) {
            if ((e = cudaMemcpyAsync(d_list, active.data(), active.size() * 4, cudaMemcpyHostToDevice, ctx->stream)) != cudaSuccess) { rc = fail(ctx, VPTC_ERR_CUDA, cudaGetErrorString(e)); break; }
            if ((e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) { rc = fail(ctx, VPTC_ERR_CUDA, cudaGetErrorString(e)); break; }
        }
        RenderArgs A{};
        RangeSpec r;
        r.film = d_film; r.sum_y2 = d_sq; r.stats = true; r.first_wave = wave; r.wave_stride = 1;
        r.tile_begin = 0; r.tile_count = active.size(); r.tile_list = active.size() == nt ? nullptr : d_list;
        r.n_jobs = (unsigned long long)active.size() * b;
        if ((rc = run_render(ctx, scene, A, r, /*trace=*/false)) != VPTC_OK) break;
        total_ms += ctx->last_ms;
        for (uint32_t t : active) waves[t] += b;
        wave += b; remaining -= b;
        if (p->rel_error > 0.0f && remaining > 0 && waves[active[0]] >= std::max(2u, p->min_waves)) { // stop_at_next_wave, per tile (tile_provider.cpp:107-111)
            if ((e = launch_tile_error(d_film, d_sq, sd.W, sd.H, sd.tw, sd.th, sd.ntx, sd.ntiles, d_err, ctx->stream)) != cudaSuccess ||
                (e = cudaMemcpyAsync(err.data(), d_err, nt * 4, cudaMemcpyDeviceToHost, ctx->stream)) != cudaSuccess || (e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) {
                rc = fail(ctx, VPTC_ERR_CUDA, std::string("vptc_render_adaptive: ") + cudaGetErrorString(e)); break;
            }
            std::vector<uint32_t> keep;
            for (uint32_t t : active) if (!(err[t] <= p->rel_error)) keep.push_back(t);
            if (keep.size() != active.size()) { active.swap(keep); list_dirty = true; }
        }
    }
    if (rc == VPTC_OK) {
        std::vector<float> h(npx * 4);
        if ((e = cudaMemcpy(h.data(), d_film, npx * 16, cudaMemcpyDeviceToHost)) != cudaSuccess) rc = fail(ctx, VPTC_ERR_CUDA, cudaGetErrorString(e));
        else host_add(film_xyzw, h.data(), npx * 4);
        if (rc == VPTC_OK && sum_y2) {
            if ((e = cudaMemcpy(h.data(), d_sq, npx * 4, cudaMemcpyDeviceToHost)) != cudaSuccess) rc = fail(ctx, VPTC_ERR_CUDA, cudaGetErrorString(e));
            else host_add(sum_y2, h.data(), npx);
        }
        if (rc == VPTC_OK && tile_waves) std::memcpy(tile_waves, waves.data(), nt * 4);
    }
    ctx->last_ms = total_ms;
    cleanup();
    return rc;
}

float vptc_last_render_ms(const vptc_ctx* ctx) { return ctx ? ctx->last_ms : 0.0f; }
uint64_t vptc_device_volume_bytes(const vptc_ctx* ctx) { return ctx && ctx->has_volume ? uint64_t(ctx->density.bytes) + (ctx->has_temperature ? uint64_t(ctx->temperature.bytes) : 0) : 0; }

int vptc_enable_counters(vptc_ctx* ctx, int enabled)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_enable_counters: ctx is null");
    if (int rc = wait_all(ctx)) return rc;
    ctx->count = enabled != 0;
    return VPTC_OK;
}
int vptc_set_kernel_variant(vptc_ctx* ctx, int variant)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_set_kernel_variant: ctx is null");
    if (variant < 0 || variant > 3) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_set_kernel_variant: 0 (nested), 1 (flat), 2 (flat + stage scheduling) or 3 (per-warp job pool)");
    if (int rc = wait_all(ctx)) return rc;
    ctx->variant = variant;
    return VPTC_OK;
}
int vptc_get_counters(const vptc_ctx* ctx, vptc_counters* out)
{
    if (!ctx || !out) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_get_counters: null argument");
    *out = ctx->last_counters;
    return VPTC_OK;
}

int vptc_trace_jobs(vptc_ctx* ctx, const vptc_scene* scene, const uint64_t* jids, size_t n_jobs, float* per_pixel_xyz, uint32_t* per_pixel_ndraws)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_jobs: ctx is null");
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    if (int rc = check_scene(ctx, scene)) return rc;
    if (!jids || !per_pixel_xyz || !per_pixel_ndraws) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_jobs: null argument");
    if (n_jobs == 0) return VPTC_OK;
    CU(cudaSetDevice(ctx->device));
    const size_t slots = size_t(scene->tile_size[0]) * size_t(scene->tile_size[1]) * n_jobs;
    uint64_t* d_jids = nullptr; float* d_xyz = nullptr; uint32_t* d_nd = nullptr;
    auto cleanup = [&]() { cudaFree(d_jids); cudaFree(d_xyz); cudaFree(d_nd); };
    cudaError_t e;
    if ((e = cudaMalloc(&d_jids, n_jobs * 8)) != cudaSuccess || (e = cudaMalloc(&d_xyz, slots * 12)) != cudaSuccess || (e = cudaMalloc(&d_nd, slots * 4)) != cudaSuccess) {
        cleanup(); return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_trace_jobs: cudaMalloc: ") + cudaGetErrorString(e));
    }
    cudaMemcpyAsync(d_jids, jids, n_jobs * 8, cudaMemcpyHostToDevice, ctx->stream);
    cudaMemsetAsync(d_xyz, 0, slots * 12, ctx->stream);
    cudaMemsetAsync(d_nd, 0, slots * 4, ctx->stream);
    RenderArgs A{};
    RangeSpec rs; rs.n_jobs = n_jobs;
    A.trace_jids = d_jids; A.trace_xyz = d_xyz; A.trace_ndraws = d_nd;
    int rc = wait_all(ctx);
    if (rc == VPTC_OK) rc = run_render(ctx, scene, A, rs, /*trace=*/true);
    if (rc == VPTC_OK) {
        e = cudaMemcpy(per_pixel_xyz, d_xyz, slots * 12, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(per_pixel_ndraws, d_nd, slots * 4, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(ctx, VPTC_ERR_CUDA, std::string("vptc_trace_jobs: copy back: ") + cudaGetErrorString(e));
    }
    cleanup();
    return rc;
}

int vptc_trace_path_log(vptc_ctx* ctx, const vptc_scene* scene_in, uint64_t jid, int64_t px, int64_t py, float* events, size_t cap, size_t* n_events)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_path_log: ctx is null");
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    if (int rc = check_scene(ctx, scene_in)) return rc;
    if (!n_events || (cap && !events)) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_path_log: null argument");
    CU(cudaSetDevice(ctx->device));
    vptc_scene scene = *scene_in; // single_pixel mode: the other pixels of the tile are skipped BEFORE drawing (worker.cpp:113-119)
    scene.worker_parameters.single_pixel_enabled = 1;
    scene.worker_parameters.single_pixel_coord[0] = px; scene.worker_parameters.single_pixel_coord[1] = py;
    const size_t slots = size_t(scene.tile_size[0]) * size_t(scene.tile_size[1]);
    uint64_t* d_jid = nullptr; float *d_xyz = nullptr, *d_ev = nullptr; uint32_t* d_nd = nullptr; unsigned long long* d_n = nullptr;
    auto cleanup = [&]() { cudaFree(d_jid); cudaFree(d_xyz); cudaFree(d_nd); cudaFree(d_ev); cudaFree(d_n); };
    cudaError_t e;
    if ((e = cudaMalloc(&d_jid, 8)) != cudaSuccess || (e = cudaMalloc(&d_xyz, slots * 12)) != cudaSuccess || (e = cudaMalloc(&d_nd, slots * 4)) != cudaSuccess ||
        (e = cudaMalloc(&d_ev, std::max<size_t>(cap, 1) * 32)) != cudaSuccess || (e = cudaMalloc(&d_n, 8)) != cudaSuccess) {
        cleanup(); return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_trace_path_log: cudaMalloc: ") + cudaGetErrorString(e));
    }
    cudaMemcpyAsync(d_jid, &jid, 8, cudaMemcpyHostToDevice, ctx->stream);
    cudaMemsetAsync(d_n, 0, 8, ctx->stream);
    RenderArgs A{};
    RangeSpec rs; rs.n_jobs = 1;
    A.trace_jids = d_jid; A.trace_xyz = d_xyz; A.trace_ndraws = d_nd;
    A.trace_events = d_ev; A.trace_event_count = d_n; A.trace_event_cap = cap;
    const int variant = ctx->variant;
    ctx->variant = 2; // the event hooks live in the default kernel
    int rc = wait_all(ctx);
    if (rc == VPTC_OK) rc = run_render(ctx, &scene, A, rs, /*trace=*/true);
    ctx->variant = variant;
    if (rc == VPTC_OK) {
        unsigned long long n = 0;
        e = cudaMemcpy(&n, d_n, 8, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess && cap) e = cudaMemcpy(events, d_ev, size_t(std::min<unsigned long long>(n, cap)) * 32, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = fail(ctx, VPTC_ERR_CUDA, std::string("vptc_trace_path_log: copy back: ") + cudaGetErrorString(e));
        *n_events = size_t(n);
    }
    cleanup();
    return rc;
}

int vptc_trace_segments(vptc_ctx* ctx, const float* origins, const float* dirs, size_t n, uint32_t max_segs, float* segs, uint32_t* n_segs)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_segments: ctx is null");
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    if (!origins || !dirs || !segs || !n_segs || max_segs == 0) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_segments: bad argument");
    if (n == 0) return VPTC_OK;
    CU(cudaSetDevice(ctx->device));
    float *d_o = nullptr, *d_d = nullptr, *d_s = nullptr; uint32_t* d_n = nullptr;
    auto cleanup = [&]() { cudaFree(d_o); cudaFree(d_d); cudaFree(d_s); cudaFree(d_n); };
    cudaError_t e;
    if ((e = cudaMalloc(&d_o, n * 12)) != cudaSuccess || (e = cudaMalloc(&d_d, n * 12)) != cudaSuccess ||
        (e = cudaMalloc(&d_s, n * size_t(max_segs) * 12)) != cudaSuccess || (e = cudaMalloc(&d_n, n * 4)) != cudaSuccess) {
        cleanup(); return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_trace_segments: cudaMalloc: ") + cudaGetErrorString(e));
    }
    cudaMemcpyAsync(d_o, origins, n * 12, cudaMemcpyHostToDevice, ctx->stream);
    cudaMemcpyAsync(d_d, dirs, n * 12, cudaMemcpyHostToDevice, ctx->stream);
    cudaMemsetAsync(d_s, 0, n * size_t(max_segs) * 12, ctx->stream);
    e = launch_segments(ctx->density.dev, d_o, d_d, n, max_segs, d_s, d_n, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(segs, d_s, n * size_t(max_segs) * 12, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(n_segs, d_n, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cleanup();
    if (e != cudaSuccess) return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_trace_segments: ") + cudaGetErrorString(e));
    return VPTC_OK;
}

int vptc_trace_pixel_csv(vptc_ctx* ctx, const vptc_scene* scene, int64_t px, int64_t py, const char* csv_path)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_pixel_csv: ctx is null");
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    if (!csv_path) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_pixel_csv: csv_path is null");
    if (int rc = check_scene(ctx, scene)) return rc;
    if (px < 0 || py < 0 || px >= scene->output_size[0] || py >= scene->output_size[1]) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_trace_pixel_csv: pixel outside the image");
    SceneDev S;
    vpt_host::prepare_scene(*scene, ctx->has_temperature, S);
    // Camera::generate_ray (camera.hpp:14-23) with sample offset 0: the kernels' arithmetic, on the host (-ffp-contract=off)
    auto sum3 = [](float a, float b, float c) { return a + (b + c); };
    const float rx = float(px) + 0.5f, ry = float(py) + 0.5f;
    float d[3], o[3] = {S.cam_pos[0], S.cam_pos[1], S.cam_pos[2]};
    for (int r = 0; r < 3; ++r) d[r] = sum3(S.cam_m[r * 3] * rx, S.cam_m[r * 3 + 1] * ry, S.cam_m[r * 3 + 2] * 0.0f) + S.cam_t[r];
    const float nrm = std::sqrt(sum3(d[0] * d[0], d[1] * d[1], d[2] * d[2]));
    for (int r = 0; r < 3; ++r) d[r] /= nrm;
    uint32_t n = 0;
    std::vector<float> segs(3);
    if (int rc = vptc_trace_segments(ctx, o, d, 1, 1, segs.data(), &n)) return rc; // count first, then fetch them all
    if (n > 1) { segs.assign(size_t(n) * 3, 0.0f); if (int rc = vptc_trace_segments(ctx, o, d, 1, n, segs.data(), &n)) return rc; }
    // idx_to_world_scale (volume.cpp:91) and world_to_density_index (volume.cpp:172-174), NanoVDB's float map chains (Math.h:905-939)
    const GridDev& G = ctx->density.dev;
    auto mat3 = [](const float* m, const float* v, float* out) {
        for (int r = 0; r < 3; ++r) out[r] = std::fmaf(v[0], m[r * 3], std::fmaf(v[1], m[r * 3 + 1], v[2] * m[r * 3 + 2]));
    };
    float di[3], q[3];
    mat3(G.invMat, d, di);
    const float len = std::sqrt(di[0] * di[0] + di[1] * di[1] + di[2] * di[2]), inv_len = 1.0f / len;
    const float dir_idx[3] = {inv_len * di[0], inv_len * di[1], inv_len * di[2]};
    mat3(G.invMat, dir_idx, q);
    const float scale = 1.0f / std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]);
    auto index_at = [&](float t_world, float* out) { // Ray::eval = origin + direction * t, then worldToIndexF
        const float w[3] = {o[0] + d[0] * t_world - G.vec[0], o[1] + d[1] * t_world - G.vec[1], o[2] + d[2] * t_world - G.vec[2]};
        mat3(G.invMat, w, out);
    };
    FILE* f = std::fopen(csv_path, "w");
    if (!f) return fail(ctx, VPTC_ERR_IO, std::string("vptc_trace_pixel_csv: cannot write ") + csv_path);
    std::fprintf(f, "X0,Y0,Z0,X1,Y1,Z1,T0,T1,Majorant\n");
    for (uint32_t i = 0; i < n; ++i) {
        const float t0 = segs[i * 3] * scale, t1 = segs[i * 3 + 1] * scale;
        float p0[3], p1[3];
        index_at(t0, p0); index_at(t1, p1);
        std::fprintf(f, "%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g\n", p0[0], p0[1], p0[2], p1[0], p1[1], p1[2], t0, t1, segs[i * 3 + 2]);
    }
    std::fclose(f);
    return VPTC_OK;
}

int vptc_sample_grid(vptc_ctx* ctx, int grid, const float* xyz, size_t n, float* values, uint32_t* dims, float* majorants)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_sample_grid: ctx is null");
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    if (grid != 0 && grid != 1) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_sample_grid: grid must be 0 (density) or 1 (temperature)");
    if (grid == 1 && !ctx->has_temperature) return fail(ctx, VPTC_ERR_NO_VOLUME, "vptc_sample_grid: no temperature grid");
    if (!xyz) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_sample_grid: xyz is null");
    if (n == 0) return VPTC_OK;
    CU(cudaSetDevice(ctx->device));
    float *d_p = nullptr, *d_v = nullptr, *d_m = nullptr; uint32_t* d_d = nullptr;
    auto cleanup = [&]() { cudaFree(d_p); cudaFree(d_v); cudaFree(d_m); cudaFree(d_d); };
    cudaError_t e;
    if ((e = cudaMalloc(&d_p, n * 12)) != cudaSuccess || (e = cudaMalloc(&d_v, n * 4)) != cudaSuccess || (e = cudaMalloc(&d_m, n * 4)) != cudaSuccess ||
        (e = cudaMalloc(&d_d, n * 4)) != cudaSuccess) {
        cleanup(); return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_sample_grid: cudaMalloc: ") + cudaGetErrorString(e));
    }
    cudaMemcpyAsync(d_p, xyz, n * 12, cudaMemcpyHostToDevice, ctx->stream);
    const GridDev& G = grid == 1 ? ctx->temperature.dev : ctx->density.dev;
    e = launch_sample(G, d_p, n, d_v, d_d, d_m, ctx->stream);
    if (e == cudaSuccess && values) e = cudaMemcpyAsync(values, d_v, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && dims) e = cudaMemcpyAsync(dims, d_d, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess && majorants) e = cudaMemcpyAsync(majorants, d_m, n * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cleanup();
    if (e != cudaSuccess) return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_sample_grid: ") + cudaGetErrorString(e));
    return VPTC_OK;
}

int vptc_blackbody(vptc_ctx* ctx, const float* temps_k, size_t n, float* xyz)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_blackbody: ctx is null");
    if (n == 0) return VPTC_OK;
    if (!temps_k || !xyz) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_blackbody: null argument");
    CU(cudaSetDevice(ctx->device));
    float *d_t = nullptr, *d_o = nullptr;
    auto cleanup = [&]() { cudaFree(d_t); cudaFree(d_o); };
    cudaError_t e;
    if ((e = cudaMalloc(&d_t, n * 4)) != cudaSuccess || (e = cudaMalloc(&d_o, n * 12)) != cudaSuccess) {
        cleanup(); return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_blackbody: cudaMalloc: ") + cudaGetErrorString(e));
    }
    e = cudaMemcpyAsync(d_t, temps_k, n * 4, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = launch_blackbody(ctx->bb_table, d_t, n, d_o, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(xyz, d_o, n * 12, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cleanup();
    if (e != cudaSuccess) return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_blackbody: ") + cudaGetErrorString(e));
    return VPTC_OK;
}

int vptc_film_to_image(vptc_ctx* ctx, const float* film_xyzw, int64_t width, int64_t height, uint8_t* rgb8)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_film_to_image: ctx is null");
    if (!film_xyzw || !rgb8 || width <= 0 || height <= 0 || width > (int64_t(1) << 20) || height > (int64_t(1) << 20) || width * height > (int64_t(1) << 32))
        return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_film_to_image: bad argument (null pointer, or a side outside 1..2^20 / more than 2^32 pixels)");
    CU(cudaSetDevice(ctx->device));
    const size_t n = size_t(width) * size_t(height);
    float* d_f = nullptr; uint8_t* d_rgb = nullptr;
    auto cleanup = [&]() { cudaFree(d_f); cudaFree(d_rgb); };
    cudaError_t e;
    if ((e = cudaMalloc(&d_f, n * 16)) != cudaSuccess || (e = cudaMalloc(&d_rgb, n * 3)) != cudaSuccess) {
        cleanup(); return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_film_to_image: cudaMalloc: ") + cudaGetErrorString(e));
    }
    e = cudaMemcpyAsync(d_f, film_xyzw, n * 16, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = launch_resolve(d_f, n, d_rgb, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgb8, d_rgb, n * 3, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cleanup();
    if (e != cudaSuccess) return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_film_to_image: ") + cudaGetErrorString(e));
    return VPTC_OK;
}

// ---- vptc_film_to_image_static: the same resolve kernel, staging in a file-scope device array instead of cudaMalloc'ed buffers ----
// For callers that may not allocate (CUDA graph capture around them, serving loops, harnesses that forbid cudaMalloc). No counterpart in the
// reference. The array lives in every GPU that loads this library until the process exits:
// VPTC_STATIC_IMAGE_MAX_PIXELS * (16 + 3) bytes = 38 MiB. Every call on a device shares it (and it keeps the last call's contents: the
// call overwrites the n pixels it uses before reading them, nothing has to start at zero), so calls that overlap on one GPU race.
} // extern "C" (the array and its helper have C++ linkage)
static __device__ __align__(256) unsigned char g_static_image_scratch[size_t(VPTC_STATIC_IMAGE_MAX_PIXELS) * 19];

namespace {
// In host code `g_static_image_scratch` is only a handle: the device address comes from cudaGetSymbolAddress, asked once per device with
// that device current.
constexpr int kMaxStaticDevices = 256;
unsigned char* g_static_image_ptr[kMaxStaticDevices] = {};
std::mutex g_static_image_mutex;
cudaError_t static_image_scratch(int device, unsigned char** out)
{
    if (device < 0 || device >= kMaxStaticDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_image_mutex);
    if (!g_static_image_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_image_scratch); // the current device's instance of the array
        if (e != cudaSuccess) return e;
        g_static_image_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_static_image_ptr[device];
    return cudaSuccess;
}
} // namespace
extern "C" {

int vptc_film_to_image_static(vptc_ctx* ctx, const float* film_xyzw, int64_t width, int64_t height, uint8_t* rgb8)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_film_to_image_static: ctx is null");
    if (!film_xyzw || !rgb8 || width <= 0 || height <= 0)
        return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_film_to_image_static: null pointer or a side < 1");
    if (width > int64_t(VPTC_STATIC_IMAGE_MAX_PIXELS) || height > int64_t(VPTC_STATIC_IMAGE_MAX_PIXELS) || width * height > int64_t(VPTC_STATIC_IMAGE_MAX_PIXELS))
        return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_film_to_image_static: " + std::to_string(width) + " x " + std::to_string(height) + " pixels exceed the static scratch of VPTC_STATIC_IMAGE_MAX_PIXELS = " +
                                                        std::to_string(VPTC_STATIC_IMAGE_MAX_PIXELS) + " pixels; use vptc_film_to_image");
    CU(cudaSetDevice(ctx->device));
    unsigned char* scratch = nullptr;
    cudaError_t e = static_image_scratch(ctx->device, &scratch);
    if (e != cudaSuccess) return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_film_to_image_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
    const size_t n = size_t(width) * size_t(height);
    float* d_f = reinterpret_cast<float*>(scratch);                                     // n float4 pixels, overwritten below
    uint8_t* d_rgb = scratch + size_t(VPTC_STATIC_IMAGE_MAX_PIXELS) * 16;               // n RGB8 pixels, every byte written by the kernel
    e = cudaMemcpyAsync(d_f, film_xyzw, n * 16, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = launch_resolve(d_f, n, d_rgb, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(rgb8, d_rgb, n * 3, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e != cudaSuccess) return fail(ctx, VPTC_ERR_CUDA, std::string("vptc_film_to_image_static: ") + cudaGetErrorString(e));
    return VPTC_OK;
}

#include "vptc_group_impl.hpp" // vptc_group_*: several devices behind one handle (same translation unit: it drives the contexts' internals)

} // extern "C"
