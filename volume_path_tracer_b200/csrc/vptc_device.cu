// vptc_device.cu — the device-facing entry points of include/vptc.h: context, volume staging, render,
// parity hooks. Plain CUDA runtime; no torch types cross this boundary.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <memory>
#include <string>
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
    void *cells = nullptr, *slots = nullptr, *tile_value = nullptr, *tile_majorant = nullptr, *bricks = nullptr;
    size_t bytes = 0;
    void release()
    {
        cudaFree(cells); cudaFree(slots); cudaFree(tile_value); cudaFree(tile_majorant); cudaFree(bricks);
        cells = slots = tile_value = tile_majorant = bricks = nullptr; bytes = 0;
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
    // scratch film for the host-buffer path
    float* film = nullptr; size_t film_floats = 0;
    float* pinned = nullptr; size_t pinned_floats = 0;
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
    cudaError_t e;
    if ((e = up(&g.cells, b.cells.data(), b.cells.size() * 4)) != cudaSuccess || (e = up(&g.slots, b.slots.data(), b.slots.size() * 4)) != cudaSuccess ||
        (e = up(&g.tile_value, b.tile_value.data(), b.tile_value.size() * 4)) != cudaSuccess ||
        (e = up(&g.tile_majorant, b.tile_majorant.data(), b.tile_majorant.size() * 4)) != cudaSuccess ||
        (e = up(&g.bricks, b.bricks.data(), b.bricks.size() * 4)) != cudaSuccess || (e = cudaStreamSynchronize(ctx->stream)) != cudaSuccess) {
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

// Launch the megakernel over n_jobs jobs and wait; fills last_ms / counters.
int run_render(vptc_ctx* ctx, const vptc_scene* scene, RenderArgs& A, bool trace)
{
    vpt_host::prepare_scene(*scene, ctx->has_temperature, A.scene);
    A.density = ctx->density.dev;
    A.temperature = ctx->has_temperature ? ctx->temperature.dev : ctx->density.dev;
    A.bb_table = ctx->bb_table;
    A.queue = ctx->queue;
    A.counters = ctx->counters;
    for (int i = 0; i < 4; ++i) A.sched_thresh[i] = std::max(1, ctx->sched_thresh[i]); // the kernel relies on >= 1
    CU(cudaMemsetAsync(ctx->queue, 0, sizeof(unsigned long long), ctx->stream));
    if (ctx->count) CU(cudaMemsetAsync(ctx->counters, 0, sizeof(CountersDev), ctx->stream));
    for (int i = 0; i < 4; ++i) A.pool_params[i] = ctx->pool_params[i];
    for (int i = 0; i < 4; ++i) A.pool_params[i] = std::max(1, A.pool_params[i]);
    // variant 3 packs pixel coordinates into 16-bit halves of its slot words: larger frames / tiles take variant 2
    const int variant = (ctx->variant == 3 && (scene->output_size[0] > 65535 || scene->output_size[1] > 65535 || scene->tile_size[0] > 65535 || scene->tile_size[1] > 65535)) ? 2 : ctx->variant;
    int per_sm = render_blocks_per_sm(variant, trace, ctx->count);
    if (const char* v = std::getenv("VPT_MAX_BLOCKS_PER_SM")) per_sm = std::min(per_sm, std::max(1, std::atoi(v))); // occupancy experiments
    if (per_sm <= 0) return fail(ctx, VPTC_ERR_CUDA, "render kernel cannot be resident on this device (occupancy query returned 0)");
    // persistent grid: every SM filled to its occupancy limit, never more blocks than jobs need
    const unsigned long long per_block = (unsigned long long)(render_jobs_per_block(variant));
    unsigned long long want = (A.n_jobs + per_block - 1) / per_block;
    int blocks = int(std::min<unsigned long long>(want ? want : 1ULL, (unsigned long long)(ctx->num_sms) * per_sm));
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    CU(launch_render(A, variant, trace, ctx->count, blocks, ctx->stream));
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    CU(cudaEventSynchronize(ctx->ev1));
    CU(cudaEventElapsedTime(&ctx->last_ms, ctx->ev0, ctx->ev1));
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
    CU(cudaStreamCreate(&ctx->stream)); // blocking stream: ordered after work on the legacy default stream (torch)
    CU(cudaEventCreate(&ctx->ev0));
    CU(cudaEventCreate(&ctx->ev1));
    CU(cudaMalloc(&ctx->bb_table, 1504 * sizeof(float))); // +1 zero entry: the reference reads bb_xyz[500] for 49 800 K < T < 49 900 K
    CU(cudaMemset(ctx->bb_table, 0, 1504 * sizeof(float)));
    CU(cudaMalloc(&ctx->queue, sizeof(unsigned long long)));
    CU(cudaMalloc(&ctx->counters, sizeof(CountersDev)));
    std::vector<float> bb(1500);
    vpt_host::compute_blackbody_table(bb.data()); // init_blackbody_radiation_xyz (main.cpp:32)
    CU(cudaMemcpy(ctx->bb_table, bb.data(), 1500 * sizeof(float), cudaMemcpyHostToDevice));
    float lambda5[471];
    vpt_host::lambda5_table(lambda5);
    CU(upload_cie_table(vpt_host::cie_table(), lambda5));
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
    cudaFree(ctx->bb_table); cudaFree(ctx->queue); cudaFree(ctx->counters); cudaFree(ctx->film);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

int vptc_set_volume(vptc_ctx* ctx, const void* density_grid, size_t density_bytes, const void* temperature_grid, size_t temperature_bytes)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_set_volume: ctx is null");
    if (!density_grid) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_set_volume: the density grid is required (volume_grids.cpp:48-56)");
    CU(cudaSetDevice(ctx->device));
    // Stage into fresh device grids and swap only when EVERYTHING succeeded: a failed call (corrupt buffer, out of device memory on a
    // larger second volume) leaves the context rendering the previous volume, never a half-released one.
    DeviceGrid new_density, new_temperature;
    bool new_has_temperature = false;
    try {
        {
            vpt_host::BakedGrid b;
            vpt_host::bake_grid(density_grid, density_bytes, /*patch_majorants=*/true, b);
            if (int rc = upload(ctx, b, new_density)) return rc;
        }
        if (temperature_grid && temperature_bytes) {
            vpt_host::BakedGrid t;
            vpt_host::bake_grid(temperature_grid, temperature_bytes, /*patch_majorants=*/false, t);
            if (int rc = upload(ctx, t, new_temperature)) { new_density.release(); return rc; }
            new_has_temperature = true;
        }
    } catch (const std::exception& e) {
        new_density.release(); new_temperature.release();
        return fail(ctx, VPTC_ERR_BAD_GRID, std::string("vptc_set_volume: ") + e.what());
    }
    ctx->density.release(); ctx->temperature.release();
    ctx->density = new_density; ctx->temperature = new_temperature;
    ctx->has_temperature = new_has_temperature;
    ctx->has_volume = true;
    return VPTC_OK;
}

int vptc_render_device(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, uint32_t wave_stride, void* d_film)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_device: ctx is null");
    if (!ctx->has_volume) return fail(ctx, VPTC_ERR_NO_VOLUME, "no volume set: call vptc_set_volume first");
    if (int rc = check_scene(ctx, scene)) return rc;
    if (!d_film || first_wave < 1 || wave_stride < 1) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_render_device: film null, first_wave < 1 or wave_stride < 1");
    CU(cudaSetDevice(ctx->device));
    ctx->last_ms = 0.0f;
    if (n_waves == 0) return VPTC_OK;
    RenderArgs A{};
    A.film = static_cast<float*>(d_film);
    A.first_wave = first_wave; A.wave_stride = wave_stride;
    vptd::SceneDev tmp; vpt_host::prepare_scene(*scene, ctx->has_temperature, tmp);
    A.n_jobs = tmp.ntiles * (unsigned long long)n_waves;
    return run_render(ctx, scene, A, /*trace=*/false);
}

int vptc_render(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, float* film_xyzw)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_render: ctx is null");
    if (int rc = check_scene(ctx, scene)) return rc;
    if (!film_xyzw) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_render: film is null");
    CU(cudaSetDevice(ctx->device));
    const size_t n = size_t(scene->output_size[0]) * size_t(scene->output_size[1]) * 4;
    if (ctx->film_floats < n) {
        cudaFree(ctx->film); ctx->film = nullptr; ctx->film_floats = 0;
        CU(cudaMalloc(&ctx->film, n * sizeof(float)));
        ctx->film_floats = n;
    }
    if (ctx->pinned_floats < n) {
        if (ctx->pinned) cudaFreeHost(ctx->pinned);
        ctx->pinned = nullptr; ctx->pinned_floats = 0;
        CU(cudaMallocHost(&ctx->pinned, n * sizeof(float)));
        ctx->pinned_floats = n;
    }
    // The device renders this call's waves into a zeroed film; the host then ADDS it into the caller's
    // film, which keeps the reference's "film accumulates" contract without uploading the old film.
    CU(cudaMemsetAsync(ctx->film, 0, n * sizeof(float), ctx->stream));
    const int rc = vptc_render_device(ctx, scene, first_wave, n_waves, 1, ctx->film);
    if (rc != VPTC_OK) return rc;
    CU(cudaMemcpyAsync(ctx->pinned, ctx->film, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    for (size_t i = 0; i < n; ++i) film_xyzw[i] += ctx->pinned[i];
    return VPTC_OK;
}

float vptc_last_render_ms(const vptc_ctx* ctx) { return ctx ? ctx->last_ms : 0.0f; }
uint64_t vptc_device_volume_bytes(const vptc_ctx* ctx) { return ctx && ctx->has_volume ? uint64_t(ctx->density.bytes) + (ctx->has_temperature ? uint64_t(ctx->temperature.bytes) : 0) : 0; }

int vptc_enable_counters(vptc_ctx* ctx, int enabled)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_enable_counters: ctx is null");
    ctx->count = enabled != 0;
    return VPTC_OK;
}
int vptc_set_kernel_variant(vptc_ctx* ctx, int variant)
{
    if (!ctx) return fail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_set_kernel_variant: ctx is null");
    if (variant < 0 || variant > 3) return fail(ctx, VPTC_ERR_INVALID_ARGUMENT, "vptc_set_kernel_variant: 0 (nested), 1 (flat), 2 (flat + stage scheduling) or 3 (per-warp job pool)");
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
    A.n_jobs = n_jobs; A.first_wave = 1; A.wave_stride = 1;
    A.trace_jids = d_jids; A.trace_xyz = d_xyz; A.trace_ndraws = d_nd;
    int rc = run_render(ctx, scene, A, /*trace=*/true);
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
    A.n_jobs = 1; A.first_wave = 1; A.wave_stride = 1;
    A.trace_jids = d_jid; A.trace_xyz = d_xyz; A.trace_ndraws = d_nd;
    A.trace_events = d_ev; A.trace_event_count = d_n; A.trace_event_cap = cap;
    const int variant = ctx->variant;
    ctx->variant = 2; // the event hooks live in the default kernel
    int rc = run_render(ctx, &scene, A, /*trace=*/true);
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

} // extern "C"
