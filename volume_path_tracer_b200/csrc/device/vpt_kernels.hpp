// vpt_kernels.hpp — host-visible interface of vpt_kernels.cu.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "layout.hpp"

#ifndef VPT_BLOCK_THREADS
#define VPT_BLOCK_THREADS 128
#endif

namespace vptd {

struct RenderArgs {
    GridDev density, temperature;
    SceneDev scene;
    const float* bb_table;        // 1500 floats: precompute_blackbody.cpp's bb_xyz[500]
    float* film;                  // H*W*4, accumulated into (render mode)
    unsigned long long* queue;    // device job counter, zeroed before launch
    unsigned long long n_jobs;    // ntiles * n_waves (render) or number of listed jobs (trace)
    uint32_t first_wave, wave_stride;
    const uint64_t* trace_jids;   // trace mode
    float* trace_xyz;
    uint32_t* trace_ndraws;
    CountersDev* counters;        // COUNT builds
    int sched_thresh[4];          // variant 2: lanes needed to run SETUP / HOT / - / EVAL in an iteration
    int pool_params[4];           // variant 3: {HOT passes per scheduling decision, queued slots needed to run HOT, EVAL, SETUP}
    // variant 2 (default kernel): the job space comes from the range ring (layout.hpp RangeDesc) instead of queue / n_jobs / film
    RangeDesc* ring;              // RING_SLOTS descriptors
    const unsigned int* ring_tail;// absolute number of ranges published so far (written by the host while the kernel runs)
    unsigned int range_begin;     // the range this launch was made for; its lanes start there
    unsigned int* host_done;      // mapped pinned, RING_SLOTS words: rid + 1 once range rid is complete (may be null)
    unsigned int* exited;         // lanes of THIS launch that have left (zeroed before the launch); with exit_limit: a launch whose lanes have
    unsigned int exit_limit;      // mostly gone stops carrying over into later ranges, so that the full grid queued behind it can take over
    float* trace_events;          // trace mode, optional: Logger<true> event log (worker.cpp:15-49), 8 floats per event
    unsigned long long* trace_event_count; unsigned long long trace_event_cap;
};

cudaError_t upload_cie_table(const float* cie471x3, const float* lambda5_471); // CIE observer + the host libm's lambda^5 per wavelength
// variant: 0 = nested megakernel (first version), 1 = flat state machine, 2 = flat + warp-level stage scheduling,
//          3 = per-warp job pool in shared memory (more jobs than lanes, stage queues compacted onto lanes)
cudaError_t launch_render(const RenderArgs& args, int variant, bool trace, bool count, int grid_blocks, cudaStream_t stream);
int render_blocks_per_sm(int variant, bool trace, bool count);
cudaError_t launch_render_stats(const RenderArgs& args, int grid_blocks, cudaStream_t stream); // default kernel + tile lists + second-moment plane (adaptive rendering)
int render_stats_blocks_per_sm();
cudaError_t launch_render_ordered(const RenderArgs& args, int ord, int grid_blocks, cudaStream_t stream); // default kernel, tile-order instantiations: 1 learn, 2 ordered
int render_ordered_blocks_per_sm(int ord);
cudaError_t launch_tile_error(const float* film, const float* sum_y2, long long W, long long H, long long tw, long long th, long long ntx, unsigned long long ntiles, float* err, cudaStream_t s);
int render_jobs_per_block(int variant); // jobs a resident block works on at a time (grid sizing)
cudaError_t launch_segments(const GridDev& G, const float* origins, const float* dirs, size_t n, uint32_t max_segs, float* segs, uint32_t* n_segs, cudaStream_t s);
cudaError_t launch_sample(const GridDev& G, const float* xyz, size_t n, float* values, uint32_t* dims, float* majorants, cudaStream_t s);
constexpr int FILM_REDUCE_MAX_SRC = 8;
struct FilmSources { const float4* p[FILM_REDUCE_MAX_SRC]; };
// dst (n_pixels float4 on the current device) += sum of k <= FILM_REDUCE_MAX_SRC source films, which may be peer-mapped memory of other devices
cudaError_t launch_film_reduce(float* dst, const float* const* srcs, int k, size_t n_pixels, cudaStream_t s);
cudaError_t launch_blackbody(const float* bb_table, const float* temps, size_t n, float* xyz, cudaStream_t s);
// device-side brick stage of the baker (device/vpt_bake.cu): apron bricks + patched leaf maxima from the serialized grid in device memory
cudaError_t launch_bake_bricks(const void* d_grid, uint64_t n_leaf, const uint32_t* d_brick_of, bool patch, float* d_bricks, float* d_max_by_brick,
                               void* d_slots, uint64_t n_slots, cudaStream_t s);
cudaError_t launch_resolve(const float* film, size_t n_pixels, uint8_t* rgb, cudaStream_t s);

} // namespace vptd
