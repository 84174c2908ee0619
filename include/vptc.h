/* vptc.h — C ABI of libvptcuda: the B200 (sm_100a) drop-in for volume_path_tracer's
 * per-pixel volumetric path integrator.
 *
 * The reference has no FFI layer; its seam is the free function
 *     void vpt::run(const WorkerParameters&, const Volume&, const Camera&, TileProvider&,
 *                   Image<float,4>& film, RandomNumberGenerator rng)     include/vpt/worker.hpp:11
 * called from num_workers jthreads (src/main.cpp:63-87), bracketed by the construction of
 * VolumeGrids/Volume/Camera/TileProvider/film (src/main.cpp:40-56) and by film_to_image + save
 * (src/main.cpp:142-143).  Each entry point below names the reference interface it replaces.
 *
 * Conventions: plain pointers and sizes, no C++/torch types; every function returns 0 on success or
 * a negative vptc_status, never throws or aborts across the boundary; the message for the last
 * failure is available from vptc_last_error().  The reference's own convention for the same errors
 * is `vptFATAL(...)` = message to std::clog + exit(1) (include/vpt/logging.hpp:16); the host `vpt`
 * CLI maps a non-zero return to exactly that.
 * A context is driven by one host thread at a time (calls are blocking unless stated).
 */
#ifndef VPTC_H
#define VPTC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VPTC_API __attribute__((visibility("default")))

typedef enum vptc_status {
    VPTC_OK = 0,
    VPTC_ERR_INVALID_ARGUMENT = -1,
    VPTC_ERR_BAD_GRID = -2,      /* buffer is not a float NanoVDB grid of ABI major 32 */
    VPTC_ERR_CUDA = -3,
    VPTC_ERR_NO_VOLUME = -4,
    VPTC_ERR_IO = -5,
    VPTC_ERR_CONFIG = -6,        /* JSON scene: missing key / wrong type (configuration.cpp:13-19) */
    VPTC_ERR_UNSUPPORTED = -7
} vptc_status;

/* ---- Scene POD: vpt::Configuration field for field (include/vpt/configuration.hpp:14-65) ---- */
typedef struct vptc_camera_parameters { /* configuration.hpp:14-21 */
    float position[3], look[3], up[3];
    float vfov_deg, imaging_ratio;
} vptc_camera_parameters;

typedef struct vptc_infinite_light { float xyz[3]; float multiplier; } vptc_infinite_light; /* :23-26 */
typedef struct vptc_distant_light {                                                         /* :28-32 */
    float xyz[3]; float multiplier; float inv_direction[3];
} vptc_distant_light;

typedef struct vptc_worker_parameters { /* configuration.hpp:34-44 */
    int32_t single_pixel_enabled;
    int64_t single_pixel_coord[2];      /* [x, y] */
    int32_t use_jitter;
    vptc_infinite_light infinite_light;
    vptc_distant_light  distant_light;
    uint32_t max_depth;
} vptc_worker_parameters;

typedef struct vptc_volume_parameters { /* configuration.hpp:46-53 */
    float henyey_greenstein_g, le_scale, sigma_a, sigma_s, temperature_offset, temperature_scale;
} vptc_volume_parameters;

typedef struct vptc_scene { /* configuration.hpp:55-65 (volume_path is returned separately) */
    uint32_t seed;
    int64_t  output_size[2];   /* [W, H] */
    int64_t  tile_size[2];     /* [w, h] */
    uint32_t num_waves;
    uint32_t num_workers;      /* mandatory key; used only by CPU baselines */
    vptc_camera_parameters camera_parameters;
    vptc_worker_parameters worker_parameters;
    vptc_volume_parameters volume_parameters;
} vptc_scene;

/* Work counters of the last render call (the algorithmic-bytes formula
 * 32*N_rho + 32*N_T + 4*N_maj + 32 per pixel-sample is evaluated from these). */
typedef struct vptc_counters {
    uint64_t pixel_samples;    /* camera paths finished                                   */
    uint64_t density_evals;    /* N_rho: trilinear density evaluations (sampler t < t1)   */
    uint64_t temperature_evals;/* N_T                                                      */
    uint64_t majorant_fetches; /* N_maj: update_current_majorant calls (volume.cpp:49,68) */
    uint64_t rng_draws;
    uint64_t scatter_events;
    uint64_t shadow_rays;
} vptc_counters;

typedef struct vptc_ctx vptc_ctx;

/* ---- context ---- */
/* Owns all device memory and the render stream of ONE CUDA device (one process per GPU). */
VPTC_API int  vptc_create(vptc_ctx** out, int device);
VPTC_API void vptc_destroy(vptc_ctx* ctx);
/* ctx may be NULL: returns the calling thread's last context-less error. */
VPTC_API const char* vptc_last_error(const vptc_ctx* ctx);

/* Replaces VolumeGrids(handle) + Volume::Volume (src/volume_grids.cpp:21-33, src/volume.cpp:162-170):
 * takes the raw serialized NanoVDB grid buffers that nanovdb::io::readGrid returns (density
 * required, temperature optional: pass NULL/0), applies fix_majorants_for_interpolation
 * (volume.cpp:104-160) to a private copy — the caller's memory is never mutated and may be freed on
 * return — bakes the brick-sparse layout and stages it to HBM once.  The tree is validated and its small tables are built on the host; the
 * apron bricks and the halo patch are computed on the device from the uploaded grid (csrc/device/vpt_bake.cu; VPT_HOST_BAKE=1 in the
 * environment keeps everything on the host threads, byte-identical result). */
VPTC_API int vptc_set_volume(vptc_ctx* ctx, const void* density_grid, size_t density_bytes,
                             const void* temperature_grid, size_t temperature_bytes);

/* Replaces the thread-spawn block + vpt::run (src/main.cpp:63-87, src/worker.cpp:92-208) for waves
 * first_wave .. first_wave + n_waves - 1 (1-based, tile_provider.cpp:30); job ids are
 * jid = (wave-1)*ntiles + tile exactly as TileProvider::next() hands them out.  ADDS
 * (imaging_ratio*L, 1) per pixel-sample into the caller-owned host film, H x W row-major float[4]
 * (worker.cpp:203-204, image.hpp:59).  Host<->device copies are inside the call.
 * WHICH lane renders a job WHEN is the library's business: after the first wave(s) of a scene it renders full-frame calls (and tile stripes)
 * longest-tiles-first, tile-major, from per-tile job times it has measured.  The job ids, hence every sample, are the
 * reference's; films differ from call to call by fp32 add order only.  VPT_TILE_ORDER=0 in the environment keeps TileProvider's order. */
VPTC_API int vptc_render(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves,
                         float* film_xyzw);

/* Same work, film resident in device memory (d_film_xyzw is a device pointer on ctx's device).
 * Waves rendered: first_wave + k*wave_stride, k = 0..n_waves-1 (wave_stride >= 1; used to shard a
 * progressive render across GPUs).  Enqueued on ctx's stream; returns after the kernels finished. */
VPTC_API int vptc_render_device(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves,
                                uint32_t wave_stride, void* d_film_xyzw);

/* A slice of the (tile, wave) job space: waves first_wave + k*wave_stride, k < n_waves, of the tiles [tile_begin, tile_begin + tile_count)
 * (row-major tile indices as TileProvider::compute_tile_rect numbers them, src/tile_provider.cpp:95-105; tile_count 0 = every tile).
 * tile_row_stride > 1 makes the slice an INTERLEAVED stripe: tile_begin and tile_count must be multiples of the tiles per row, and the
 * slice's rows of tiles are tile_begin/ntx, + tile_row_stride, + 2*tile_row_stride, ... (G devices taking every G-th row share the
 * expensive part of a frame evenly; 0 and 1 both mean a contiguous slice).
 * Job ids stay (wave-1)*ntiles + tile whatever the slice, so any partition of the job space — by waves, by tile stripes, across calls or
 * across GPUs — renders exactly the samples one vptc_render call over the whole space would. */
typedef struct vptc_job_range {
    uint32_t first_wave, n_waves, wave_stride;
    uint32_t tile_row_stride;
    uint64_t tile_begin, tile_count;
} vptc_job_range;

/* vptc_render_device over an arbitrary job slice (blocking). */
VPTC_API int vptc_render_device_range(vptc_ctx* ctx, const vptc_scene* scene, const vptc_job_range* range, void* d_film_xyzw);

/* Asynchronous forms: enqueue and return. The reference's workers never stop between waves (TileProvider hands out the next wave's jobs
 * as soon as a worker is free, src/tile_provider.cpp:27-67); a sequence of blocking calls would instead let the GPU drain at the end of
 * every call (a job is a whole tile rendered by one lane). Calls made before vptc_wait are published to a device-side ring: the lanes of
 * the kernel already running carry on into the next call's jobs, so many small calls run at the rate of one large one.
 *   vptc_render_device_async: film resident on the device (may be the same film for every call).
 *   vptc_render_async: host film; returns once the PREVIOUS async call's film has been added into its host film (that copy and add overlap
 *     this call's rendering). The film passed to a call is complete after the next vptc_render_async or vptc_wait returns; do not read it before.
 *   vptc_wait: everything enqueued is finished and accumulated; vptc_last_render_ms then covers the whole series.
 * Any other call on the context waits first. A call whose scene differs from the running series' waits for the series, then starts its own. */
VPTC_API int vptc_render_device_async(vptc_ctx* ctx, const vptc_scene* scene, const vptc_job_range* range, void* d_film_xyzw);
VPTC_API int vptc_render_async(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, float* film_xyzw);
VPTC_API int vptc_wait(vptc_ctx* ctx);

/* Variance-driven adaptive wave count. The reference renders every tile for num_waves waves and can only be stopped as a
 * whole (TileProvider::stop_at_next_wave, src/tile_provider.cpp:107-111); here each TILE stops on its own once the relative standard error of
 * its mean luminance (Y of the XYZ film) falls to rel_error: waves are rendered in batches of batch_waves, and after a batch a tile leaves the
 * job space when  sqrt(mean over its pixels of s^2/n) / (mean over its pixels of y-bar) <= rel_error  (s^2 = unbiased sample variance of a
 * pixel's n luminance samples, kept in a second-moment plane beside the film), never before min_waves and never after max_waves.
 * Job ids are unchanged, so a tile's film is exactly the plain render truncated at the wave the tile stopped: film.w holds each pixel's sample
 * count, film.xyz the sums (ADDED into film_xyzw like vptc_render). rel_error = 0 never stops: the plain render plus its second moments.
 * Outside the bit-exact contract by construction (the reference has no such mode): validated against the CPU oracle by Monte-Carlo statistics.
 * sum_y2 (H*W floats, ADDED into; may be NULL) receives sum of (imaging_ratio*L.y)^2 per pixel; tile_waves (ntiles, may be NULL) the waves per tile. */
typedef struct vptc_adaptive_params { uint32_t max_waves, batch_waves, min_waves; float rel_error; } vptc_adaptive_params;
VPTC_API int vptc_render_adaptive(vptc_ctx* ctx, const vptc_scene* scene, uint32_t first_wave, const vptc_adaptive_params* params,
                                  float* film_xyzw, float* sum_y2, uint32_t* tile_waves);

/* ---- several GPUs behind one handle ----
 * Replaces the reference's worker pool as a whole: src/main.cpp:63-87 spawns num_workers threads that share ONE TileProvider, ONE Volume and
 * ONE film. A group is that, with GPUs for workers: one host thread drives every listed device, the baked volume is replicated in each
 * device's HBM (baked once on the host), the (tile, wave) job space is partitioned statically, and the films are combined on the first
 * device over peer memory (NVLink / NVSwitch):
 *   VPTC_SHARD_WAVES  device g of G renders waves first + g, first + g + G, ... of the whole frame; the partial films are SUMMED by a kernel on
 *                     the first device that reads the other films in place through peer mappings (progressive rendering, high spp);
 *   VPTC_SHARD_TILES  device g renders all waves of every G-th ROW of tiles, g, g + G, ... (src/tile_provider.cpp:95-105; interleaved so that the
 *                     devices share the expensive part of the frame); the rows are GATHERED with peer copies: disjoint pixels, 1/G of the film
 *                     each, no arithmetic (low spp, large frames).
 * Job ids are those of the single-device render either way, so sample counts are equal and radiance differs only by fp32 add order.
 * vptc_group_render_async enqueues on every device and returns; vptc_group_wait finishes everything enqueued since the last wait, combines
 * and ADDS the result into the caller's host film. vptc_group_render = both. Frame size and sharding must not change between waits. */
typedef struct vptc_group vptc_group;
enum { VPTC_SHARD_WAVES = 0, VPTC_SHARD_TILES = 1 };
VPTC_API int  vptc_group_create(vptc_group** out, const int* devices, int n_devices);
VPTC_API void vptc_group_destroy(vptc_group* group);
VPTC_API const char* vptc_group_last_error(const vptc_group* group);
VPTC_API int  vptc_group_size(const vptc_group* group);
VPTC_API int  vptc_group_peer_mapped(const vptc_group* group); /* devices the first one reads directly; the others are staged by a peer copy */
VPTC_API int  vptc_group_set_volume(vptc_group* group, const void* density_grid, size_t density_bytes, const void* temperature_grid, size_t temperature_bytes);
VPTC_API int  vptc_group_render(vptc_group* group, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, int sharding, float* film_xyzw);
VPTC_API int  vptc_group_render_async(vptc_group* group, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, int sharding);
VPTC_API int  vptc_group_wait(vptc_group* group, float* film_xyzw);
VPTC_API float vptc_group_last_render_ms(const vptc_group* group); /* slowest device of the last vptc_group_wait */
VPTC_API vptc_ctx* vptc_group_context(vptc_group* group, int index); /* one device's context (hooks, counters); owned by the group */

/* Device time (CUDA events on ctx's stream) of the kernels of the last render call, milliseconds. */
VPTC_API float vptc_last_render_ms(const vptc_ctx* ctx);
/* Bytes of HBM the baked volume (density + temperature: cells, lower slots, tiles, apron bricks) occupies; 0 before vptc_set_volume. */
VPTC_API uint64_t vptc_device_volume_bytes(const vptc_ctx* ctx);
VPTC_API int   vptc_get_counters(const vptc_ctx* ctx, vptc_counters* out);
/* Counting costs atomics; off by default. */
VPTC_API int   vptc_enable_counters(vptc_ctx* ctx, int enabled);

/* Which render kernel runs: 2 (default) = flat state machine with warp-level stage scheduling, 1 = flat state
 * machine, 0 = the first nested megakernel (kept so the design choice stays measurable: profiles/). All produce
 * the same film bit for bit per sample.
 * The environment variable VPT_KERNEL_VARIANT sets the default at vptc_create. */
VPTC_API int   vptc_set_kernel_variant(vptc_ctx* ctx, int variant);

/* Parity hook = the reference's single_pixel / Logger<true> debugging aids (worker.cpp:15-49,113-119):
 * runs the given jobs and returns, per job and per pixel of its tile (row-major inside the tile,
 * tile_w*tile_h slots per job, unused slots of clipped tiles zero), the un-scaled radiance L (XYZ)
 * and the number of RNG draws the job had consumed when the pixel finished. */
VPTC_API int vptc_trace_jobs(vptc_ctx* ctx, const vptc_scene* scene, const uint64_t* jids, size_t n_jobs,
                             float* per_pixel_xyz, uint32_t* per_pixel_ndraws);

/* Debug front-end = Logger<true> + single_pixel (worker.cpp:15-49,113-119; the reference ships Logger<false>, :102): the event log
 * of the ONE camera path that job `jid` traces through pixel (px, py) when every other pixel of its tile is skipped before drawing.
 * 8 floats per event {type, 6 arguments, 0}: 0 new_ray(origin, direction)  1 sampled_point(point, density)  2 null
 * 3 scatter(origin, new direction)  4 absorbed  (5 scatter_terminated: unreachable).
 * At most cap_events are stored; *n_events receives the true count. Runs the default kernel (variant 2). */
VPTC_API int vptc_trace_path_log(vptc_ctx* ctx, const vptc_scene* scene, uint64_t jid, int64_t px, int64_t py,
                                 float* events, size_t cap_events, size_t* n_events);

/* Parity hook = Volume::log_majorant_trace (volume.cpp:176-192): merged majorant segments
 * {t0, t1, d_maj} (index-space t) of n world-space rays; at most max_segs per ray are stored,
 * n_segs[i] receives the true count. */
VPTC_API int vptc_trace_segments(vptc_ctx* ctx, const float* origins_xyz, const float* dirs_xyz, size_t n_rays,
                                 uint32_t max_segs, float* segs_t0_t1_dmaj, uint32_t* n_segs);

/* Debug front-end = Volume::log_majorant_trace (volume.cpp:176-192) as the reference's ray_visualizer drives it, headless:
 * traces the UN-JITTERED camera ray through the centre of pixel (px, py) on the device and writes `csv_path` with the
 * reference's columns X0,Y0,Z0,X1,Y1,Z1,T0,T1,Majorant — index-space end points of each merged majorant segment
 * (world_to_density_index(ray.eval(t * idx_to_world_scale))), world-space T0/T1, and d_maj — so scripts/plot_majorant_trace.py
 * keeps working. A ray that misses the volume writes the header only. (log_dda_trace's NodeInfo columns need the NanoVDB
 * tree itself and stay with the oracle.) */
VPTC_API int vptc_trace_pixel_csv(vptc_ctx* ctx, const vptc_scene* scene, int64_t px, int64_t py, const char* csv_path);

/* Parity hook for the trilinear sampler (SampleFromVoxels<Acc,1>): grid 0 = density, 1 = temperature;
 * index-space positions. Also returns getDim/majorant per position when the pointers are non-NULL. */
VPTC_API int vptc_sample_grid(vptc_ctx* ctx, int grid, const float* index_xyz, size_t n, float* values,
                              uint32_t* dims, float* majorants);

/* Parity hook for blackbody_radiation_xyz (src/precompute_blackbody.cpp:24-52, src/spectral.cpp:7-20): the DEVICE's lookup — table lerp,
 * exact hits, T <= 0 -> 0, non-finite -> NaN, and the direct 471-term Planck x CIE integration for T >= 49 900 K — for n temperatures
 * in kelvin; xyz receives n x {X, Y, Z}. Needs no volume. */
VPTC_API int vptc_blackbody(vptc_ctx* ctx, const float* temps_k, size_t n, float* xyz);

/* Replaces film_to_image (src/main.cpp:12-24, include/vpt/color.hpp:8-30): XYZ/N -> linear sRGB ->
 * sRGB OETF -> clamp -> *255 -> truncate, on the device, byte-identical to the host code (glibc-exact powf). Host film in, host RGB8 out. */
VPTC_API int vptc_film_to_image(vptc_ctx* ctx, const float* film_xyzw, int64_t width, int64_t height, uint8_t* rgb8);

/* vptc_film_to_image without any allocation. NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). Same kernel, same bytes out;
 * the device staging (film in, RGB8 out) is a file-scope static __device__ array of VPTC_STATIC_IMAGE_MAX_PIXELS * 19 bytes (38 MiB) that
 * every GPU which loads the library holds until the process exits. The call makes no cudaMalloc / cudaFree / allocating library call.
 * width * height > VPTC_STATIC_IMAGE_MAX_PIXELS (1920 x 1080 fits) -> VPTC_ERR_INVALID_ARGUMENT, with the limit in vptc_last_error.
 * All calls on one GPU share the array (it keeps the previous call's contents; the call overwrites what it reads, nothing needs clearing):
 * calls that overlap on the same GPU — two host threads, two contexts on one device — race; serialise them. */
#define VPTC_STATIC_IMAGE_MAX_PIXELS (1u << 21)
VPTC_API int vptc_film_to_image_static(vptc_ctx* ctx, const float* film_xyzw, int64_t width, int64_t height, uint8_t* rgb8);

/* ---- host-side helpers (no GPU needed) ----
 * Everything from here on is plain host code. It is exported by libvptcuda.so and, without any CUDA dependency, by libvpthost.so (the same
 * objects linked alone), so that tools which only need the scene reader, the .nvdb reader/writer, the procedural volumes or the PNG writer —
 * e.g. the CPU baseline arm of bench.py — never map the CUDA library. The message of a failed helper is in the calling thread's context-less
 * error slot: vptc_last_error(NULL), or vptc_host_last_error() which both libraries export. */
VPTC_API const char* vptc_host_last_error(void);
/* Replaces read_configuration (src/configuration.cpp:8-22): all keys mandatory. volume_path is
 * copied (relative, as written in the file) into volume_path_out. */
VPTC_API int vptc_scene_from_json_file(const char* path, vptc_scene* out, char* volume_path_out, size_t volume_path_cap);
VPTC_API int vptc_scene_from_json(const char* json_text, vptc_scene* out, char* volume_path_out, size_t volume_path_cap);

/* Replaces nanovdb::io::readGrid(path, name) as used by src/volume_grids.cpp:39-67. On success *grid is
 * malloc'ed (release with vptc_free). A missing grid name returns VPTC_ERR_BAD_GRID. */
VPTC_API int  vptc_nvdb_read_grid(const char* path, const char* grid_name, void** grid, size_t* bytes);
VPTC_API int  vptc_nvdb_write_grids(const char* path, const void* const* grids, size_t n_grids);
VPTC_API void vptc_free(void* p);

/* Synthetic procedural volumes (BASELINE.json configs), returned as serialized NanoVDB float grids.
 * kind: "fire" (density+temperature), "cloud" (WDAS-shaped sparse density), "dense" (all-leaf noise),
 *       "torus" (fog torus), "tiles" (test volume with active tiles at every tree level).
 * dims = index-space extent; voxel_size/seed as named; anchor_xyz (NULL = world origin) = world position
 * of the box centre (centre of the base for "fire"). temperature may be NULL. */
VPTC_API int vptc_synth_volume(const char* kind, const int32_t dims[3], double voxel_size, uint32_t seed, const double* anchor_xyz,
                               void** density, size_t* density_bytes, void** temperature, size_t* temperature_bytes);

/* Debug front-end = Volume::log_dda_trace (src/volume.cpp:194-225; the third of the reference's CSV traces, beside vptc_trace_pixel_csv and
 * vptc_trace_path_log): the per-VOXEL ground truth along a ray — X,Y,Z,T,Value,Dim_getdim,Dim_nodeinfo,Active,Maximum, one row per voxel a
 * dim-1 DDA visits from 16 index units before the ray enters the grid's index bounding box to 16 after it leaves — in the reference's number
 * formatting, so scripts/plot_dda_trace.py reads it. Host code on the caller's NanoVDB buffer (value masks and node statistics are not in the
 * baked device layout). The pixel form traces the un-jittered camera ray through the centre of pixel (px, py). A ray that misses the grid
 * writes no file, like the reference; *n_rows (may be NULL) receives the number of voxel rows. */
VPTC_API int vptc_dda_trace_csv(const void* density_grid, size_t density_bytes, const vptc_scene* scene, int64_t px, int64_t py, const char* csv_path, size_t* n_rows);
VPTC_API int vptc_dda_trace_ray_csv(const void* density_grid, size_t density_bytes, const float origin_xyz[3], const float direction_xyz[3], const char* csv_path, size_t* n_rows);

/* 8-bit RGB PNG (replaces Image<uchar,3>::save, src/image_io.cpp:18-58). */
VPTC_API int vptc_write_png(const char* path, const uint8_t* rgb8, int64_t width, int64_t height);

/* Progressive render checkpoint: the float4 film plus the next wave to render. The reference keeps both
 * only in memory (film: src/main.cpp:58-60; the wave counter: TileProvider, src/tile_provider.cpp:27-38, stopped early by
 * stop_at_next_wave :107-110 when the preview window closes); a headless multi-hour render needs them on disk.
 * The file records a fingerprint of every scene field that determines the film (everything except num_waves and
 * num_workers, so a resumed render may ask for more waves); vptc_film_load refuses a file whose size or fingerprint
 * differs (VPTC_ERR_CONFIG) and returns VPTC_ERR_IO when the file is missing or truncated. Save is atomic (tmp + rename). */
VPTC_API int vptc_film_save(const char* path, const vptc_scene* scene, const float* film_xyzw, uint32_t next_wave);
VPTC_API int vptc_film_load(const char* path, const vptc_scene* scene, float* film_xyzw, uint32_t* next_wave);

VPTC_API const char* vptc_version(void);

#ifdef __cplusplus
}
#endif
#endif /* VPTC_H */
