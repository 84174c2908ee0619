// layout.hpp — the brick-sparse HBM layout a NanoVDB float grid is baked into (shared host/device).
//
// What it must reproduce, and how:
//   getValue(ijk)  (NV/NanoVDB.h GetValue)      cells -> lower slots -> 9^3 apron bricks / tile values
//   getDim(ijk)    (NV/NanoVDB.h:3071-3083,3564-3577) ONE load: the 128^3-cell entry carries the dim code
//   majorant(ijk)  (src/volume.cpp:18-36)        lower-slot entry carries the patched leaf maximum
//
// Level 0 "cells": a DENSE 3-D array of 128^3-voxel cells over the span of the grid's root entries
//   (root keys = ijk >> 12; each root entry = 32^3 cells). The sparse root table and the 32^3 upper
//   tables of NanoVDB collapse into this one array (1 MB for a 2k^3 cloud), so a lookup is
//   cell -> lower slot -> voxel: two dependent 4/8-byte loads before the voxel data instead of three
//   pointer hops plus a linear root search.
// Level 1 "lower slots": per lower node 16^3 entries of 8 bytes {code, majorant}.
// Level 2 "bricks": per leaf 9x9x9 floats = the 8^3 voxels plus a +1 apron in x, y, z (values read
//   through getValue at bake time, so whatever lies beyond the leaf — another leaf, a tile, background —
//   is already folded in). A trilinear stencil whose base voxel lies in a leaf touches exactly one brick.
//   Brick stride is padded to 736 floats (2944 B = 92 sectors of 32 B).
#pragma once
#include <cstdint>

namespace vptd {

constexpr uint32_t CELL_KIND_TILE = 0x80000000u;  // else: lower-node index
constexpr uint32_t CELL_DIM_SHIFT = 28;           // 2 bits: 0 -> 8, 1 -> 128, 2 -> 4096
constexpr uint32_t CELL_DIM_MASK = 0x30000000u;
constexpr uint32_t CELL_INDEX_MASK = 0x0fffffffu;
constexpr uint32_t SLOT_KIND_TILE = 0x80000000u;  // else: brick index
constexpr uint32_t SLOT_INDEX_MASK = 0x7fffffffu;

struct Slot { uint32_t code; float majorant; }; // 8 bytes, loaded as one 64-bit word

constexpr int BRICK_EDGE = 9;
// Fetch-layout switch (profiles/r02_fetch_layout.txt). 0, the default: rows of 9 floats, a trilinear request = 8 scalar loads touching
// 4 rows. 1: every (x, y) row stores its 8 OVERLAPPING z-pairs {v[z], v[z+1]}, z = 0..7, as 8-byte-aligned float2 — a request is
// 4 x LDG.64 that never straddle a 32-byte sector, at 1.76x the brick bytes. 2: see below.
#ifndef VPT_BRICK_PAIRS
#define VPT_BRICK_PAIRS 0
#endif
#if defined(__CUDACC__)
#define VPT_LAYOUT_HD __host__ __device__
#else
#define VPT_LAYOUT_HD
#endif
#if VPT_BRICK_PAIRS == 2
// 2: every voxel of the leaf stores its whole 2x2x2 stencil {v000, v001, v010, v011, v100, v101, v110, v111} as one 32-byte-aligned
// group — a trilinear request is ONE sector (2 x LDG.128), at 5.6x the brick bytes (no apron: the +1 neighbours are folded in at bake time).
constexpr int BRICK_ROW = 64;                 // floats per (x, y) row: 8 voxels x 8 corners
constexpr int BRICK_STRIDE = 8 * 8 * 8 * 8;   // 4096 floats = 16 KiB = 512 sectors
VPT_LAYOUT_HD constexpr int brick_offset(int x, int y, int z) { return ((x * 8 + y) * 8 + z) * 8; } // x, y, z in 0..7: the voxel's own value comes first
#elif VPT_BRICK_PAIRS
constexpr int BRICK_ROW = 16;                 // floats per (x, y) row: 8 float2
constexpr int BRICK_STRIDE = 9 * 9 * 16;      // 1296 floats = 5184 B = 162 sectors
VPT_LAYOUT_HD constexpr int brick_offset(int x, int y, int z) { return (x * 9 + y) * BRICK_ROW + 2 * z; } // x, y in 0..8, z in 0..7: first float of the pair
#else
constexpr int BRICK_ROW = 9;
constexpr int BRICK_STRIDE = 736;             // floats; >= 9^3 = 729 (2944 B = 92 sectors of 32 B)
VPT_LAYOUT_HD constexpr int brick_offset(int x, int y, int z) { return (x * 9 + y) * BRICK_ROW + z; }       // x, y, z in 0..8
#endif

// Majorant-table switch (profiles/r02_kernel_ab.txt). 0, the default: the HDDA reads the majorant out of the 8-byte {code, majorant} slot.
// 1: a second table of 4-byte majorants in the same order (x << 8 | y << 4 | z inside a lower node): twice the entries per line.
// 2: that table in 2x4x4 blocks per 128-byte line, so that a step to a neighbouring leaf in ANY axis mostly stays in the line.
#ifndef VPT_SPLIT_MAJ
#define VPT_SPLIT_MAJ 0
#endif
VPT_LAYOUT_HD constexpr uint32_t maj_index(uint32_t xs, uint32_t ys, uint32_t zs) // leaf position inside its lower node, 0..15 each
{
#if VPT_SPLIT_MAJ == 2
    return (xs >> 2) << 10 | (ys >> 2) << 8 | (zs >> 2) << 6 | (xs & 3u) << 4 | (ys & 3u) << 2 | (zs & 3u);
#else
    return xs << 8 | ys << 4 | zs;
#endif
}

struct GridDev {
    // NanoVDB Map, float members verbatim (NV/NanoVDB.h:1375-1381)
    float mat[9], invMat[9], vec[3];
    int32_t bbox_min[3], bbox_max[3]; // indexBBox(): AABB of active values, inclusive
    int32_t cell_org[3];              // index-space origin of cell (0,0,0): multiple of 4096
    int32_t cell_dim[3];              // number of 128^3 cells per axis (multiple of 32)
    uint32_t background_tile;         // tile index of {background, inactive}: always 0
    const uint32_t* cells;            // [cell_dim x*y*z]
    const Slot* slots;                // [nLower * 4096]
    const float* tile_value;          // [nTiles] value getValue returns inside the tile
    const float* tile_majorant;       // [nTiles] active ? value : 0        (volume.cpp:28-34)
    const float* bricks;              // [nBricks * BRICK_STRIDE]
    uint32_t n_lower, n_bricks, n_tiles, pad_;
    const float* slot_majorant;       // [nLower * 4096] the slots' majorants alone, maj_index order (VPT_SPLIT_MAJ builds; else null)
};

// vpt::Configuration flattened to what the kernels read, plus host-precomputed constants.
struct SceneDev {
    // camera (src/camera.cpp:45-57): raster -> world direction affine, row-major linear + translation
    float cam_m[9], cam_t[3], cam_pos[3];
    float imaging_ratio;
    // lights
    float inf_light[3];      // infinite_light.xyz * multiplier         (worker.cpp:199)
    float dist_Li[3];        // distant_light.xyz * multiplier          (worker.cpp:55)
    float dist_wi[3];        // distant_light.inv_direction.normalized() (worker.cpp:54)
    int32_t dist_light_on;   // Li != 0                                  (worker.cpp:57)
    // volume parameters
    float sigma_a, sigma_s, sigma_t, g, le_scale, temperature_offset, temperature_scale;
    uint32_t max_depth;
    int32_t use_jitter, has_temperature;
    int32_t single_pixel_enabled; int64_t single_pixel_x, single_pixel_y;
    // tiling (src/tile_provider.cpp:14-31,95-105)
    int64_t W, H, tw, th, ntx, nty;
    uint64_t ntiles;
    uint32_t seed;
};

// One render call's share of the (tile, wave) job space. Calls are published into a small device-side ring; the persistent render kernel
// takes jobs from its own call's range and, when that is exhausted, carries on into the next published one instead of draining — so a
// progressive render made of many small calls keeps every lane busy across call boundaries (DESIGN.md "launch drain"). A range may also
// cover only a stripe of tiles [tile_begin, tile_begin + tile_count) (multi-GPU tile sharding, src/tile_provider.cpp:95-105).
//   job j of the range (no tile_list):  k = j / tile_count,  i = j % tile_count,  tile = tile_begin + (i / ntx) * tile_row_stride * ntx + i % ntx,
//                        jid = (first_wave - 1 + k * wave_stride) * ntiles + tile
//   (tile_row_stride 1: the contiguous stripe; G: every G-th ROW of tiles starting at row tile_begin / ntx — interleaved stripes balance the load)
constexpr unsigned RING_SLOTS = 64;
struct RangeDesc { // 96 bytes
    unsigned long long counter;  // next job index of the range = TileProvider::m_job_idx (tile_provider.cpp:28); device atomics only
    unsigned long long done;     // jobs finished; the lane that makes it n_jobs raises host_done[slot]
    unsigned long long n_jobs;   // tile_count * n_waves
    float* film;                 // H*W*4 floats this range accumulates into
    unsigned long long tile_begin, tile_count;
    uint32_t first_wave, wave_stride;
    uint32_t rid, tile_row_stride; // absolute range number (slot = rid % RING_SLOTS); rows of tiles to skip between the stripe's rows (>= 1)
    // statistics launches only (the STATS instantiation of the kernel; adaptive rendering):
    const uint32_t* tile_list;   // when set: the range covers tile_list[0 .. tile_count) instead of a contiguous stripe
    float* sum_y2;               // when set: H*W floats, += (imaging_ratio * L.y)^2 per pixel-sample (second moment of the luminance)
    // tile order (longest jobs first, DESIGN.md "launch drain"): with a tile_list the range is walked TILE-MAJOR — job j is wave j % n_waves of
    // tile_list[j / n_waves] — so that every wave of the expensive tiles starts at once and the cheap tiles fill the tail;
    uint32_t* tile_cost;         // when set: [ntiles] += SM clock ticks >> 10 a job of the tile took (how the order is learnt)
    uint32_t n_waves, pad_;      // n_jobs / tile_count
};
static_assert(sizeof(RangeDesc) == 96, "RangeDesc");

struct CountersDev {
    unsigned long long pixel_samples, density_evals, temperature_evals, majorant_fetches, rng_draws, scatter_events, shadow_rays;
    // warp-scheduling diagnostics of the sched kernel (COUNT builds): iterations, and per stage {times run, lanes served}
    unsigned long long sched_iters, sched_runs[3], sched_lanes[3], sched_waiting, sched_exited;
};

} // namespace vptd
