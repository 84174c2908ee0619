// bake.hpp — host-side result of baking a NanoVDB float grid into the device layout (device/layout.hpp).
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

#include "../device/layout.hpp"

namespace vpt_host {

struct BakedGrid {
    vptd::GridDev header;             // pointers left null; filled by the uploader
    std::vector<uint32_t> cells;
    std::vector<uint32_t> slots;      // 2 x uint32 per slot: {code, majorant bits}
    std::vector<float> tile_value, tile_majorant;
    std::vector<float> bricks;        // n_bricks * BRICK_STRIDE
    std::vector<float> slot_majorant; // VPT_SPLIT_MAJ builds: the slots' majorants alone, maj_index order (else empty)
    // device_bricks builds (the default of vptc_set_volume): `bricks` and `brick_max` stay empty and the slots carry the leaves' UNPATCHED
    // maxima; the uploader stages the caller's grid and runs device/vpt_bake.cu, which needs:
    bool device_bricks = false, patch_majorants = false;
    const void* raw = nullptr; size_t raw_bytes = 0;   // the caller's serialized grid (must outlive the staging)
    std::vector<uint32_t> brick_of;                    // leaf index -> brick index (empty = identity)
    std::vector<float> brick_max;     // patched leaf maxima, leaf-array order (also folded into `slots`)
    size_t device_bytes() const { return cells.size() * 4 + slots.size() * 4 + slot_majorant.size() * 4 + (tile_value.size() + tile_majorant.size()) * 4 + size_t(header.n_bricks) * size_t(vptd::BRICK_STRIDE) * 4; }
};

// Throws std::runtime_error on a malformed grid. patch_majorants = apply fix_majorants_for_interpolation
// (density grid: yes; temperature grid: maxima unused).
void bake_grid(const void* grid, size_t bytes, bool patch_majorants, BakedGrid& out, bool device_bricks = false);

} // namespace vpt_host
