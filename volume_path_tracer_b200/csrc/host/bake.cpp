// bake.cpp — NanoVDB float grid -> brick-sparse device layout (device/layout.hpp), incl. the one-time
// halo-majorant patch. Replaces Volume::Volume / fix_majorants_for_interpolation (src/volume.cpp:104-170):
// the reference mutates leaf.mMaximum inside the caller's grid buffer; we leave the caller's buffer
// untouched and write the patched maxima into the baked lower-slot table.
#include "bake.hpp"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <thread>

#include "nvdb_build.hpp" // check_grid
#include "nvdb_format.h"

namespace vpt_host {
using namespace nvdbf;
using namespace vptd;

namespace {

// Stateless top-down value lookup on the serialized tree (what ReadAccessor::getValue returns:
// the leaf voxel, else the tile value of the slot where the descent ends, else the background).
struct TreeView {
    const GridData* grid; const TreeData* tree; const RootData* root; const RootTile* tiles;
    const UpperData* upper0; const LowerData* lower0; const LeafData* leaf0;
    explicit TreeView(const void* data)
    {
        grid = static_cast<const GridData*>(data);
        tree = tree_of(grid); root = root_of(grid); tiles = root_tiles(root);
        const char* t = reinterpret_cast<const char*>(tree);
        leaf0 = reinterpret_cast<const LeafData*>(t + tree->nodeOffset[0]);
        lower0 = reinterpret_cast<const LowerData*>(t + tree->nodeOffset[1]);
        upper0 = reinterpret_cast<const UpperData*>(t + tree->nodeOffset[2]);
    }
    const UpperData* upper_of(const RootTile& rt) const { return reinterpret_cast<const UpperData*>(reinterpret_cast<const char*>(root) + rt.child); }
    const LowerData* lower_of(const UpperData* u, uint32_t n) const { return reinterpret_cast<const LowerData*>(reinterpret_cast<const char*>(u) + u->table[n].child); }
    const LeafData* leaf_of(const LowerData* l, uint32_t n) const { return reinterpret_cast<const LeafData*>(reinterpret_cast<const char*>(l) + l->table[n].child); }
    float value(int32_t x, int32_t y, int32_t z) const
    {
        const uint64_t key = root_key(x, y, z);
        const RootTile* rt = nullptr;
        for (uint32_t i = 0; i < root->tableSize; ++i) if (tiles[i].key == key) { rt = &tiles[i]; break; }
        if (!rt) return root->background;
        if (rt->child == 0) return rt->value;
        const UpperData* u = upper_of(*rt);
        const uint32_t un = upper_offset(x, y, z);
        if (!mask_on(u->childMask, un)) return u->table[un].value;
        const LowerData* l = lower_of(u, un);
        const uint32_t ln = lower_offset(x, y, z);
        if (!mask_on(l->childMask, ln)) return l->table[ln].value;
        return leaf_of(l, ln)->values[leaf_offset(x, y, z)];
    }
};

// The buffer comes from the caller across the C ABI: every offset the baker is about to follow is checked first, so a corrupt or
// hostile grid is an error (VPTC_ERR_BAD_GRID), never an out-of-bounds read. check_grid() has validated the node ARRAYS; this
// validates the root table and that every child offset lands exactly on an element of the next level's array.
void validate_tree(const TreeView& tv, size_t bytes)
{
    const char* base = reinterpret_cast<const char*>(tv.grid);
    const char* end = base + std::min<uint64_t>(tv.grid->gridSize, bytes);
    auto inside = [&](const void* p, uint64_t n) { const char* c = static_cast<const char*>(p); return c >= base && c <= end && n <= uint64_t(end - c); };
    for (int l = 0; l < 4; ++l)
        if (tv.tree->nodeOffset[l] % 8 != 0) throw std::runtime_error("bake: misaligned tree node offset");
    if (!inside(tv.root, sizeof(RootData))) throw std::runtime_error("bake: root node out of range");
    if (!inside(tv.tiles, uint64_t(tv.root->tableSize) * sizeof(RootTile))) throw std::runtime_error("bake: root table out of range");
    const uint64_t nUpper = tv.tree->nodeCount[2], nLower = tv.tree->nodeCount[1], nLeaf = tv.tree->nodeCount[0];
    // the target must be element k < count of the array starting at `first` (element size `size`)
    auto element = [&](const void* parent, int64_t off, const void* first, uint64_t size, uint64_t count) {
        if (off > int64_t(1) << 46 || off < -(int64_t(1) << 46)) return false; // keeps the sum below in range
        const int64_t rel = (reinterpret_cast<const char*>(parent) - reinterpret_cast<const char*>(first)) + off; // parent and first lie inside the buffer
        return rel >= 0 && uint64_t(rel) % size == 0 && uint64_t(rel) / size < count;
    };
    for (uint32_t i = 0; i < tv.root->tableSize; ++i)
        if (tv.tiles[i].child != 0 && !element(tv.root, tv.tiles[i].child, tv.upper0, sizeof(UpperData), nUpper))
            throw std::runtime_error("bake: root tile child offset does not address an upper node");
    for (uint64_t ui = 0; ui < nUpper; ++ui) {
        const UpperData* u = tv.upper0 + ui;
        for (uint32_t w = 0; w < 32768 / 64; ++w)
            for (uint64_t bits = u->childMask[w]; bits; bits &= bits - 1) { // set bits only
                const uint32_t n = w * 64 + uint32_t(__builtin_ctzll(bits));
                if (!element(u, u->table[n].child, tv.lower0, sizeof(LowerData), nLower)) throw std::runtime_error("bake: upper-node child offset does not address a lower node");
            }
    }
    for (uint64_t li = 0; li < nLower; ++li) {
        const LowerData* l = tv.lower0 + li;
        for (uint32_t w = 0; w < 4096 / 64; ++w)
            for (uint64_t bits = l->childMask[w]; bits; bits &= bits - 1) {
                const uint32_t n = w * 64 + uint32_t(__builtin_ctzll(bits));
                if (!element(l, l->table[n].child, tv.leaf0, sizeof(LeafData), nLeaf)) throw std::runtime_error("bake: lower-node child offset does not address a leaf");
            }
    }
}

template <class F>
void parallel_for(size_t n, F&& f)
{
    unsigned nt = std::max(1u, std::min(64u, std::thread::hardware_concurrency()));
    if (n < 256) nt = 1;
    std::atomic<size_t> next{0};
    auto work = [&]() {
        for (;;) {
            const size_t b = next.fetch_add(64);
            if (b >= n) break;
            for (size_t i = b; i < std::min(n, b + 64); ++i) f(i);
        }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; ++t) th.emplace_back(work);
    work();
    for (auto& t : th) t.join();
}

} // namespace

void bake_grid(const void* data, size_t bytes, bool patch_majorants, BakedGrid& out, bool device_bricks)
{
    const bool timing = std::getenv("VPT_BAKE_TIMING") != nullptr; // development aid: stage times on stderr
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto t_last = now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        const auto t = now();
        std::fprintf(stderr, "[bake] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t - t_last).count());
        t_last = t;
    };
    if (const char* bad = check_grid(data, bytes)) throw std::runtime_error(bad);
    const TreeView tv(data);
    validate_tree(tv, bytes);
    const GridData* g = tv.grid;
    const size_t nUpper = tv.tree->nodeCount[2], nLower = tv.tree->nodeCount[1], nLeaf = tv.tree->nodeCount[0];

    GridDev& h = out.header;
    std::memset(&h, 0, sizeof(h));
    std::memcpy(h.mat, g->map.matF, sizeof(h.mat));
    std::memcpy(h.invMat, g->map.invMatF, sizeof(h.invMat));
    std::memcpy(h.vec, g->map.vecF, sizeof(h.vec));
    for (int i = 0; i < 9; ++i) // a NaN / infinite transform turns every ray time into NaN, and `t0 >= t1` never ends a march on NaN
        if (!std::isfinite(h.mat[i]) || !std::isfinite(h.invMat[i]) || (i < 3 && !std::isfinite(h.vec[i]))) throw std::runtime_error("bake: the grid's index<->world transform is not finite");
    {   // ... and so does a finite but SINGULAR map (all zeros in a corrupt header): the index-space direction is 0, its normalisation 0 * inf = NaN.
        // Require both 3x3 maps to be invertible and each other's inverse to within fp32 slack.
        auto det3 = [](const float* m) {
            return double(m[0]) * (double(m[4]) * m[8] - double(m[5]) * m[7]) - double(m[1]) * (double(m[3]) * m[8] - double(m[5]) * m[6]) +
                   double(m[2]) * (double(m[3]) * m[7] - double(m[4]) * m[6]);
        };
        const double d_fwd = det3(h.mat), d_inv = det3(h.invMat);
        bool ok = std::isfinite(d_fwd) && std::isfinite(d_inv) && std::fabs(d_fwd) > 1e-30 && std::fabs(d_inv) > 1e-30;
        for (int r = 0; ok && r < 3; ++r)
            for (int c = 0; c < 3; ++c) { // mat * invMat ~= I (NanoVDB applies them as row-vector maps; the product test is the same either way)
                double acc = 0.0;
                for (int k = 0; k < 3; ++k) acc += double(h.mat[r * 3 + k]) * double(h.invMat[k * 3 + c]);
                if (!(std::fabs(acc - (r == c ? 1.0 : 0.0)) < 1e-3)) ok = false;
            }
        if (!ok) throw std::runtime_error("bake: the grid's index<->world transform is singular or inconsistent with its stored inverse");
    }
    for (int a = 0; a < 3; ++a) { h.bbox_min[a] = tv.root->bbox[a]; h.bbox_max[a] = tv.root->bbox[3 + a]; }
    for (int a = 0; a < 3; ++a) // keeps every device-side coordinate sum in int32 and the cell caches' INT_MIN sentinel outside the grid
        if (h.bbox_min[a] <= h.bbox_max[a] && (h.bbox_min[a] < -(1 << 30) || h.bbox_max[a] > (1 << 30)))
            throw std::runtime_error("bake: grid index bounding box exceeds +-2^30");

    // ---- tiles: index 0 = {background, inactive} ----
    out.tile_value.assign(1, tv.root->background);
    out.tile_majorant.assign(1, 0.0f);
    std::map<std::pair<uint32_t, bool>, uint32_t> tileIndex; // (value bits, active) -> index
    auto tile_of = [&](float v, bool active) -> uint32_t {
        uint32_t bits; std::memcpy(&bits, &v, 4);
        if (!active && v == tv.root->background) { uint32_t bb; std::memcpy(&bb, &tv.root->background, 4); if (bb == bits) return 0u; }
        auto key = std::make_pair(bits, active);
        auto it = tileIndex.find(key);
        if (it != tileIndex.end()) return it->second;
        const uint32_t idx = uint32_t(out.tile_value.size());
        if (idx > CELL_INDEX_MASK) throw std::runtime_error("bake: too many distinct tile values");
        out.tile_value.push_back(v);
        out.tile_majorant.push_back(active ? v : 0.0f); // probeValue(ijk, v) ? v : 0   (volume.cpp:28-34)
        tileIndex.emplace(key, idx);
        return idx;
    };

    // ---- cells: dense over the span of the root keys ----
    int32_t kmin[3] = {0, 0, 0}, kmax[3] = {0, 0, 0};
    for (uint32_t i = 0; i < tv.root->tableSize; ++i) {
        int32_t o[3]; root_key_origin(tv.tiles[i].key, o);
        for (int a = 0; a < 3; ++a) {
            const int32_t k = o[a] >> 12;
            if (i == 0) { kmin[a] = kmax[a] = k; } else { kmin[a] = std::min(kmin[a], k); kmax[a] = std::max(kmax[a], k); }
        }
    }
    uint64_t nRootCells = 1;
    for (int a = 0; a < 3; ++a) {
        h.cell_org[a] = kmin[a] * 4096;
        h.cell_dim[a] = (kmax[a] - kmin[a] + 1) * 32;
        nRootCells *= uint64_t(kmax[a] - kmin[a] + 1);
    }
    if (nRootCells > 4096) throw std::runtime_error("bake: grid spans more than 4096 root entries (index extent too large for the dense cell table)");
    const uint32_t NONE = CELL_KIND_TILE | (2u << CELL_DIM_SHIFT) | 0u; // background, dim 4096 (NanoVDB.h:3081)
    out.cells.assign(size_t(nRootCells) * 32768, NONE);
    auto cell_at = [&](int32_t x, int32_t y, int32_t z) -> uint32_t& {
        const size_t cx = size_t((x - h.cell_org[0]) >> 7), cy = size_t((y - h.cell_org[1]) >> 7), cz = size_t((z - h.cell_org[2]) >> 7);
        return out.cells[(cx * size_t(h.cell_dim[1]) + cy) * size_t(h.cell_dim[2]) + cz];
    };
    if (nLower > CELL_INDEX_MASK || nLeaf > SLOT_INDEX_MASK) throw std::runtime_error("bake: too many nodes");
    (void)nUpper;
    for (uint32_t i = 0; i < tv.root->tableSize; ++i) {
        const RootTile& rt = tv.tiles[i];
        int32_t o[3]; root_key_origin(rt.key, o);
        if (rt.child == 0) { // constant root tile: dim 4096 (NanoVDB.h:3079)
            const uint32_t code = CELL_KIND_TILE | (2u << CELL_DIM_SHIFT) | tile_of(rt.value, rt.state > 0u);
            for (int a = 0; a < 32; ++a) for (int b = 0; b < 32; ++b) for (int c = 0; c < 32; ++c) cell_at(o[0] + 128 * a, o[1] + 128 * b, o[2] + 128 * c) = code;
            continue;
        }
        const UpperData* u = tv.upper_of(rt);
        const bool uSkip = (u->flags & 1u) != 0; // node without active values: getDim returns its own dim (NanoVDB.h:3567-3568)
        for (uint32_t n = 0; n < 32768; ++n) {
            const int32_t x = o[0] + int32_t(n >> 10) * 128, y = o[1] + int32_t((n >> 5) & 31) * 128, z = o[2] + int32_t(n & 31) * 128;
            uint32_t code;
            if (mask_on(u->childMask, n)) {
                const LowerData* l = tv.lower_of(u, n);
                const uint32_t li = uint32_t(l - tv.lower0);
                const uint32_t dimc = uSkip ? 2u : ((l->flags & 1u) ? 1u : 0u);
                code = (dimc << CELL_DIM_SHIFT) | li;
            } else {
                code = CELL_KIND_TILE | ((uSkip ? 2u : 1u) << CELL_DIM_SHIFT) | tile_of(u->table[n].value, mask_on(u->valueMask, n));
            }
            cell_at(x, y, z) = code;
        }
    }

    // ---- brick order: leaf-array order (NanoVDB's breadth-first order: leaves of one lower node are contiguous, z fastest), or, for the
    // fetch-layout experiment (VPT_BAKE_BRICK_ORDER=morton, profiles/r02_fetch_layout.txt), Morton order of the leaf origins ----
    std::vector<uint32_t> brick_of(nLeaf);
    for (size_t i = 0; i < nLeaf; ++i) brick_of[i] = uint32_t(i);
    if (const char* ord = std::getenv("VPT_BAKE_BRICK_ORDER"); ord && std::string(ord) == "morton" && nLeaf > 1) {
        auto spread = [](uint64_t v) { // 21 bits -> every third bit
            v &= 0x1fffffull;
            v = (v | v << 32) & 0x1f00000000ffffull; v = (v | v << 16) & 0x1f0000ff0000ffull; v = (v | v << 8) & 0x100f00f00f00f00full;
            v = (v | v << 4) & 0x10c30c30c30c30c3ull; v = (v | v << 2) & 0x1249249249249249ull;
            return v;
        };
        std::vector<std::pair<uint64_t, uint32_t>> keyed(nLeaf);
        for (size_t i = 0; i < nLeaf; ++i) {
            const LeafData* lf = tv.leaf0 + i;
            const uint64_t x = uint64_t(uint32_t(lf->bboxMin[0] >> 3) + (1u << 20)), y = uint64_t(uint32_t(lf->bboxMin[1] >> 3) + (1u << 20)), z = uint64_t(uint32_t(lf->bboxMin[2] >> 3) + (1u << 20));
            keyed[i] = {spread(x) << 2 | spread(y) << 1 | spread(z), uint32_t(i)};
        }
        std::sort(keyed.begin(), keyed.end());
        for (size_t r = 0; r < nLeaf; ++r) brick_of[keyed[r].second] = uint32_t(r);
    }

    lap("validate + tiles + cells");
    // ---- lower slots (codes now, majorants after the patch) ----
    out.slots.assign(nLower * 4096 * 2, 0u);
    for (size_t li = 0; li < nLower; ++li) {
        const LowerData* l = tv.lower0 + li;
        for (uint32_t n = 0; n < 4096; ++n) {
            uint32_t code; float maj;
            if (mask_on(l->childMask, n)) {
                const LeafData* lf = tv.leaf_of(l, n);
                code = brick_of[size_t(lf - tv.leaf0)];
                maj = lf->maximum; // patched below
            } else {
                const uint32_t ti = tile_of(l->table[n].value, mask_on(l->valueMask, n));
                code = SLOT_KIND_TILE | ti;
                maj = out.tile_majorant[ti];
            }
            out.slots[(li * 4096 + n) * 2] = code;
            std::memcpy(&out.slots[(li * 4096 + n) * 2 + 1], &maj, 4);
        }
    }

    lap("lower slots");
    h.n_lower = uint32_t(nLower); h.n_bricks = uint32_t(nLeaf); h.n_tiles = uint32_t(out.tile_value.size());
    h.background_tile = 0;
    out.device_bricks = device_bricks; out.patch_majorants = patch_majorants;
    out.raw = data; out.raw_bytes = size_t(std::min<uint64_t>(g->gridSize, bytes));
    bool identity = true;
    for (size_t i = 0; i < nLeaf && identity; ++i) identity = brick_of[i] == uint32_t(i);
    if (!identity) out.brick_of = brick_of; else out.brick_of.clear();
    if (device_bricks) { out.bricks.clear(); out.brick_max.clear(); return; } // the brick stage runs on the device (device/vpt_bake.cu)
    // ---- bricks with +1 apron, and patched maxima ----
    out.bricks.assign(nLeaf * size_t(BRICK_STRIDE), 0.0f);
    out.brick_max.assign(nLeaf, 0.0f);
    parallel_for(nLeaf, [&](size_t bi) {
        const LeafData* lf = tv.leaf0 + bi;
        const int32_t o[3] = {lf->bboxMin[0] & ~7, lf->bboxMin[1] & ~7, lf->bboxMin[2] & ~7};
        float dst[729]; // the 9^3 apron brick, x-major; laid out into the device format below
        for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) for (int k = 0; k < 9; ++k) {
            const bool core = i < 8 && j < 8 && k < 8;
            dst[(i * 9 + j) * 9 + k] = core ? lf->values[(i << 6) | (j << 3) | k] : tv.value(o[0] + i, o[1] + j, o[2] + k);
        }
        float* dev = &out.bricks[size_t(brick_of[bi]) * size_t(BRICK_STRIDE)];
#if VPT_BRICK_PAIRS == 2
        for (int i = 0; i < 8; ++i) for (int j = 0; j < 8; ++j) for (int k = 0; k < 8; ++k)
            for (int c = 0; c < 8; ++c) dev[brick_offset(i, j, k) + c] = dst[((i + (c >> 2)) * 9 + (j + ((c >> 1) & 1))) * 9 + (k + (c & 1))];
#else
        for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) {
#if VPT_BRICK_PAIRS
            for (int k = 0; k < 8; ++k) { dev[brick_offset(i, j, k)] = dst[(i * 9 + j) * 9 + k]; dev[brick_offset(i, j, k) + 1] = dst[(i * 9 + j) * 9 + k + 1]; }
#else
            for (int k = 0; k < 9; ++k) dev[brick_offset(i, j, k)] = dst[(i * 9 + j) * 9 + k];
#endif
        }
#endif
        float m = lf->maximum; // stats over ACTIVE voxels only (NV/tools/GridStats.h:716-730)
        if (patch_majorants) {
            // max over the 1-voxel shell [o-1, o+8]^3 \ leaf, read through getValue (volume.cpp:128-156)
            for (int i = -1; i <= 8; ++i) for (int j = -1; j <= 8; ++j) for (int k = -1; k <= 8; ++k) {
                if (i >= 0 && i < 8 && j >= 0 && j < 8 && k >= 0 && k < 8) continue;
                float v;
                if (i >= 0 && j >= 0 && k >= 0) v = dst[(i * 9 + j) * 9 + k]; // already fetched for the apron
                else v = tv.value(o[0] + i, o[1] + j, o[2] + k);
                m = std::max(m, v);
            }
        }
        out.brick_max[bi] = m;
    });
    lap("bricks + halo patch");
    std::vector<float> max_by_brick(nLeaf);
    for (size_t i = 0; i < nLeaf; ++i) max_by_brick[brick_of[i]] = out.brick_max[i];
    for (size_t e = 0; e < nLower * 4096; ++e) {
        const uint32_t code = out.slots[e * 2];
        if (!(code & SLOT_KIND_TILE)) std::memcpy(&out.slots[e * 2 + 1], &max_by_brick[code], 4);
    }
#if VPT_SPLIT_MAJ
    out.slot_majorant.assign(nLower * 4096, 0.0f);
    for (size_t li = 0; li < nLower; ++li)
        for (uint32_t n = 0; n < 4096; ++n)
            std::memcpy(&out.slot_majorant[li * 4096 + maj_index(n >> 8, (n >> 4) & 15u, n & 15u)], &out.slots[(li * 4096 + n) * 2 + 1], 4);
#endif
    lap("majorants into slots");
}

} // namespace vpt_host
