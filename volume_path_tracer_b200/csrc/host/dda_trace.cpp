// dda_trace.cpp — Volume::log_dda_trace (src/volume.cpp:194-225), host side: the per-VOXEL ground truth along one ray, written as the
// reference's dda_trace.csv so that scripts/plot_dda_trace.py keeps working without the window:
//     X,Y,Z,T,Value,Dim_getdim,Dim_nodeinfo,Active,Maximum
// one row per voxel a dim-1 DDA visits between 16 index units before the ray enters the density grid's index bounding box and 16 after it
// leaves (world-space maxTime 10 000 before the transform, as the reference sets it).
// The baked device layout keeps neither value masks nor node statistics, so this walks the caller's NanoVDB buffer itself
// (include/nvdb_format.h): ReadAccessor::getValue / isActive / getDim / getNodeInfo as NV/NanoVDB.h:3071-3083,3564-3577,4534-4542,
// 6096-6105,6205-6245 define them; the ray set-up is NV/math/Ray.h:64-74,154-159,258-268,398-424 and the stepping
// nanovdb::math::DDA<Ray, Coord, 1> (NV/math/HDDA.h:210-330) — the same float operations in the same order, so the rows equal the
// reference's (tests/test_reference_compiled.py compares the files).
#include "dda_trace.hpp"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <stdexcept>

#include "nvdb_build.hpp" // check_grid
#include "nvdb_format.h"

namespace vpt_host {
using namespace nvdbf;

namespace {
struct Probe { float value; uint32_t dim_getdim, dim_nodeinfo; bool active; float maximum; const LeafData* leaf; };

struct Tree {
    const GridData* grid; const RootData* root; const RootTile* tiles;
    explicit Tree(const void* data) : grid(static_cast<const GridData*>(data)), root(root_of(grid)), tiles(root_tiles(root)) {}
    Probe probe(int32_t x, int32_t y, int32_t z) const
    {
        const uint64_t key = root_key(x, y, z);
        const RootTile* rt = nullptr;
        for (uint32_t i = 0; i < root->tableSize; ++i) if (tiles[i].key == key) { rt = &tiles[i]; break; }
        // background: GetNodeInfo::get(root) = {level 3, dim 4096, root statistics}; GetState(root) = false; getDim = 4096
        if (!rt) return {root->background, 4096u, 4096u, false, root->maximum, nullptr};
        if (rt->child == 0) return {rt->value, 4096u, 4096u, rt->state > 0, rt->value, nullptr}; // root tile: min = max = value
        const auto* u = reinterpret_cast<const UpperData*>(reinterpret_cast<const char*>(root) + rt->child);
        const uint32_t un = upper_offset(x, y, z);
        if (!mask_on(u->childMask, un)) // tile of an upper node: NodeInfo of the NODE (level 2, dim 4096); getDim: 4096 if the node is flagged "skip", else 128
            return {u->table[un].value, (u->flags & 1u) ? 4096u : 128u, 4096u, mask_on(u->valueMask, un), u->maximum, nullptr};
        const auto* l = reinterpret_cast<const LowerData*>(reinterpret_cast<const char*>(u) + u->table[un].child);
        const uint32_t ln = lower_offset(x, y, z);
        const uint32_t dim_u = (u->flags & 1u) ? 4096u : 0u; // a flagged upper node answers getDim itself, whatever lies below
        if (!mask_on(l->childMask, ln))
            return {l->table[ln].value, dim_u ? dim_u : ((l->flags & 1u) ? 128u : 8u), 128u, mask_on(l->valueMask, ln), l->maximum, nullptr};
        const auto* leaf = reinterpret_cast<const LeafData*>(reinterpret_cast<const char*>(l) + l->table[ln].child);
        const uint32_t n = leaf_offset(x, y, z);
        const uint32_t dim_l = dim_u ? dim_u : ((l->flags & 1u) ? 128u : 0u);
        return {leaf->values[n], dim_l ? dim_l : ((leaf->flags & 1u) ? 8u : 1u), 8u, mask_on(leaf->valueMask, n), leaf->maximum, leaf};
    }
};

// The reference traces through a Volume, whose constructor has already run fix_majorants_for_interpolation on the grid (volume.cpp:104-170):
// a leaf's Maximum column is max(stored maximum, getValue over the 1-voxel shell around the leaf). The caller's buffer here is untouched,
// so the same maximum is formed on the fly (488 lookups per distinct leaf on the ray).
float patched_leaf_max(const Tree& tree, const LeafData* leaf)
{
    float m = leaf->maximum;
    const int32_t ox = leaf->bboxMin[0] & ~7, oy = leaf->bboxMin[1] & ~7, oz = leaf->bboxMin[2] & ~7;
    for (int32_t x = ox - 1; x <= ox + 8; ++x)
        for (int32_t y = oy - 1; y <= oy + 8; ++y)
            for (int32_t z = oz - 1; z <= oz + 8; ++z) {
                if (x >= ox && x < ox + 8 && y >= oy && y < oy + 8 && z >= oz && z < oz + 8) continue;
                m = std::max(m, tree.probe(x, y, z).value);
            }
    return m;
}

inline int32_t ifloor(float v) { return int32_t(std::floor(v)); }
} // namespace

std::string write_dda_trace(const void* grid_data, size_t bytes, const float origin[3], const float direction[3], const std::string& csv_path, size_t* rows_out)
{
    if (rows_out) *rows_out = 0;
    if (const char* bad = check_grid(grid_data, bytes)) return bad;
    const Tree tree(grid_data);
    const Map& map = tree.grid->map;
    auto mat3 = [](const float* m, const float* v, float* out) { // Map::applyInverseMapF / applyInverseJacobianF: fmaf chains (NV/math/Math.h:905-939)
        for (int r = 0; r < 3; ++r) out[r] = std::fmaf(v[0], m[r * 3], std::fmaf(v[1], m[r * 3 + 1], v[2] * m[r * 3 + 2]));
    };
    // Ray(eye, dir) has t0 = Delta = 1e-5, t1 = Maximum; the reference then sets t1 = 10000 and transforms to index space
    const float w[3] = {origin[0] - map.vecF[0], origin[1] - map.vecF[1], origin[2] - map.vecF[2]};
    float eye[3], di[3];
    mat3(map.invMatF, w, eye);
    mat3(map.invMatF, direction, di);
    const float length = std::sqrt(di[0] * di[0] + di[1] * di[1] + di[2] * di[2]), invLength = 1.0f / length;
    const float dir[3] = {di[0] * invLength, di[1] * invLength, di[2] * invLength};
    const float inv[3] = {1.0f / dir[0], 1.0f / dir[1], 1.0f / dir[2]};
    float t0 = length * 1e-5f, t1 = 10000.0f * length;
    for (int a = 0; a < 3; ++a) { // clip(CoordBBox) with max + 1
        float lo = float(tree.root->bbox[a]), hi = float(tree.root->bbox[3 + a] + 1);
        if (lo >= hi) return ""; // empty box: log_dda_trace returns before opening the file
        lo = (lo - eye[a]) * inv[a]; hi = (hi - eye[a]) * inv[a];
        if (lo > hi) std::swap(lo, hi);
        if (lo > t0) t0 = lo;
        if (hi < t1) t1 = hi;
        if (t0 > t1) return "";
    }
    t0 -= 16.0f; t1 += 16.0f;
    // stdio, not iostreams (the library links libstdc++ statically and no other host code pulls the stream locale machinery in);
    // "%g" is what an ostream prints for a float at its default precision of 6, which is what the reference's print_csv does
    std::FILE* out = std::fopen(csv_path.c_str(), "w");
    if (!out) return "cannot write \"" + csv_path + "\"";
    std::fputs("X,Y,Z,T,Value,Dim_getdim,Dim_nodeinfo,Active,Maximum\n", out);
    // DDA<Ray, Coord, 1>::init
    auto at = [&](float t, int a) { return std::fmaf(t, dir[a], eye[a]); };
    float time = t0, next[3], delta[3];
    int32_t voxel[3], step[3];
    for (int a = 0; a < 3; ++a) {
        const float pos = at(t0, a);
        voxel[a] = ifloor(pos);
        if (dir[a] == 0.0f) { next[a] = FLT_MAX; step[a] = 0; delta[a] = 0.0f; }
        else if (inv[a] > 0.0f) { step[a] = 1; next[a] = time + (float(voxel[a] + 1) - pos) * inv[a]; delta[a] = inv[a]; }
        else { step[a] = -1; next[a] = time + (float(voxel[a]) - pos) * inv[a]; delta[a] = -inv[a]; }
    }
    size_t rows = 0;
    const LeafData* last_leaf = nullptr;
    float last_leaf_max = 0.0f;
    do {
        Probe p = tree.probe(voxel[0], voxel[1], voxel[2]);
        if (p.leaf) {
            if (p.leaf != last_leaf) { last_leaf = p.leaf; last_leaf_max = patched_leaf_max(tree, p.leaf); }
            p.maximum = last_leaf_max;
        }
        std::fprintf(out, "%d,%d,%d,%g,%g,%u,%u,%d,%g\n", voxel[0], voxel[1], voxel[2], double(time), double(p.value), p.dim_getdim, p.dim_nodeinfo, p.active ? 1 : 0, double(p.maximum));
        ++rows;
        const int axis = (next[0] < next[1] && next[0] < next[2]) ? 0 : (next[1] < next[2] ? 1 : 2); // MinIndex (NV/math/Math.h:285-298)
        if (next[axis] <= time) next[axis] += time - 0.999999f * next[axis] + 1.0e-6f;              // ENFORCE_FORWARD_STEPPING
        time = next[axis];
        next[axis] += delta[axis];
        voxel[axis] += step[axis];
    } while (time <= t1 && rows < (size_t(1) << 26));
    if (rows_out) *rows_out = rows;
    const bool bad = std::ferror(out) != 0;
    return (std::fclose(out) != 0 || bad) ? "write to \"" + csv_path + "\" failed" : "";
}

} // namespace vpt_host
