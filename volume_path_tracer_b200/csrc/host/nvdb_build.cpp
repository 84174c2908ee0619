// nvdb_build.cpp — see nvdb_build.hpp.
#include "nvdb_build.hpp"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <numeric>
#include <stdexcept>

namespace vpt_host {
using namespace nvdbf;

namespace {

struct Acc { // statistics of active values below a node
    double  mn = std::numeric_limits<double>::infinity(), mx = -std::numeric_limits<double>::infinity();
    double  sum = 0, sum2 = 0;
    uint64_t n = 0;
    int32_t bmin[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, bmax[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
    void add(double v, uint64_t count = 1)
    {
        mn = std::min(mn, v); mx = std::max(mx, v);
        sum += v * double(count); sum2 += v * v * double(count); n += count;
    }
    void box(const int32_t lo[3], const int32_t hi[3])
    {
        for (int a = 0; a < 3; ++a) { bmin[a] = std::min(bmin[a], lo[a]); bmax[a] = std::max(bmax[a], hi[a]); }
    }
    void merge(const Acc& o)
    {
        if (o.n == 0) return;
        mn = std::min(mn, o.mn); mx = std::max(mx, o.mx); sum += o.sum; sum2 += o.sum2; n += o.n;
        box(o.bmin, o.bmax);
    }
    float avg() const { return n ? float(sum / double(n)) : 0.0f; }
    float dev() const
    {
        if (!n) return 0.0f;
        const double m = sum / double(n), var = std::max(0.0, sum2 / double(n) - m * m);
        return float(std::sqrt(var));
    }
};

struct Key3 {
    int32_t x, y, z;
    bool operator<(const Key3& o) const { return x != o.x ? x < o.x : (y != o.y ? y < o.y : z < o.z); }
    bool operator==(const Key3& o) const { return x == o.x && y == o.y && z == o.z; }
};

size_t find_key(const std::vector<Key3>& v, Key3 k)
{
    auto it = std::lower_bound(v.begin(), v.end(), k);
    if (it == v.end() || !(*it == k)) throw std::runtime_error("nvdb_build: internal key lookup failed");
    return size_t(it - v.begin());
}

void sort_unique(std::vector<Key3>& v)
{
    std::sort(v.begin(), v.end());
    v.erase(std::unique(v.begin(), v.end()), v.end());
}

template <class NodeT>
void init_internal(NodeT* n, float background)
{
    std::memset(n, 0, sizeof(NodeT));
    for (int i = 0; i < NodeT::SIZE; ++i) n->table[i].value = background; // upper 4 bytes of the union stay 0
}

template <class NodeT>
void finish_internal(NodeT* n, const Acc& a, float background)
{
    if (a.n == 0) {
        n->flags = 1u; // no active value below: "disable rendering" (GridStats.h:699-702)
        n->minimum = n->maximum = background;
        n->average = n->stdDevi = 0.0f;
        // empty CoordBBox = (max int, min int) as nanovdb::CoordBBox() constructs it
        for (int c = 0; c < 3; ++c) { n->bbox[c] = INT32_MAX; n->bbox[3 + c] = INT32_MIN; }
    } else {
        n->flags = 2u;
        n->minimum = float(a.mn); n->maximum = float(a.mx); n->average = a.avg(); n->stdDevi = a.dev();
        for (int c = 0; c < 3; ++c) { n->bbox[c] = a.bmin[c]; n->bbox[3 + c] = a.bmax[c]; }
    }
}

} // namespace

std::vector<uint8_t> build_grid(std::vector<BuildLeaf>& leaves, const std::vector<BuildTile>& tiles,
                                const BuildParams& P)
{
    // ---- topology: which lower / upper nodes and root entries exist ----
    std::vector<Key3> lowerKeys, upperKeys, rootTileKeys;
    lowerKeys.reserve(leaves.size() / 64 + 16);
    {
        Key3 last{INT32_MIN, INT32_MIN, INT32_MIN};
        for (const auto& l : leaves) {
            if ((l.origin[0] | l.origin[1] | l.origin[2]) & 7) throw std::runtime_error("nvdb_build: leaf origin not 8-aligned");
            Key3 k{l.origin[0] & ~127, l.origin[1] & ~127, l.origin[2] & ~127};
            if (!(k == last)) { lowerKeys.push_back(k); last = k; }
        }
    }
    for (const auto& t : tiles) {
        if (t.level == 1) lowerKeys.push_back({t.origin[0] & ~127, t.origin[1] & ~127, t.origin[2] & ~127});
    }
    sort_unique(lowerKeys);
    for (const auto& k : lowerKeys) upperKeys.push_back({k.x & ~4095, k.y & ~4095, k.z & ~4095});
    for (const auto& t : tiles) {
        if (t.level == 2) upperKeys.push_back({t.origin[0] & ~4095, t.origin[1] & ~4095, t.origin[2] & ~4095});
        if (t.level == 3) rootTileKeys.push_back({t.origin[0] & ~4095, t.origin[1] & ~4095, t.origin[2] & ~4095});
        if (t.level < 1 || t.level > 3) throw std::runtime_error("nvdb_build: tile level must be 1..3");
    }
    sort_unique(upperKeys);
    sort_unique(rootTileKeys);
    for (const auto& k : rootTileKeys)
        if (std::binary_search(upperKeys.begin(), upperKeys.end(), k)) throw std::runtime_error("nvdb_build: root tile overlaps a child node");

    // ---- leaf order: by lower node, then by slot ----
    std::vector<uint32_t> order(leaves.size());
    std::iota(order.begin(), order.end(), 0u);
    std::vector<uint64_t> sortKey(leaves.size());
    for (size_t i = 0; i < leaves.size(); ++i) {
        const auto& l = leaves[i];
        const size_t li = find_key(lowerKeys, {l.origin[0] & ~127, l.origin[1] & ~127, l.origin[2] & ~127});
        sortKey[i] = (uint64_t(li) << 12) | lower_offset(l.origin[0], l.origin[1], l.origin[2]);
    }
    std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return sortKey[a] < sortKey[b]; });
    for (size_t i = 1; i < order.size(); ++i)
        if (sortKey[order[i]] == sortKey[order[i - 1]]) throw std::runtime_error("nvdb_build: duplicate leaf");

    const size_t nLeaf = leaves.size(), nLower = lowerKeys.size(), nUpper = upperKeys.size();
    const size_t nRoot = nUpper + rootTileKeys.size();
    const size_t offTree = sizeof(GridData), offRoot = offTree + sizeof(TreeData);
    const size_t offUpper = offRoot + sizeof(RootData) + nRoot * sizeof(RootTile);
    const size_t offLower = offUpper + nUpper * sizeof(UpperData);
    const size_t offLeaf = offLower + nLower * sizeof(LowerData);
    const size_t total = offLeaf + nLeaf * sizeof(LeafData);

    std::vector<uint8_t> buf(total);
    uint8_t* base = buf.data();
    auto* grid = reinterpret_cast<GridData*>(base);
    auto* tree = reinterpret_cast<TreeData*>(base + offTree);
    auto* root = reinterpret_cast<RootData*>(base + offRoot);
    auto* rtiles = reinterpret_cast<RootTile*>(root + 1);
    auto* uppers = reinterpret_cast<UpperData*>(base + offUpper);
    auto* lowers = reinterpret_cast<LowerData*>(base + offLower);
    auto* oleaves = reinterpret_cast<LeafData*>(base + offLeaf);

    for (size_t i = 0; i < nUpper; ++i) init_internal(&uppers[i], P.background);
    for (size_t i = 0; i < nLower; ++i) init_internal(&lowers[i], P.background);

    std::vector<Acc> lowerAcc(nLower), upperAcc(nUpper);
    uint64_t activeVoxels = 0;
    uint32_t tileCount[3] = {0, 0, 0};

    // ---- leaves ----
    for (size_t k = 0; k < nLeaf; ++k) {
        const BuildLeaf& src = leaves[order[k]];
        LeafData* dst = &oleaves[k];
        std::memset(dst, 0, sizeof(LeafData));
        std::memcpy(dst->valueMask, src.mask, sizeof(src.mask));
        std::memcpy(dst->values, src.values, sizeof(src.values));
        Acc a;
        int lo[3] = {8, 8, 8}, hi[3] = {-1, -1, -1};
        for (uint32_t n = 0; n < 512; ++n) {
            if (!mask_on(src.mask, n)) continue;
            a.add(src.values[n]);
            const int c[3] = {int(n >> 6), int((n >> 3) & 7), int(n & 7)};
            for (int ax = 0; ax < 3; ++ax) { lo[ax] = std::min(lo[ax], c[ax]); hi[ax] = std::max(hi[ax], c[ax]); }
        }
        if (a.n) {
            for (int ax = 0; ax < 3; ++ax) {
                dst->bboxMin[ax] = src.origin[ax] + lo[ax];
                dst->bboxDif[ax] = uint8_t(hi[ax] - lo[ax]);
            }
            dst->flags = 2u | 16u; // has bbox, has stats
            dst->minimum = float(a.mn); dst->maximum = float(a.mx); dst->average = a.avg(); dst->stdDevi = a.dev();
            int32_t blo[3], bhi[3];
            for (int ax = 0; ax < 3; ++ax) { blo[ax] = src.origin[ax] + lo[ax]; bhi[ax] = src.origin[ax] + hi[ax]; }
            a.box(blo, bhi);
        } else {
            for (int ax = 0; ax < 3; ++ax) dst->bboxMin[ax] = src.origin[ax];
            dst->flags = 0u;
            dst->minimum = dst->maximum = P.background;
        }
        activeVoxels += a.n;
        const size_t li = size_t(sortKey[order[k]] >> 12);
        const uint32_t slot = uint32_t(sortKey[order[k]] & 4095);
        LowerData* lw = &lowers[li];
        mask_set(lw->childMask, slot);
        lw->table[slot].child = int64_t(reinterpret_cast<uint8_t*>(dst) - reinterpret_cast<uint8_t*>(lw));
        lowerAcc[li].merge(a);
    }

    // ---- level-1 and level-2 tiles ----
    for (const auto& t : tiles) {
        if (t.level == 1) {
            if ((t.origin[0] | t.origin[1] | t.origin[2]) & 7) throw std::runtime_error("nvdb_build: level-1 tile not 8-aligned");
            const size_t li = find_key(lowerKeys, {t.origin[0] & ~127, t.origin[1] & ~127, t.origin[2] & ~127});
            const uint32_t slot = lower_offset(t.origin[0], t.origin[1], t.origin[2]);
            LowerData* lw = &lowers[li];
            if (mask_on(lw->childMask, slot)) throw std::runtime_error("nvdb_build: level-1 tile overlaps a leaf");
            lw->table[slot].child = 0; lw->table[slot].value = t.value;
            if (t.active) {
                mask_set(lw->valueMask, slot);
                lowerAcc[li].add(t.value, 512);
                const int32_t hi[3] = {t.origin[0] + 7, t.origin[1] + 7, t.origin[2] + 7};
                lowerAcc[li].box(t.origin, hi);
                activeVoxels += 512; ++tileCount[0];
            }
        }
    }
    // ---- lower nodes into upper nodes ----
    for (size_t li = 0; li < nLower; ++li) {
        const Key3 k = lowerKeys[li];
        finish_internal(&lowers[li], lowerAcc[li], P.background);
        const size_t ui = find_key(upperKeys, {k.x & ~4095, k.y & ~4095, k.z & ~4095});
        const uint32_t slot = upper_offset(k.x, k.y, k.z);
        UpperData* up = &uppers[ui];
        mask_set(up->childMask, slot);
        up->table[slot].child = int64_t(reinterpret_cast<uint8_t*>(&lowers[li]) - reinterpret_cast<uint8_t*>(up));
        upperAcc[ui].merge(lowerAcc[li]);
    }
    for (const auto& t : tiles) {
        if (t.level == 2) {
            if ((t.origin[0] | t.origin[1] | t.origin[2]) & 127) throw std::runtime_error("nvdb_build: level-2 tile not 128-aligned");
            const size_t ui = find_key(upperKeys, {t.origin[0] & ~4095, t.origin[1] & ~4095, t.origin[2] & ~4095});
            const uint32_t slot = upper_offset(t.origin[0], t.origin[1], t.origin[2]);
            UpperData* up = &uppers[ui];
            if (mask_on(up->childMask, slot)) throw std::runtime_error("nvdb_build: level-2 tile overlaps a lower node");
            up->table[slot].child = 0; up->table[slot].value = t.value;
            if (t.active) {
                mask_set(up->valueMask, slot);
                upperAcc[ui].add(t.value, uint64_t(128) * 128 * 128);
                const int32_t hi[3] = {t.origin[0] + 127, t.origin[1] + 127, t.origin[2] + 127};
                upperAcc[ui].box(t.origin, hi);
                activeVoxels += uint64_t(128) * 128 * 128; ++tileCount[1];
            }
        }
    }
    // ---- root ----
    Acc rootAcc;
    std::memset(root, 0, sizeof(RootData));
    std::memset(rtiles, 0, nRoot * sizeof(RootTile));
    for (size_t ui = 0; ui < nUpper; ++ui) {
        finish_internal(&uppers[ui], upperAcc[ui], P.background);
        RootTile& rt = rtiles[ui];
        rt.key = root_key(upperKeys[ui].x, upperKeys[ui].y, upperKeys[ui].z);
        rt.child = int64_t(reinterpret_cast<uint8_t*>(&uppers[ui]) - reinterpret_cast<uint8_t*>(root));
        rt.state = 0; rt.value = P.background;
        rootAcc.merge(upperAcc[ui]);
    }
    {
        size_t ri = nUpper;
        for (const auto& t : tiles) {
            if (t.level != 3) continue;
            if ((t.origin[0] | t.origin[1] | t.origin[2]) & 4095) throw std::runtime_error("nvdb_build: root tile not 4096-aligned");
            const uint64_t key = root_key(t.origin[0], t.origin[1], t.origin[2]);
            for (size_t j = nUpper; j < ri; ++j) if (rtiles[j].key == key) throw std::runtime_error("nvdb_build: duplicate root tile");
            RootTile& rt = rtiles[ri++];
            rt.key = key; rt.child = 0; rt.state = t.active ? 1u : 0u; rt.value = t.value;
            if (t.active) {
                rootAcc.add(t.value, uint64_t(4096) * 4096 * 4096);
                const int32_t hi[3] = {t.origin[0] + 4095, t.origin[1] + 4095, t.origin[2] + 4095};
                rootAcc.box(t.origin, hi);
                activeVoxels += uint64_t(4096) * 4096 * 4096; ++tileCount[2];
            }
        }
    }
    root->tableSize = uint32_t(nRoot);
    root->background = P.background;
    if (rootAcc.n) {
        root->minimum = float(rootAcc.mn); root->maximum = float(rootAcc.mx);
        root->average = rootAcc.avg(); root->stdDevi = rootAcc.dev();
        for (int c = 0; c < 3; ++c) { root->bbox[c] = rootAcc.bmin[c]; root->bbox[3 + c] = rootAcc.bmax[c]; }
    } else {
        root->minimum = root->maximum = P.background;
        for (int c = 0; c < 3; ++c) { root->bbox[c] = INT32_MAX; root->bbox[3 + c] = INT32_MIN; }
    }

    // ---- tree ----
    tree->nodeOffset[0] = int64_t(offLeaf - offTree);
    tree->nodeOffset[1] = int64_t(offLower - offTree);
    tree->nodeOffset[2] = int64_t(offUpper - offTree);
    tree->nodeOffset[3] = int64_t(offRoot - offTree);
    tree->nodeCount[0] = uint32_t(nLeaf); tree->nodeCount[1] = uint32_t(nLower); tree->nodeCount[2] = uint32_t(nUpper);
    tree->tileCount[0] = tileCount[0]; tree->tileCount[1] = tileCount[1]; tree->tileCount[2] = tileCount[2];
    tree->voxelCount = activeVoxels;

    // ---- grid ----
    std::memset(grid, 0, sizeof(GridData));
    grid->magic = MAGIC_GRID;
    grid->checksum = ~uint64_t(0); // disabled
    grid->version = pack_version(ABI_MAJOR, ABI_MINOR, ABI_PATCH);
    grid->flags = GF_HAS_BBOX | GF_HAS_MINMAX | GF_HAS_AVERAGE | GF_HAS_STDDEV | GF_BREADTH_FIRST;
    grid->gridIndex = 0; grid->gridCount = 1;
    grid->gridSize = total;
    std::strncpy(grid->gridName, P.name.c_str(), 255);
    {
        const double s = P.voxel_size, is = 1.0 / s;
        Map& m = grid->map;
        for (int i = 0; i < 9; ++i) { m.matF[i] = m.invMatF[i] = 0.0f; m.matD[i] = m.invMatD[i] = 0.0; }
        for (int i = 0; i < 3; ++i) {
            m.matD[4 * i] = s; m.invMatD[4 * i] = is;
            m.matF[4 * i] = float(s); m.invMatF[4 * i] = float(is); // Map::set casts the doubles (NanoVDB.h:1521-1538)
            m.vecD[i] = P.translation[i]; m.vecF[i] = float(P.translation[i]);
        }
        m.taperF = 1.0f; m.taperD = 1.0;
        for (int i = 0; i < 3; ++i) grid->voxelSize[i] = s;
    }
    if (rootAcc.n) {
        for (int c = 0; c < 3; ++c) {
            grid->worldBBox[c] = double(root->bbox[c]) * P.voxel_size + P.translation[c];
            grid->worldBBox[3 + c] = double(root->bbox[3 + c] + 1) * P.voxel_size + P.translation[c];
        }
    } else {
        for (int c = 0; c < 3; ++c) { grid->worldBBox[c] = std::numeric_limits<double>::max(); grid->worldBBox[3 + c] = -std::numeric_limits<double>::max(); }
    }
    grid->gridClass = P.grid_class;
    grid->gridType = GRIDTYPE_FLOAT;
    grid->blindMetadataOffset = int64_t(total);
    grid->blindMetadataCount = 0;
    return buf;
}

const char* check_grid(const void* data, size_t bytes)
{
    if (!data || bytes < sizeof(GridData) + sizeof(TreeData) + sizeof(RootData)) return "buffer too small for a NanoVDB grid";
    if (reinterpret_cast<uintptr_t>(data) % 8 != 0) return "grid buffer must be 8-byte aligned";
    const auto* g = static_cast<const GridData*>(data);
    if (g->magic != MAGIC_GRID && g->magic != MAGIC_NUMB) return "bad magic number (not a NanoVDB grid)";
    if (version_major(g->version) != ABI_MAJOR) return "unsupported NanoVDB ABI version (need major 32)";
    if (g->gridType != GRIDTYPE_FLOAT) return "grid value type is not float";
    if (g->gridSize > bytes) return "grid size field exceeds the buffer";
    const TreeData* t = tree_of(g);
    const int64_t limit = int64_t(g->gridSize) - int64_t(sizeof(GridData));
    for (int l = 0; l < 4; ++l)
        if (t->nodeOffset[l] < 0 || t->nodeOffset[l] > limit) return "tree node offset out of range";
    if (uint64_t(t->nodeOffset[0]) + uint64_t(t->nodeCount[0]) * sizeof(LeafData) > uint64_t(limit)) return "leaf array out of range";
    if (uint64_t(t->nodeOffset[1]) + uint64_t(t->nodeCount[1]) * sizeof(LowerData) > uint64_t(limit)) return "lower-node array out of range";
    if (uint64_t(t->nodeOffset[2]) + uint64_t(t->nodeCount[2]) * sizeof(UpperData) > uint64_t(limit)) return "upper-node array out of range";
    return nullptr;
}

} // namespace vpt_host
