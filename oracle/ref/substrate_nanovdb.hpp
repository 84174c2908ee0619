// oracle/ref/substrate_nanovdb.hpp — TEST INFRASTRUCTURE (CPU oracle, `oracle/_ref` build only).
//
// Substrate for vpt_restated.hpp that is the reference's OWN vendored arithmetic: the NanoVDB 32.8 and
// pcg-cpp headers are #included unchanged from where they lie under the reference tree
// (-I$REF/external/openvdb/nanovdb -I$REF/external/pcg-cpp/include; nothing is copied into this
// repository). Built only where /root/reference exists (oracle/Makefile target `ref`), output
// oracle/_ref/libvpt_oracle_ref.so. Used to pin substrate_own.hpp and as the "reference" CPU baseline.
#pragma once
#include <nanovdb/NanoVDB.h>
#include <nanovdb/math/HDDA.h>
#include <nanovdb/math/Ray.h>
#include <nanovdb/math/SampleFromVoxels.h>
#include <pcg/pcg_random.hpp>

#include "nvdb_format.h"

namespace vpto_nano {

using GridT = nanovdb::NanoGrid<float>;
using AccT = GridT::AccessorType;

// ---- the layout header must describe exactly what the real headers define ----
static_assert(sizeof(nvdbf::GridData) == sizeof(nanovdb::GridData), "GridData size");
static_assert(sizeof(nvdbf::Map) == sizeof(nanovdb::Map), "Map size");
static_assert(sizeof(nvdbf::TreeData) == sizeof(nanovdb::TreeData), "TreeData size");
static_assert(sizeof(nvdbf::RootData) == sizeof(GridT::TreeType::RootType::DataType), "RootData size");
static_assert(sizeof(nvdbf::RootTile) == sizeof(GridT::TreeType::RootType::DataType::Tile), "RootTile size");
static_assert(sizeof(nvdbf::UpperData) == sizeof(nanovdb::NanoUpper<float>::DataType), "Upper size");
static_assert(sizeof(nvdbf::LowerData) == sizeof(nanovdb::NanoLower<float>::DataType), "Lower size");
static_assert(sizeof(nvdbf::LeafData) == sizeof(nanovdb::NanoLeaf<float>::DataType), "Leaf size");
static_assert(sizeof(nvdbf::FileHeader) == sizeof(nanovdb::io::FileHeader), "FileHeader size");
static_assert(sizeof(nvdbf::FileMetaData) == sizeof(nanovdb::io::FileMetaData), "FileMetaData size");
static_assert(nvdbf::ABI_MAJOR == NANOVDB_MAJOR_VERSION_NUMBER, "ABI major");

// Thin by-value handle with the member names vpt_restated.hpp uses.
class Grid {
public:
    explicit Grid(void* data) : g(reinterpret_cast<GridT*>(data)) {}
    nanovdb::Vec3f worldToIndexF(const nanovdb::Vec3f& w) const { return g->worldToIndexF(w); }
    nanovdb::Vec3f worldToIndexDirF(const nanovdb::Vec3f& d) const { return g->worldToIndexDirF(d); }
    nanovdb::Vec3f indexToWorldF(const nanovdb::Vec3f& p) const { return g->indexToWorldF(p); }
    auto indexBBox() const { return g->indexBBox(); }
    AccT getAccessor() const { return g->getAccessor(); }
    // so that nanovdb::math::Ray::worldToIndexF(grid) accepts this handle
    GridT* g;
};

class Ray : public nanovdb::math::Ray<float> {
public:
    using Base = nanovdb::math::Ray<float>;
    Ray(const nanovdb::Vec3f& eye, const nanovdb::Vec3f& dir) : Base(eye, dir) {}
    Ray(const Base& b) : Base(b) {}
    Ray worldToIndexF(const Grid& grid) const { return Ray(Base::worldToIndexF(*grid.g)); }
};

struct Substrate {
    using Vec3f = nanovdb::Vec3f;
    using Coord = nanovdb::Coord;
    using Grid = vpto_nano::Grid;
    using Accessor = AccT;
    using Ray = vpto_nano::Ray;
    using HDDA = nanovdb::math::HDDA<nanovdb::math::Ray<float>>;
    using Sampler = nanovdb::math::SampleFromVoxels<AccT, 1>;
    using Engine = pcg32_fast;
    static const char* name() { return "nanovdb"; }
    static bool probe_leaf_max(const Accessor& acc, const Coord& ijk, float& v)
    {
        if (const auto* leaf = acc.probeLeaf(ijk)) { v = leaf->getMax(); return true; }
        return false;
    }
    // src/volume.cpp:104-160, through the same NanoVDB calls the reference makes.
    static void fix_majorants_for_interpolation(Grid& dens, unsigned order)
    {
        GridT& density = *dens.g;
        auto* leaf_begin = density.tree().getFirstLeaf();
        auto* leaf_end = leaf_begin + density.tree().nodeCount<GridT::LeafNodeType>();
        auto acc = density.getAccessor();
        for (auto* lp = leaf_begin; lp != leaf_end; ++lp) {
            auto& leaf = *lp;
            float majorant_density = leaf.getMax();
            constexpr auto LEAF_DIM = GridT::LeafNodeType::DIM;
            nanovdb::math::BBox<nanovdb::math::Coord> leaf_bbox(leaf.origin(), leaf.origin().offsetBy(LEAF_DIM - 1));
            auto aoe_bbox = leaf_bbox.expandBy(order);
            for (int i = -1; i <= 1; ++i) for (int j = -1; j <= 1; ++j) for (int k = -1; k <= 1; ++k) {
                if (i == 0 && j == 0 && k == 0) continue;
                auto neighbour_bbox = leaf_bbox;
                neighbour_bbox.translate(nanovdb::Coord(i * int(LEAF_DIM), j * int(LEAF_DIM), k * int(LEAF_DIM)));
                neighbour_bbox.intersect(aoe_bbox);
                for (const nanovdb::math::Coord& c : neighbour_bbox) majorant_density = std::max(majorant_density, acc.getValue(c));
            }
            leaf.setMax(majorant_density);
        }
    }
};

} // namespace vpto_nano
