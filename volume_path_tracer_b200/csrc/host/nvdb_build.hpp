// nvdb_build.hpp — serialize a sparse float volume as a NanoVDB grid buffer (ABI 32) without the
// NanoVDB headers. Used by the synthetic-volume generators so that the product and its tests run
// from `.nvdb`-format grids on machines where the reference tree is absent.
//
// The byte layout follows include/nvdb_format.h. Node statistics follow what
// nanovdb::tools::GridStats writes (NV/tools/GridStats.h:617-735): min/max/avg/stddev over ACTIVE
// values only, index bbox of active values, flag bit0 = "no active value below this node",
// bit1 = "has active values"; world bbox = map(indexBBox.min .. indexBBox.max+1) (GridStats.h:596).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "nvdb_format.h"

namespace vpt_host {

struct BuildLeaf {
    int32_t  origin[3];   // multiple of 8
    uint64_t mask[8];     // active voxels, bit n = (x&7)<<6 | (y&7)<<3 | (z&7)
    float    values[512]; // values of active AND inactive voxels (inactive normally = background)
};

// A constant region stored in an internal-node slot or in the root table.
struct BuildTile {
    int     level;      // 1: 8^3 slot of a lower node, 2: 128^3 slot of an upper node, 3: 4096^3 root tile
    int32_t origin[3];  // aligned to the tile size
    float   value;
    bool    active;
};

struct BuildParams {
    double      voxel_size = 1.0;
    double      translation[3] = {0, 0, 0};
    float       background = 0.0f;
    std::string name = "density";
    uint32_t    grid_class = nvdbf::GRIDCLASS_FOG;
};

// Leaves without any active voxel are kept (NanoVDB keeps them too);
// filter beforehand if unwanted. Throws std::runtime_error on overlapping leaves/tiles.
std::vector<uint8_t> build_grid(std::vector<BuildLeaf>& leaves, const std::vector<BuildTile>& tiles,
                                const BuildParams& params);

// Validates magic/version/type/size of a serialized grid. Returns nullptr when fine, else a message.
const char* check_grid(const void* data, size_t bytes);

} // namespace vpt_host
