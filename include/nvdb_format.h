/* nvdb_format.h — byte layout of a serialized NanoVDB float grid (ABI major 32).
 *
 * This is a DATA-FORMAT description, not an algorithm: plain structs whose sizes and
 * member offsets restate the published NanoVDB layout, so that this repository can read
 * and write `.nvdb` grids without the NanoVDB headers being present (they are absent on
 * the GPU box).  Every struct cites the upstream definition it mirrors; the `oracle/_ref`
 * build static_asserts each size/offset below against the real vendored headers.
 *
 * Upstream (read for behaviour only, `NV/` = external/openvdb/nanovdb/nanovdb/):
 *   GridData      NV/NanoVDB.h:1894-1920   (672 B)
 *   Map           NV/NanoVDB.h:1375-1385   (264 B)
 *   TreeData      NV/NanoVDB.h:2342-2348   (64 B)
 *   RootData/Tile NV/NanoVDB.h:2566-2640   (64 B header, 32 B tiles, 64-bit packed key)
 *   InternalData  NV/NanoVDB.h:3127-3165   (upper 32^3: 270400 B, lower 16^3: 33856 B)
 *   LeafData      NV/NanoVDB.h:3626-3636   (2144 B; voxel n = (x&7)<<6 | (y&7)<<3 | (z&7))
 *   FileHeader / FileMetaData NV/NanoVDB.h:5817-5856
 */
#ifndef VPT_NVDB_FORMAT_H
#define VPT_NVDB_FORMAT_H

#include <cstddef>
#include <cstdint>

namespace nvdbf {

/* the bit helpers below are also used by the device-side baker (csrc/device/vpt_bake.cu) */
#if defined(__CUDACC__)
#define NVDBF_HD __host__ __device__
#else
#define NVDBF_HD
#endif
constexpr uint64_t MAGIC_NUMB = 0x304244566f6e614eULL; /* "NanoVDB0" */
constexpr uint64_t MAGIC_GRID = 0x314244566f6e614eULL; /* "NanoVDB1" */
constexpr uint64_t MAGIC_FILE = 0x324244566f6e614eULL; /* "NanoVDB2" */
constexpr uint32_t ABI_MAJOR = 32, ABI_MINOR = 8, ABI_PATCH = 0;
/* Version packs major<<21 | minor<<10 | patch (NV/NanoVDB.h Version class). */
constexpr uint32_t pack_version(uint32_t ma, uint32_t mi, uint32_t pa) { return (ma << 21) | (mi << 10) | pa; }
constexpr uint32_t version_major(uint32_t v) { return v >> 21; }

constexpr uint32_t GRIDTYPE_FLOAT = 1;      /* GridType::Float */
constexpr uint32_t GRIDCLASS_UNKNOWN = 0;   /* GridClass::Unknown */
constexpr uint32_t GRIDCLASS_FOG = 2;       /* GridClass::FogVolume */
/* GridFlags bits (NV/NanoVDB.h GridFlags) */
constexpr uint32_t GF_HAS_LONG_NAME = 1u << 0, GF_HAS_BBOX = 1u << 1, GF_HAS_MINMAX = 1u << 2,
                   GF_HAS_AVERAGE = 1u << 3, GF_HAS_STDDEV = 1u << 4, GF_BREADTH_FIRST = 1u << 5;

struct Map { /* 264 B */
    float  matF[9], invMatF[9], vecF[3], taperF;
    double matD[9], invMatD[9], vecD[3], taperD;
};
static_assert(sizeof(Map) == 264, "Map");

struct GridData { /* 672 B */
    uint64_t magic;            /*   0 */
    uint64_t checksum;         /*   8  all ones = disabled */
    uint32_t version;          /*  16 */
    uint32_t flags;            /*  20 */
    uint32_t gridIndex;        /*  24 */
    uint32_t gridCount;        /*  28 */
    uint64_t gridSize;         /*  32 */
    char     gridName[256];    /*  40 */
    Map      map;              /* 296 */
    double   worldBBox[6];     /* 560  min xyz, max xyz */
    double   voxelSize[3];     /* 608 */
    uint32_t gridClass;        /* 632 */
    uint32_t gridType;         /* 636 */
    int64_t  blindMetadataOffset; /* 640 */
    uint32_t blindMetadataCount;  /* 648 */
    uint32_t data0;            /* 652 */
    uint64_t data1;            /* 656 */
    uint64_t data2;            /* 664 */
};
static_assert(sizeof(GridData) == 672 && offsetof(GridData, map) == 296 &&
              offsetof(GridData, worldBBox) == 560 && offsetof(GridData, gridClass) == 632, "GridData");

struct TreeData { /* 64 B, immediately after GridData */
    int64_t  nodeOffset[4];  /* bytes from TreeData to first leaf, lower, upper, root */
    uint32_t nodeCount[3];   /* leaf, lower, upper */
    uint32_t tileCount[3];   /* active tiles at lower, upper, root level */
    uint64_t voxelCount;     /* active voxels */
};
static_assert(sizeof(TreeData) == 64, "TreeData");

struct RootData { /* 64 B, followed by tableSize RootTile */
    int32_t  bbox[6];        /* index AABB of active values: min xyz, max xyz (inclusive) */
    uint32_t tableSize;
    float    background, minimum, maximum, average, stdDevi;
    uint8_t  pad[16];
};
static_assert(sizeof(RootData) == 64 && offsetof(RootData, tableSize) == 24, "RootData");

struct RootTile { /* 32 B */
    uint64_t key;    /* (x>>12)<<42 | (y>>12)<<21 | (z>>12), each 21 bits of the uint32 coord */
    int64_t  child;  /* byte offset RootData -> upper node; 0 = constant tile */
    uint32_t state;  /* tile active? */
    float    value;
    uint8_t  pad[8];
};
static_assert(sizeof(RootTile) == 32, "RootTile");

NVDBF_HD inline uint64_t root_key(int32_t x, int32_t y, int32_t z)
{
    return (uint64_t(uint32_t(z)) >> 12) | ((uint64_t(uint32_t(y)) >> 12) << 21) | ((uint64_t(uint32_t(x)) >> 12) << 42);
}
NVDBF_HD inline void root_key_origin(uint64_t key, int32_t o[3])
{
    const uint64_t M = (1u << 21) - 1;
    o[0] = int32_t(uint32_t(((key >> 42) & M) << 12));
    o[1] = int32_t(uint32_t(((key >> 21) & M) << 12));
    o[2] = int32_t(uint32_t((key & M) << 12));
}

union InternalTile { float value; int64_t child; }; /* child: byte offset from the node itself */

template <int LOG2DIM>
struct InternalData {
    static constexpr int SIZE = 1 << (3 * LOG2DIM);
    int32_t  bbox[6];
    uint64_t flags;                 /* bit0: "skip/disable rendering", bit1: has active values */
    uint64_t valueMask[SIZE / 64];
    uint64_t childMask[SIZE / 64];
    float    minimum, maximum, average, stdDevi;
    alignas(32) InternalTile table[SIZE];
};
using UpperData = InternalData<5>; /* 32^3 children of 128^3 voxels, spans 4096^3 */
using LowerData = InternalData<4>; /* 16^3 children of   8^3 voxels, spans  128^3 */
static_assert(sizeof(UpperData) == 270400 && offsetof(UpperData, table) == 8256, "Upper");
static_assert(sizeof(LowerData) == 33856 && offsetof(LowerData, table) == 1088, "Lower");

struct LeafData { /* 2144 B */
    int32_t  bboxMin[3];
    uint8_t  bboxDif[3];
    uint8_t  flags;         /* bit0 skip render, bit1 has bbox, bit4 has stats */
    uint64_t valueMask[8];
    float    minimum, maximum, average, stdDevi;
    alignas(32) float values[512];
};
static_assert(sizeof(LeafData) == 2144 && offsetof(LeafData, values) == 96 &&
              offsetof(LeafData, maximum) == 84, "Leaf");

NVDBF_HD inline uint32_t upper_offset(int32_t x, int32_t y, int32_t z)
{
    return (uint32_t((x & 4095) >> 7) << 10) | (uint32_t((y & 4095) >> 7) << 5) | uint32_t((z & 4095) >> 7);
}
NVDBF_HD inline uint32_t lower_offset(int32_t x, int32_t y, int32_t z)
{
    return (uint32_t((x & 127) >> 3) << 8) | (uint32_t((y & 127) >> 3) << 4) | uint32_t((z & 127) >> 3);
}
NVDBF_HD inline uint32_t leaf_offset(int32_t x, int32_t y, int32_t z)
{
    return (uint32_t(x & 7) << 6) | (uint32_t(y & 7) << 3) | uint32_t(z & 7);
}
NVDBF_HD inline bool mask_on(const uint64_t* words, uint32_t n) { return (words[n >> 6] >> (n & 63)) & 1u; }
NVDBF_HD inline void mask_set(uint64_t* words, uint32_t n) { words[n >> 6] |= uint64_t(1) << (n & 63); }

/* ---- .nvdb file segments (uncompressed or ZIP codec) ---- */
struct FileHeader { uint64_t magic; uint32_t version; uint16_t gridCount; uint16_t codec; };
static_assert(sizeof(FileHeader) == 16, "FileHeader");
struct FileMetaData { /* 176 B, followed by nameSize chars (incl. NUL) */
    uint64_t gridSize, fileSize, nameKey, voxelCount;
    uint32_t gridType, gridClass;
    double   worldBBox[6];
    int32_t  indexBBox[6];
    double   voxelSize[3];
    uint32_t nameSize;
    uint32_t nodeCount[4];
    uint32_t tileCount[3];
    uint16_t codec, padding;
    uint32_t version;
};
static_assert(sizeof(FileMetaData) == 176, "FileMetaData");

NVDBF_HD inline const TreeData* tree_of(const GridData* g) { return reinterpret_cast<const TreeData*>(g + 1); }
NVDBF_HD inline const RootData* root_of(const GridData* g)
{
    const TreeData* t = tree_of(g);
    return reinterpret_cast<const RootData*>(reinterpret_cast<const char*>(t) + t->nodeOffset[3]);
}
NVDBF_HD inline const RootTile* root_tiles(const RootData* r) { return reinterpret_cast<const RootTile*>(r + 1); }

} // namespace nvdbf
#endif
