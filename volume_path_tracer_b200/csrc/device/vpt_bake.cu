// vpt_bake.cu — the heavy half of the baker on the device: apron bricks and the halo-majorant patch.
//
// Volume::Volume runs fix_majorants_for_interpolation over every leaf (src/volume.cpp:104-170: max over the 1-voxel shell of the leaf, 488
// getValue descents each); the host baker (csrc/host/bake.cpp) does that and fills the 9^3 apron bricks on the host threads, which is 97 % of
// its time (2.5 s of 2.7 s for the 3.58 GB cloud) and is followed by a 5.6 GB upload of the result. Here the caller's serialized grid is
// uploaded as it is (3.58 GB) and one block per leaf reads the 10^3 neighbourhood [o-1, o+8]^3 through the same top-down descent
// (ReadAccessor::getValue, NV/NanoVDB.h GetValue: leaf voxel, else the tile value where the descent ends, else the background), writes the brick
// in the device layout (layout.hpp brick_offset) and reduces the patched maximum; a second pass folds the maxima into the lower-slot table.
// The host still validates every offset of the tree first (bake.cpp validate_tree), builds the cell / slot / tile tables (1 % of the work)
// and decides the brick order. Output is byte-identical to the host baker's (tests/test_gpu_parity.py::test_device_bake_equals_host_bake).
#include <cstdint>

#include <cuda_runtime.h>

#include "layout.hpp"
#include "nvdb_format.h"
#include "vpt_kernels.hpp"

namespace vptd {
using namespace nvdbf;

namespace {
struct TreeDev {
    const RootData* root; const RootTile* tiles; const LeafData* leaf0;
};

__device__ __forceinline__ float tree_value(const TreeDev& t, int32_t x, int32_t y, int32_t z)
{
    const uint64_t key = root_key(x, y, z);
    const RootTile* rt = nullptr;
    for (uint32_t i = 0; i < t.root->tableSize; ++i)
        if (t.tiles[i].key == key) { rt = &t.tiles[i]; break; }
    if (!rt) return t.root->background;
    if (rt->child == 0) return rt->value;
    const UpperData* u = reinterpret_cast<const UpperData*>(reinterpret_cast<const char*>(t.root) + rt->child);
    const uint32_t un = upper_offset(x, y, z);
    if (!mask_on(u->childMask, un)) return u->table[un].value;
    const LowerData* l = reinterpret_cast<const LowerData*>(reinterpret_cast<const char*>(u) + u->table[un].child);
    const uint32_t ln = lower_offset(x, y, z);
    if (!mask_on(l->childMask, ln)) return l->table[ln].value;
    const LeafData* lf = reinterpret_cast<const LeafData*>(reinterpret_cast<const char*>(l) + l->table[ln].child);
    return lf->values[leaf_offset(x, y, z)];
}

// std::max(m, v) as the host baker applies it: (m < v) ? v : m — a NaN never replaces the running maximum
__device__ __forceinline__ float max_like_std(float m, float v) { return (m < v) ? v : m; }

constexpr int BAKE_THREADS = 256;

// One block per leaf. The 10^3 neighbourhood is staged in shared memory (index (i+1, j+1, k+1)), then written out in the device layout.
__global__ void __launch_bounds__(BAKE_THREADS) vpt_bake_bricks_kernel(const char* grid, uint64_t n_leaf, const uint32_t* brick_of, int patch, float* bricks, float* max_by_brick)
{
    __shared__ float nb[1000];
    __shared__ float red[BAKE_THREADS / 32];
    const GridData* g = reinterpret_cast<const GridData*>(grid);
    const TreeData* tree = tree_of(g);
    TreeDev t;
    t.root = root_of(g); t.tiles = root_tiles(t.root);
    t.leaf0 = reinterpret_cast<const LeafData*>(reinterpret_cast<const char*>(tree) + tree->nodeOffset[0]);
    for (uint64_t li = blockIdx.x; li < n_leaf; li += gridDim.x) {
        const LeafData* lf = t.leaf0 + li;
        const int32_t ox = lf->bboxMin[0] & ~7, oy = lf->bboxMin[1] & ~7, oz = lf->bboxMin[2] & ~7;
        float m = -__int_as_float(0x7f800000); // -inf: neutral for the maximum of the shell values
        for (int p = threadIdx.x; p < 1000; p += BAKE_THREADS) {
            const int i = p / 100 - 1, j = (p / 10) % 10 - 1, k = p % 10 - 1;
            const bool core = i >= 0 && i < 8 && j >= 0 && j < 8 && k >= 0 && k < 8;
            // bricks only need [0, 8]^3; the i/j/k = -1 faces are needed by the patch alone
            const bool apron = i >= 0 && j >= 0 && k >= 0;
            float v = 0.0f;
            if (core) v = lf->values[(i << 6) | (j << 3) | k];
            else if (apron || patch) v = tree_value(t, ox + i, oy + j, oz + k);
            nb[p] = v;
            if (!core && patch) m = max_like_std(m, v); // a NaN shell value is dropped, as std::max(m, NaN) keeps m
        }
        // block maximum of the shell (no NaN left: fmaxf is exact)
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
        __syncthreads();
        const uint32_t brick = brick_of ? brick_of[li] : uint32_t(li);
        if (threadIdx.x == 0) {
            float s = red[0];
            for (int w = 1; w < BAKE_THREADS / 32; ++w) s = fmaxf(s, red[w]);
            max_by_brick[brick] = patch ? max_like_std(lf->maximum, s) : lf->maximum; // stats over ACTIVE voxels only (NV/tools/GridStats.h:716-730), then the shell
        }
        float* dev = bricks + size_t(brick) * size_t(BRICK_STRIDE);
#if VPT_BRICK_PAIRS == 2
        for (int q = threadIdx.x; q < 512 * 8; q += BAKE_THREADS) {
            const int c = q & 7, v = q >> 3, i = v >> 6, j = (v >> 3) & 7, k = v & 7;
            dev[brick_offset(i, j, k) + c] = nb[(i + 1 + (c >> 2)) * 100 + (j + 1 + ((c >> 1) & 1)) * 10 + (k + 1 + (c & 1))];
        }
#elif VPT_BRICK_PAIRS
        for (int q = threadIdx.x; q < 81 * 16; q += BAKE_THREADS) {
            const int r = q >> 4, e = q & 15, i = r / 9, j = r % 9, k = e >> 1;
            dev[brick_offset(i, j, k) + (e & 1)] = nb[(i + 1) * 100 + (j + 1) * 10 + (k + 1 + (e & 1))];
        }
#else
        for (int q = threadIdx.x; q < 729; q += BAKE_THREADS) {
            const int i = q / 81, j = (q / 9) % 9, k = q % 9;
            dev[brick_offset(i, j, k)] = nb[(i + 1) * 100 + (j + 1) * 10 + (k + 1)];
        }
#endif
        __syncthreads();
    }
}

// The lower-slot table was uploaded with the leaves' UNPATCHED maxima: fold the patched ones in.
__global__ void vpt_bake_slot_majorants_kernel(Slot* slots, uint64_t n_slots, const float* max_by_brick)
{
    for (uint64_t e = blockIdx.x * uint64_t(blockDim.x) + threadIdx.x; e < n_slots; e += uint64_t(gridDim.x) * blockDim.x) {
        const uint32_t code = slots[e].code;
        if (!(code & SLOT_KIND_TILE)) slots[e].majorant = max_by_brick[code];
    }
}
} // namespace

cudaError_t launch_bake_bricks(const void* d_grid, uint64_t n_leaf, const uint32_t* d_brick_of, bool patch, float* d_bricks, float* d_max_by_brick,
                               void* d_slots, uint64_t n_slots, cudaStream_t s)
{
    if (n_leaf) {
        const unsigned blocks = unsigned(n_leaf < (1u << 20) ? n_leaf : (1u << 20));
        vpt_bake_bricks_kernel<<<blocks, BAKE_THREADS, 0, s>>>(static_cast<const char*>(d_grid), n_leaf, d_brick_of, patch ? 1 : 0, d_bricks, d_max_by_brick);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (n_slots && n_leaf) {
        vpt_bake_slot_majorants_kernel<<<unsigned((n_slots + 255) / 256 < 65535 * 16 ? (n_slots + 255) / 256 : 65535 * 16), 256, 0, s>>>(static_cast<Slot*>(d_slots), n_slots, d_max_by_brick);
        return cudaGetLastError();
    }
    return cudaSuccess;
}

} // namespace vptd
