// oracle/substrate_own.hpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product.
//
// Self-contained restatement of the vendored NanoVDB 32.8 arithmetic the reference's hot path rests
// on, operating directly on a serialized grid buffer (include/nvdb_format.h), so the oracle can be
// built where /root/reference is absent. Class and method names follow NanoVDB so that the shared
// restatement (vpt_restated.hpp) compiles unchanged against the REAL headers too (substrate_nanovdb.hpp,
// the `oracle/_ref` build), which is how this file is pinned: tests/test_oracle_ref.py compares both
// builds on identical inputs, bit for bit.
//
// Restated from (NV/ = external/openvdb/nanovdb/nanovdb/):
//   Vec3/Coord ops        NV/math/Math.h:183-186,260-298,630-690
//   matMult fmaf chains   NV/math/Math.h:905-939        Map apply*  NV/NanoVDB.h:1439-1497
//   Ray                   NV/math/Ray.h:64-74,154-159,258-268,398-424,507-515
//   HDDA                  NV/math/HDDA.h:42-160
//   ReadAccessor ops      NV/NanoVDB.h:5330-5397, GetValue/GetLeaf/ProbeValue :6105-6205
//   getDimAndCache        NV/NanoVDB.h:3071-3083,3564-3577,4534-4542
//   trilinear sampler     NV/math/SampleFromVoxels.h:62-69,202-242,370-407
//   pcg32_fast            PCG/pcg_random.hpp:1535,1673 (mcg_xsh_rs_64_32)
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>

#include "nvdb_format.h"

namespace vpto_own {

struct Coord {
    int32_t v[3];
    Coord() : v{0, 0, 0} {}
    Coord(int32_t i, int32_t j, int32_t k) : v{i, j, k} {}
    int32_t& operator[](int i) { return v[i]; }
    const int32_t& operator[](int i) const { return v[i]; }
    Coord operator&(int32_t m) const { return Coord(v[0] & m, v[1] & m, v[2] & m); }
    bool operator==(const Coord& o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2]; }
    bool operator!=(const Coord& o) const { return !(*this == o); }
    static Coord max() { return Coord(INT32_MAX, INT32_MAX, INT32_MAX); }
};

inline int32_t Floor(float x) { return int32_t(floorf(x)); }

struct Vec3f {
    float v[3];
    Vec3f() : v{0, 0, 0} {}
    Vec3f(float a, float b, float c) : v{a, b, c} {}
    float& operator[](int i) { return v[i]; }
    const float& operator[](int i) const { return v[i]; }
    float lengthSqr() const { return v[0] * v[0] + v[1] * v[1] + v[2] * v[2]; }
    float length() const { return sqrtf(lengthSqr()); }
    Vec3f operator*(float s) const { return Vec3f(s * v[0], s * v[1], s * v[2]); }
    Coord floor() const { return Coord(Floor(v[0]), Floor(v[1]), Floor(v[2])); }
};

struct CoordBBox {
    Coord mn, mx;
    const Coord& min() const { return mn; }
    const Coord& max() const { return mx; }
};

inline Vec3f matMult(const float* m, const Vec3f& p)
{
    return Vec3f(fmaf(p[0], m[0], fmaf(p[1], m[1], p[2] * m[2])), fmaf(p[0], m[3], fmaf(p[1], m[4], p[2] * m[5])),
                 fmaf(p[0], m[6], fmaf(p[1], m[7], p[2] * m[8])));
}
inline Vec3f matMult(const float* m, const float* t, const Vec3f& p)
{
    return Vec3f(fmaf(p[0], m[0], fmaf(p[1], m[1], fmaf(p[2], m[2], t[0]))), fmaf(p[0], m[3], fmaf(p[1], m[4], fmaf(p[2], m[5], t[1]))),
                 fmaf(p[0], m[6], fmaf(p[1], m[7], fmaf(p[2], m[8], t[2]))));
}

using Leaf = nvdbf::LeafData;

class Accessor;

// View of a serialized float grid.
class Grid {
public:
    explicit Grid(void* data) : g(static_cast<nvdbf::GridData*>(data)) {}
    Vec3f worldToIndexF(const Vec3f& w) const
    {
        const auto& m = g->map;
        return matMult(m.invMatF, Vec3f(w[0] - m.vecF[0], w[1] - m.vecF[1], w[2] - m.vecF[2]));
    }
    Vec3f worldToIndexDirF(const Vec3f& d) const { return matMult(g->map.invMatF, d); }
    Vec3f indexToWorldF(const Vec3f& p) const { return matMult(g->map.matF, g->map.vecF, p); }
    CoordBBox indexBBox() const
    {
        const nvdbf::RootData* r = nvdbf::root_of(g);
        CoordBBox b;
        b.mn = Coord(r->bbox[0], r->bbox[1], r->bbox[2]);
        b.mx = Coord(r->bbox[3], r->bbox[4], r->bbox[5]);
        return b;
    }
    inline Accessor getAccessor() const;
    uint32_t leafCount() const { return nvdbf::tree_of(g)->nodeCount[0]; }
    Leaf* firstLeaf() const
    {
        const nvdbf::TreeData* t = nvdbf::tree_of(g);
        return reinterpret_cast<Leaf*>(const_cast<char*>(reinterpret_cast<const char*>(t)) + t->nodeOffset[0]);
    }
    nvdbf::GridData* g;
};

// ReadAccessor<float, 0, 1, 2>: caches the last leaf / lower / upper node. Values returned do not
// depend on the cache; getDim is evaluated from the root (see DESIGN.md "getDim and the accessor cache").
class Accessor {
public:
    explicit Accessor(const nvdbf::GridData* g) : root(nvdbf::root_of(g)) { clear(); }
    void clear() { key0 = key1 = key2 = Coord::max(); leaf = nullptr; lower = nullptr; upper = nullptr; }

    float getValue(const Coord& c) const
    {
        if (const Leaf* l = descend(c)) return l->values[nvdbf::leaf_offset(c[0], c[1], c[2])];
        return tileValue;
    }
    const Leaf* probeLeaf(const Coord& c) const { return descend(c); }
    bool probeValue(const Coord& c, float& v) const
    {
        if (const Leaf* l = descend(c)) {
            const uint32_t n = nvdbf::leaf_offset(c[0], c[1], c[2]);
            v = l->values[n];
            return nvdbf::mask_on(l->valueMask, n);
        }
        v = tileValue;
        return tileActive;
    }
    template <class RayT>
    uint32_t getDim(const Coord& c, const RayT&) const
    {
        const nvdbf::RootTile* t = probeTile(c);
        if (!t) return 4096u;          // background: ChildNodeType::dim()
        if (t->child == 0) return 4096u; // root tile: 1 << ChildT::TOTAL
        const auto* up = reinterpret_cast<const nvdbf::UpperData*>(reinterpret_cast<const char*>(root) + t->child);
        if (up->flags & 1u) return 4096u;
        const uint32_t n = nvdbf::upper_offset(c[0], c[1], c[2]);
        if (!nvdbf::mask_on(up->childMask, n)) return 128u;
        const auto* lw = reinterpret_cast<const nvdbf::LowerData*>(reinterpret_cast<const char*>(up) + up->table[n].child);
        if (lw->flags & 1u) return 128u;
        const uint32_t m = nvdbf::lower_offset(c[0], c[1], c[2]);
        if (!nvdbf::mask_on(lw->childMask, m)) return 8u;
        const auto* lf = reinterpret_cast<const Leaf*>(reinterpret_cast<const char*>(lw) + lw->table[m].child);
        return (lf->flags & 1u) ? 8u : 1u;
    }

private:
    const nvdbf::RootTile* probeTile(const Coord& c) const
    {
        const uint64_t key = nvdbf::root_key(c[0], c[1], c[2]);
        const nvdbf::RootTile* t = nvdbf::root_tiles(root);
        for (uint32_t i = 0; i < root->tableSize; ++i) if (t[i].key == key) return &t[i];
        return nullptr;
    }
    // Returns the leaf containing c, or nullptr after setting tileValue/tileActive to what terminates the descent.
    const Leaf* descend(const Coord& c) const
    {
        if ((c & ~7) == key0) return leaf;
        const nvdbf::LowerData* lw = nullptr;
        if ((c & ~127) == key1) {
            lw = lower;
        } else {
            const nvdbf::UpperData* up = nullptr;
            if ((c & ~4095) == key2) {
                up = upper;
            } else {
                const nvdbf::RootTile* t = probeTile(c);
                if (!t) { tileValue = root->background; tileActive = false; return nullptr; }
                if (t->child == 0) { tileValue = t->value; tileActive = t->state > 0u; return nullptr; }
                up = reinterpret_cast<const nvdbf::UpperData*>(reinterpret_cast<const char*>(root) + t->child);
                upper = up; key2 = c & ~4095;
            }
            const uint32_t n = nvdbf::upper_offset(c[0], c[1], c[2]);
            if (!nvdbf::mask_on(up->childMask, n)) { tileValue = up->table[n].value; tileActive = nvdbf::mask_on(up->valueMask, n); return nullptr; }
            lw = reinterpret_cast<const nvdbf::LowerData*>(reinterpret_cast<const char*>(up) + up->table[n].child);
            lower = lw; key1 = c & ~127;
        }
        const uint32_t m = nvdbf::lower_offset(c[0], c[1], c[2]);
        if (!nvdbf::mask_on(lw->childMask, m)) { tileValue = lw->table[m].value; tileActive = nvdbf::mask_on(lw->valueMask, m); return nullptr; }
        leaf = reinterpret_cast<const Leaf*>(reinterpret_cast<const char*>(lw) + lw->table[m].child);
        key0 = c & ~7;
        return leaf;
    }
    const nvdbf::RootData* root;
    mutable Coord key0, key1, key2;
    mutable const Leaf* leaf;
    mutable const nvdbf::LowerData* lower;
    mutable const nvdbf::UpperData* upper;
    mutable float tileValue = 0.0f;
    mutable bool  tileActive = false;
};
inline Accessor Grid::getAccessor() const { return Accessor(g); }

class Ray {
public:
    Ray(const Vec3f& eye, const Vec3f& dir, float t0 = 1e-5f /*Delta<float>*/, float t1 = FLT_MAX /*host Maximum<float>*/)
        : mEye(eye), mDir(dir), mInv(1 / dir[0], 1 / dir[1], 1 / dir[2]), mT0(t0), mT1(t1) {}
    Ray worldToIndexF(const Grid& grid) const
    {
        const Vec3f eye = grid.worldToIndexF(mEye);
        const Vec3f dir = grid.worldToIndexDirF(mDir);
        const float length = dir.length(), invLength = 1.0f / length;
        float t1 = mT1;
        if (mT1 < FLT_MAX) t1 *= length;
        return Ray(eye, dir * invLength, length * mT0, t1);
    }
    bool clip(const CoordBBox& b)
    {
        float t0 = mT0, t1 = mT1;
        for (int i = 0; i < 3; ++i) {
            float a = float(b.min()[i]), c = float(b.max()[i] + 1);
            if (a >= c) return false;
            a = (a - mEye[i]) * mInv[i];
            c = (c - mEye[i]) * mInv[i];
            if (a > c) { const float tmp = a; a = c; c = tmp; }
            if (a > t0) t0 = a;
            if (c < t1) t1 = c;
            if (t0 > t1) return false;
        }
        mT0 = t0; mT1 = t1;
        return true;
    }
    Vec3f operator()(float t) const { return Vec3f(fmaf(t, mDir[0], mEye[0]), fmaf(t, mDir[1], mEye[1]), fmaf(t, mDir[2], mEye[2])); }
    Vec3f start() const { return (*this)(mT0); }
    const Vec3f& eye() const { return mEye; }
    const Vec3f& dir() const { return mDir; }
    const Vec3f& invDir() const { return mInv; }
    float t0() const { return mT0; }
    float t1() const { return mT1; }
private:
    Vec3f mEye, mDir, mInv;
    float mT0, mT1;
};

inline int MinIndex(const Vec3f& v)
{
    if (v[0] < v[1] && v[0] < v[2]) return 0;
    if (v[1] < v[2]) return 1;
    return 2;
}

class HDDA {
public:
    HDDA(const Ray& ray, int dim) { init(ray, ray.t0(), ray.t1(), dim); }
    void init(const Ray& ray, float startTime, float maxTime, int dim)
    {
        mDim = dim; mT0 = startTime; mT1 = maxTime;
        const Vec3f pos = ray(mT0), dir = ray.dir(), inv = ray.invDir();
        mVoxel = pos.floor() & (~(dim - 1));
        for (int a = 0; a < 3; ++a) {
            if (dir[a] == 0.0f) { mNext[a] = FLT_MAX; mStep[a] = 0; }
            else if (inv[a] > 0) { mStep[a] = 1; mNext[a] = mT0 + (mVoxel[a] + dim - pos[a]) * inv[a]; mDelta[a] = inv[a]; }
            else { mStep[a] = -1; mNext[a] = mT0 + (mVoxel[a] - pos[a]) * inv[a]; mDelta[a] = -inv[a]; }
        }
    }
    bool update(const Ray& ray, int dim)
    {
        if (mDim == dim) return false;
        mDim = dim;
        const Vec3f pos = ray(mT0), inv = ray.invDir();
        mVoxel = pos.floor() & (~(dim - 1));
        for (int a = 0; a < 3; ++a) {
            if (mStep[a] == 0) continue;
            mNext[a] = mT0 + (mVoxel[a] - pos[a]) * inv[a];
            if (mStep[a] > 0) mNext[a] += dim * inv[a];
        }
        return true;
    }
    bool step()
    {
        const int a = MinIndex(mNext);
        if (mNext[a] <= mT0) mNext[a] += mT0 - 0.999999f * mNext[a] + 1.0e-6f; // ENFORCE_FORWARD_STEPPING
        mT0 = mNext[a];
        mNext[a] += mDim * mDelta[a];
        mVoxel[a] += mDim * mStep[a];
        return mT0 <= mT1;
    }
    int dim() const { return mDim; }
    const Coord& voxel() const { return mVoxel; }
    float time() const { return mT0; }
    float maxTime() const { return mT1; }
private:
    int32_t mDim;
    float mT0, mT1;
    Coord mVoxel, mStep;
    Vec3f mDelta, mNext;
};

// SampleFromVoxels<Accessor, 1, true>
class TrilinearSampler {
public:
    explicit TrilinearSampler(const Accessor& a) : acc(&a), pos(Coord::max()) {}
    float operator()(Vec3f xyz) const
    {
        const float f[3] = {floorf(xyz[0]), floorf(xyz[1]), floorf(xyz[2])};
        xyz[0] -= f[0]; xyz[1] -= f[1]; xyz[2] -= f[2];
        Coord ijk{int32_t(f[0]), int32_t(f[1]), int32_t(f[2])};
        if (ijk != pos) {
            pos = ijk;
            val[0][0][0] = acc->getValue(ijk); ijk[2] += 1;
            val[0][0][1] = acc->getValue(ijk); ijk[1] += 1;
            val[0][1][1] = acc->getValue(ijk); ijk[2] -= 1;
            val[0][1][0] = acc->getValue(ijk); ijk[0] += 1; ijk[1] -= 1;
            val[1][0][0] = acc->getValue(ijk); ijk[2] += 1;
            val[1][0][1] = acc->getValue(ijk); ijk[1] += 1;
            val[1][1][1] = acc->getValue(ijk); ijk[2] -= 1;
            val[1][1][0] = acc->getValue(ijk);
        }
        auto lerp = [](float a, float b, float w) { return a + w * (b - a); };
        return lerp(lerp(lerp(val[0][0][0], val[0][0][1], xyz[2]), lerp(val[0][1][0], val[0][1][1], xyz[2]), xyz[1]),
                    lerp(lerp(val[1][0][0], val[1][0][1], xyz[2]), lerp(val[1][1][0], val[1][1][1], xyz[2]), xyz[1]), xyz[0]);
    }
private:
    const Accessor* acc;
    mutable Coord pos;
    mutable float val[2][2][2];
};

// pcg32_fast = mcg_xsh_rs_64_32: state *= 6364136223846793005; output from the OLD state.
struct Pcg32Fast {
    uint64_t state = 0xcafef00dd15ea5e5ULL | 3u;
    void seed(uint64_t s) { state = s | 3u; }
    uint32_t operator()()
    {
        const uint64_t old = state;
        state = old * 6364136223846793005ULL;
        const unsigned rshift = unsigned(old >> 61);
        const uint64_t x = old ^ (old >> 22);
        return uint32_t(x >> (22 + rshift));
    }
};

// Trait bundle consumed by vpt_restated.hpp.
struct Substrate {
    using Vec3f = vpto_own::Vec3f;
    using Coord = vpto_own::Coord;
    using Grid = vpto_own::Grid;
    using Accessor = vpto_own::Accessor;
    using Ray = vpto_own::Ray;
    using HDDA = vpto_own::HDDA;
    using Sampler = vpto_own::TrilinearSampler;
    using Engine = vpto_own::Pcg32Fast;
    static const char* name() { return "own"; }
    static bool probe_leaf_max(const Accessor& acc, const Coord& ijk, float& v)
    {
        if (const Leaf* leaf = acc.probeLeaf(ijk)) { v = leaf->maximum; return true; }
        return false;
    }
    // src/volume.cpp:104-160: leaf maximum := max(leaf.getMax(), getValue over the 1-voxel shell around the
    // leaf), for every leaf; values are read through getValue (inactive voxels and tiles included).
    static void fix_majorants_for_interpolation(Grid& density, unsigned order)
    {
        Leaf* leaf = density.firstLeaf();
        const uint32_t n = density.leafCount();
        Accessor acc = density.getAccessor();
        const int o = int(order);
        for (uint32_t l = 0; l < n; ++l, ++leaf) {
            float majorant = leaf->maximum;
            const int32_t org[3] = {leaf->bboxMin[0] & ~7, leaf->bboxMin[1] & ~7, leaf->bboxMin[2] & ~7};
            for (int i = -1; i <= 1; ++i) for (int j = -1; j <= 1; ++j) for (int k = -1; k <= 1; ++k) {
                if (i == 0 && j == 0 && k == 0) continue;
                const int d[3] = {i, j, k};
                int32_t lo[3], hi[3];
                for (int a = 0; a < 3; ++a) {
                    const int32_t nlo = org[a] + 8 * d[a], nhi = nlo + 7; // neighbour leaf-sized bbox
                    lo[a] = std::max(nlo, org[a] - o);                    // intersected with the stencil's reach
                    hi[a] = std::min(nhi, org[a] + 7 + o);
                }
                for (int32_t x = lo[0]; x <= hi[0]; ++x) for (int32_t y = lo[1]; y <= hi[1]; ++y) for (int32_t z = lo[2]; z <= hi[2]; ++z)
                    majorant = std::max(majorant, acc.getValue(Coord(x, y, z)));
            }
            leaf->maximum = majorant;
        }
    }
};

} // namespace vpto_own
