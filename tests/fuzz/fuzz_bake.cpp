// ASan/UBSan fuzz harness for vpt_host::bake_grid (csrc/host/bake.cpp): the C ABI takes raw NanoVDB buffers from the caller, so a
// corrupt or hostile buffer must be REJECTED (std::runtime_error -> VPTC_ERR_BAD_GRID), never followed out of bounds.
// Built and run by tests/test_host.py::test_baker_rejects_corrupt_grids; usage: fuzz_bake <seed> <iterations> <fire|cloud|dense>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <stdexcept>
#include "host/bake.hpp"
#include "host/synth.hpp"
using namespace vpt_host;
int main(int argc, char** argv)
{
    const unsigned seed = argc > 1 ? unsigned(std::atoi(argv[1])) : 1u;
    const int iters = argc > 2 ? std::atoi(argv[2]) : 2000;
    const char* kind = argc > 3 ? argv[3] : "cloud";
    const int32_t dims[3] = {48, 40, 56};
    SynthResult s = synth_volume(kind, dims, 1.0, 1, nullptr);
    std::mt19937_64 rng(seed);
    int ok = 0, thrown = 0;
    const size_t n = s.density.size();
    { // the pristine grid must bake
        void* p = nullptr;
        if (posix_memalign(&p, 64, n)) return 2;
        std::memcpy(p, s.density.data(), n);
        BakedGrid out;
        bake_grid(p, n, true, out);
        if (out.bricks.empty() || out.cells.empty()) { std::printf("pristine grid baked to nothing\n"); return 3; }
        std::free(p);
    }
    { // fixed corpus: finite but SINGULAR / inconsistent index<->world maps must be rejected (a zero inverse map makes every ray direction
      // NaN on the device, whose march then never ends): all-zero invMatF, all-zero matF, a rank-2 map, matF that is not invMatF's inverse
        const size_t MAT = 296, INV = 296 + 36; // nvdbf::GridData::map.matF / .invMatF
        for (int c = 0; c < 4; ++c) {
            void* p = nullptr;
            if (posix_memalign(&p, 64, n)) return 2;
            std::memcpy(p, s.density.data(), n);
            uint8_t* b = static_cast<uint8_t*>(p);
            float m[9];
            if (c == 0) std::memset(b + INV, 0, 36);
            else if (c == 1) std::memset(b + MAT, 0, 36);
            else if (c == 2) { std::memcpy(m, b + INV, 36); m[6] = m[3]; m[7] = m[4]; m[8] = m[5]; std::memcpy(b + INV, m, 36); }
            else { std::memcpy(m, b + MAT, 36); for (float& v : m) v *= 3.0f; std::memcpy(b + MAT, m, 36); }
            bool rejected = false;
            try { BakedGrid out; bake_grid(p, n, true, out); } catch (const std::exception&) { rejected = true; }
            std::free(p);
            if (!rejected) { std::printf("singular-map case %d was baked\n", c); return 4; }
        }
    }
    // regions worth hitting: grid/tree/root headers (first ~1 KB), root tiles, and random places in the node tables
    for (int it = 0; it < iters; ++it) {
        void* p = nullptr;
        if (posix_memalign(&p, 64, n)) return 2;
        std::memcpy(p, s.density.data(), n);
        uint8_t* b = static_cast<uint8_t*>(p);
        const int k = 1 + int(rng() % 4);
        for (int j = 0; j < k; ++j) {
            const double r = double(rng() % 1000) / 1000.0;
            size_t pos = r < 0.5 ? size_t(rng() % 1200) : (r < 0.8 ? size_t(rng() % std::min<size_t>(n, 300000)) : size_t(rng() % n));
            pos = std::min(pos, n - 8);
            const int mode = int(rng() % 4);
            if (mode == 0) b[pos] = uint8_t(rng());
            else if (mode == 1) { const uint32_t v[] = {0u, 1u, 0xffffffffu, 0x7fffffffu, 0x80000000u, uint32_t(n), uint32_t(n * 2)}; const uint32_t x = v[rng() % 7]; std::memcpy(b + (pos & ~size_t(3)), &x, 4); }
            else if (mode == 2) { const uint64_t v[] = {0ull, ~0ull, 1ull << 63, uint64_t(n), uint64_t(n) * 16, 0x7fffffffffffffffull, uint64_t(-64ll)}; const uint64_t x = v[rng() % 7]; std::memcpy(b + (pos & ~size_t(7)), &x, 8); }
            else { uint64_t x; std::memcpy(&x, b + (pos & ~size_t(7)), 8); x ^= 1ull << (rng() % 64); std::memcpy(b + (pos & ~size_t(7)), &x, 8); }
        }
        size_t bytes = n;
        if (rng() % 10 == 0) bytes = size_t(rng() % n);
        try {
            BakedGrid out;
            bake_grid(p, bytes, /*patch_majorants=*/(rng() & 1) != 0, out);
            ++ok;
        } catch (const std::exception&) { ++thrown; }
        std::free(p);
    }
    std::printf("seed %u: baked %d rejected %d\n", seed, ok, thrown);
    return 0;
}
