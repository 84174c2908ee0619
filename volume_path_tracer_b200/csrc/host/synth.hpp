// synth.hpp — seeded procedural test/benchmark volumes as serialized NanoVDB float grids (see synth.cpp).
#pragma once
#include <array>
#include <cstdint>
#include <string>
#include <vector>

#include "nvdb_build.hpp"

namespace vpt_host {
struct SynthResult {
    std::vector<uint8_t> density;
    std::vector<uint8_t> temperature; // empty when the kind has none
};
// kind: fire | cloud | dense | torus | tiles. anchor (may be null = origin): world position of the volume's
// anchor point — centre of the box, or centre of the base for "fire". Throws std::runtime_error on bad arguments.
SynthResult synth_volume(const std::string& kind, const int32_t dims[3], double voxel_size, uint32_t seed, const double* anchor);
} // namespace vpt_host
