// dda_trace.hpp — host-side Volume::log_dda_trace (see dda_trace.cpp).
#pragma once
#include <cstddef>
#include <string>

namespace vpt_host {
// Writes dda_trace.csv for the world-space ray (origin, direction) through the serialized NanoVDB float grid. Returns "" on success (also when
// the ray misses the grid: no file is written then, like the reference), else the error text. *rows_out = voxel rows written.
std::string write_dda_trace(const void* grid, size_t bytes, const float origin[3], const float direction[3], const std::string& csv_path, size_t* rows_out);
} // namespace vpt_host
