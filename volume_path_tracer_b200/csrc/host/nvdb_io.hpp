// nvdb_io.hpp — `.nvdb` file access without the NanoVDB headers (see nvdb_io.cpp).
#pragma once
#include <cstddef>
#include <string>
#include <vector>

namespace vpt_host {
// Both return "" on success, else the error message. *out is 64-byte aligned, release with free().
std::string read_nvdb_grid(const std::string& path, const std::string& grid_name, void** out, size_t* out_bytes);
std::string write_nvdb_grids(const std::string& path, const void* const* grids, size_t n);
} // namespace vpt_host
