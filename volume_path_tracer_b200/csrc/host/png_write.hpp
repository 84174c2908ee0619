// png_write.hpp — 8-bit RGB truecolor PNG writer (see png_write.cpp).
#pragma once
#include <cstdint>
#include <string>

namespace vpt_host {
bool write_png_rgb8(const std::string& path, const uint8_t* rgb, int64_t width, int64_t height, std::string& err);
} // namespace vpt_host
