// png_write.cpp — replaces Image<unsigned char,3>::save (include/vpt/image.hpp:53-56, src/image_io.cpp:18-58),
// which encodes 8-bit RGB truecolor through the vendored libspng with no gamma/ICC chunks. Ours writes the
// same kind of file (IHDR colour type 2, bit depth 8, no ancillary chunks) with zlib's deflate directly.
#include "png_write.hpp"

#include <cstdio>
#include <vector>
#include <zlib.h>

namespace vpt_host {
namespace {
void put32(std::vector<uint8_t>& v, uint32_t x) { v.push_back(uint8_t(x >> 24)); v.push_back(uint8_t(x >> 16)); v.push_back(uint8_t(x >> 8)); v.push_back(uint8_t(x)); }
bool chunk(FILE* f, const char type[4], const std::vector<uint8_t>& data)
{
    std::vector<uint8_t> head; put32(head, uint32_t(data.size()));
    uint32_t crc = uint32_t(crc32(0L, reinterpret_cast<const Bytef*>(type), 4));
    if (!data.empty()) crc = uint32_t(crc32(crc, data.data(), uInt(data.size())));
    std::vector<uint8_t> tail; put32(tail, crc);
    return std::fwrite(head.data(), 1, 4, f) == 4 && std::fwrite(type, 1, 4, f) == 4 &&
           (data.empty() || std::fwrite(data.data(), 1, data.size(), f) == data.size()) && std::fwrite(tail.data(), 1, 4, f) == 4;
}
} // namespace

bool write_png_rgb8(const std::string& path, const uint8_t* rgb, int64_t W, int64_t H, std::string& err)
{
    std::vector<uint8_t> raw(size_t(H) * (size_t(W) * 3 + 1));
    for (int64_t y = 0; y < H; ++y) {
        uint8_t* row = raw.data() + size_t(y) * (size_t(W) * 3 + 1);
        row[0] = 0; // filter: none
        std::copy(rgb + size_t(y) * size_t(W) * 3, rgb + size_t(y + 1) * size_t(W) * 3, row + 1);
    }
    uLongf clen = compressBound(uLong(raw.size()));
    std::vector<uint8_t> comp(clen);
    if (compress2(comp.data(), &clen, raw.data(), uLong(raw.size()), 6) != Z_OK) { err = "PNG deflate failed"; return false; }
    comp.resize(clen);
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) { err = "cannot open \"" + path + "\" for writing"; return false; }
    static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};
    std::vector<uint8_t> ihdr; put32(ihdr, uint32_t(W)); put32(ihdr, uint32_t(H));
    ihdr.push_back(8); ihdr.push_back(2); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);
    const bool ok = std::fwrite(sig, 1, 8, f) == 8 && chunk(f, "IHDR", ihdr) && chunk(f, "IDAT", comp) && chunk(f, "IEND", {});
    std::fclose(f);
    if (!ok) err = "write to \"" + path + "\" failed";
    return ok;
}
} // namespace vpt_host
