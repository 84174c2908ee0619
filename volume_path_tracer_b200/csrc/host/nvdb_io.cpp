#define _FILE_OFFSET_BITS 64
// nvdb_io.cpp — read/write `.nvdb` files without the NanoVDB headers.
//
// Replaces nanovdb::io::readGrid<HostBuffer>(path, name) / io::writeGrids as the reference uses them
// (src/volume_grids.cpp:39-67). File = sequence of segments; a segment = FileHeader, gridCount x
// (FileMetaData + name), then the grid buffers (codec NONE: raw bytes; ZIP: u64 size + zlib stream)
// (NV/io/IO.h:204-320, NV/NanoVDB.h:5817-5856). A file that starts with the grid magic is a raw dump
// of concatenated grid buffers (NV/NanoVDB.h:5861-5865).
#include "nvdb_io.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>
#include <sys/types.h>
#include <vector>
#include <zlib.h>

#include "nvdb_format.h"

namespace vpt_host {
using namespace nvdbf;

namespace {
struct File {
    FILE* f = nullptr;
    ~File() { if (f) std::fclose(f); }
};
bool rd(FILE* f, void* p, size_t n) { return std::fread(p, 1, n, f) == n; }

void* alloc_grid(size_t bytes)
{
    void* p = nullptr;
    if (posix_memalign(&p, 64, bytes ? bytes : 64) != 0) return nullptr;
    return p;
}
} // namespace

// NanoVDB's grid-name hash (NV/io/IO.h:717-728): h = 67*h + c + (h >> 56), 64-bit wrap.
static uint64_t name_key(const char* s)
{
    uint64_t h = 0;
    for (; *s; ++s) {
        const uint64_t overflow = h >> 56;
        h = h * 67 + uint64_t(uint8_t(*s)) + overflow;
    }
    return h;
}

// Every size field comes from the file: each is checked against the bytes that remain before it is used to allocate, read or
// skip, and every skip moves strictly forward (64-bit offsets), so a hostile file ends in an error — never in a rewind loop,
// a giant allocation or a short read taken for a grid.
std::string read_nvdb_grid(const std::string& path, const std::string& name, void** out, size_t* outBytes)
{
    *out = nullptr; *outBytes = 0;
    File file; file.f = std::fopen(path.c_str(), "rb");
    if (!file.f) return "cannot open \"" + path + "\"";
    FILE* f = file.f;
    if (fseeko(f, 0, SEEK_END) != 0) return "cannot seek in \"" + path + "\"";
    const uint64_t file_size = uint64_t(ftello(f));
    if (fseeko(f, 0, SEEK_SET) != 0) return "cannot seek in \"" + path + "\"";
    uint64_t pos = 0; // offset of the next unread byte; only ever grows
    auto remaining = [&]() { return file_size - pos; };
    auto read_at_pos = [&](void* p, uint64_t n) { // n <= remaining() is the caller's check
        if (!rd(f, p, size_t(n))) return false;
        pos += n;
        return true;
    };
    auto skip = [&](uint64_t n) {
        if (n > remaining()) return false;
        pos += n;
        return fseeko(f, off_t(pos), SEEK_SET) == 0;
    };
    uint64_t magic = 0;
    if (file_size < 8 || !rd(f, &magic, 8)) return "file too short";
    if (fseeko(f, 0, SEEK_SET) != 0) return "cannot seek in \"" + path + "\"";

    if (magic == MAGIC_GRID) { // raw concatenated grid buffers
        while (remaining() >= sizeof(GridData)) {
            GridData g;
            const uint64_t grid_pos = pos;
            if (!read_at_pos(&g, sizeof(g))) return "truncated grid header";
            if (g.magic != MAGIC_GRID || g.gridSize < sizeof(GridData) || g.gridSize - sizeof(GridData) > remaining()) return "corrupt raw grid buffer";
            if (std::memchr(g.gridName, 0, sizeof(g.gridName)) && name == g.gridName) {
                void* p = alloc_grid(g.gridSize);
                if (!p) return "out of memory";
                pos = grid_pos;
                if (fseeko(f, off_t(pos), SEEK_SET) != 0 || !read_at_pos(p, g.gridSize)) { std::free(p); return "truncated grid"; }
                *out = p; *outBytes = g.gridSize;
                return "";
            }
            if (!skip(g.gridSize - sizeof(GridData))) return "corrupt raw grid buffer";
        }
        return "Grid name '" + name + "' not found in file";
    }

    while (remaining() >= sizeof(FileHeader)) { // segments
        FileHeader h;
        if (!read_at_pos(&h, sizeof(h))) return "truncated segment header";
        if (h.magic != MAGIC_NUMB && h.magic != MAGIC_FILE) return "bad segment magic in \"" + path + "\"";
        if (version_major(h.version) != ABI_MAJOR) return "unsupported NanoVDB file version";
        if (uint64_t(h.gridCount) * sizeof(FileMetaData) > remaining()) return "truncated metadata";
        std::vector<FileMetaData> metas(h.gridCount);
        std::vector<std::string> names(h.gridCount);
        for (uint16_t i = 0; i < h.gridCount; ++i) {
            if (remaining() < sizeof(FileMetaData) || !read_at_pos(&metas[i], sizeof(FileMetaData))) return "truncated metadata";
            if (metas[i].nameSize > remaining()) return "truncated grid name";
            std::string nm(metas[i].nameSize, '\0');
            if (metas[i].nameSize && !read_at_pos(nm.data(), metas[i].nameSize)) return "truncated grid name";
            names[i] = nm.c_str(); // strip NUL
        }
        for (uint16_t i = 0; i < h.gridCount; ++i) {
            const FileMetaData& m = metas[i];
            if (m.fileSize == 0 || m.fileSize > remaining()) return "corrupt segment: grid payload size exceeds the file";
            if (names[i] != name) { if (!skip(m.fileSize)) return "corrupt segment"; continue; }
            if (m.gridSize < sizeof(GridData) || m.gridSize > (uint64_t(1) << 40)) return "corrupt segment: implausible grid size";
            std::string err;
            void* p = nullptr;
            if (h.codec == 0) {
                if (m.gridSize > remaining()) return "truncated grid";
                if (!(p = alloc_grid(m.gridSize))) return "out of memory";
                if (!read_at_pos(p, m.gridSize)) err = "truncated grid";
            } else if (h.codec == 1) { // ZIP (codec is taken from the segment header, NV/io/IO.h:512)
                uint64_t csize = 0;
                if (remaining() < 8 || !read_at_pos(&csize, 8) || csize == 0 || csize > remaining()) return "ZIP decode failed: bad compressed size";
                if (m.gridSize / 1100 > csize) return "ZIP decode failed: grid size impossible for the compressed size"; // deflate expands <= ~1032x
                std::unique_ptr<uint8_t[]> tmp(new (std::nothrow) uint8_t[csize]);
                if (!tmp) return "out of memory";
                if (!read_at_pos(tmp.get(), csize)) return "ZIP decode failed";
                if (!(p = alloc_grid(m.gridSize))) return "out of memory";
                uLongf n = uLongf(m.gridSize);
                if (!(uncompress(static_cast<Bytef*>(p), &n, tmp.get(), uLong(csize)) == Z_OK && n == m.gridSize)) err = "ZIP decode failed";
            } else {
                err = "BLOSC-compressed .nvdb is not supported";
            }
            if (!err.empty()) { std::free(p); return err; }
            *out = p; *outBytes = m.gridSize;
            return "";
        }
    }
    return "Grid name '" + name + "' not found in file";
}

std::string write_nvdb_grids(const std::string& path, const void* const* grids, size_t n)
{
    File file; file.f = std::fopen(path.c_str(), "wb");
    if (!file.f) return "cannot open \"" + path + "\" for writing";
    FILE* f = file.f;
    // one segment per grid, as io::writeGrids does (NV/io/IO.h:470-500)
    for (size_t i = 0; i < n; ++i) {
        const auto* g = static_cast<const GridData*>(grids[i]);
        if (g->magic != MAGIC_GRID && g->magic != MAGIC_NUMB) return "not a NanoVDB grid buffer";
        const TreeData* t = tree_of(g);
        const RootData* r = root_of(g);
        FileHeader h{MAGIC_FILE, pack_version(ABI_MAJOR, ABI_MINOR, ABI_PATCH), 1, 0};
        FileMetaData m;
        std::memset(&m, 0, sizeof(m));
        m.gridSize = m.fileSize = g->gridSize;
        m.nameKey = name_key(g->gridName);
        m.voxelCount = t->voxelCount;
        m.gridType = g->gridType; m.gridClass = g->gridClass;
        std::memcpy(m.worldBBox, g->worldBBox, sizeof(m.worldBBox));
        std::memcpy(m.indexBBox, r->bbox, sizeof(m.indexBBox));
        std::memcpy(m.voxelSize, g->voxelSize, sizeof(m.voxelSize));
        m.nameSize = uint32_t(std::strlen(g->gridName) + 1);
        m.nodeCount[0] = t->nodeCount[0]; m.nodeCount[1] = t->nodeCount[1]; m.nodeCount[2] = t->nodeCount[2]; m.nodeCount[3] = 1;
        m.tileCount[0] = t->tileCount[0]; m.tileCount[1] = t->tileCount[1]; m.tileCount[2] = t->tileCount[2];
        m.codec = 0; m.version = g->version;
        if (std::fwrite(&h, sizeof(h), 1, f) != 1 || std::fwrite(&m, sizeof(m), 1, f) != 1 ||
            std::fwrite(g->gridName, 1, m.nameSize, f) != m.nameSize ||
            std::fwrite(g, 1, g->gridSize, f) != g->gridSize)
            return "write failed";
    }
    return "";
}

} // namespace vpt_host
