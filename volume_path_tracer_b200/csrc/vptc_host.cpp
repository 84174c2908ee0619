// vptc_host.cpp — the host-only entry points of include/vptc.h (no CUDA calls here):
// .nvdb I/O, synthetic volumes, scene JSON, PNG. The device side lives in vptc_device.cu.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include <vector>

#include "host/dda_trace.hpp"
#include "host/json_scene.hpp"
#include "host/nvdb_build.hpp"
#include "host/nvdb_io.hpp"
#include "host/png_write.hpp"
#include "host/scene_prep.hpp"
#include "host/synth.hpp"
#include "vptc.h"
#include "vptc_internal.hpp"

namespace vptc_internal {
static thread_local std::string g_tls_error;
void set_global_error(const std::string& msg) { g_tls_error = msg; }
const char* global_error() { return g_tls_error.c_str(); }
} // namespace vptc_internal
using vptc_internal::set_global_error;

namespace {
void* dup_aligned(const std::vector<uint8_t>& v)
{
    void* p = nullptr;
    if (posix_memalign(&p, 64, v.size() ? v.size() : 64) != 0) return nullptr;
    std::memcpy(p, v.data(), v.size());
    return p;
}
} // namespace

extern "C" {

const char* vptc_host_last_error(void) { return vptc_internal::global_error(); }

const char* vptc_version(void) { return "vpt-b200 0.1 (sm_100a)"; }

void vptc_free(void* p) { std::free(p); }

int vptc_nvdb_read_grid(const char* path, const char* grid_name, void** grid, size_t* bytes)
{
    if (!path || !grid_name || !grid || !bytes) { set_global_error("vptc_nvdb_read_grid: null argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    try {
        const std::string err = vpt_host::read_nvdb_grid(path, grid_name, grid, bytes);
        if (!err.empty()) {
            set_global_error(err);
            return err.rfind("Grid name", 0) == 0 ? VPTC_ERR_BAD_GRID : VPTC_ERR_IO;
        }
        if (const char* bad = vpt_host::check_grid(*grid, *bytes)) {
            std::free(*grid); *grid = nullptr; *bytes = 0;
            set_global_error(std::string("grid \"") + grid_name + "\": " + bad);
            return VPTC_ERR_BAD_GRID;
        }
        return VPTC_OK;
    } catch (const std::exception& e) { set_global_error(e.what()); return VPTC_ERR_IO; }
}

int vptc_nvdb_write_grids(const char* path, const void* const* grids, size_t n_grids)
{
    if (!path || !grids) { set_global_error("vptc_nvdb_write_grids: null argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    try {
        const std::string err = vpt_host::write_nvdb_grids(path, grids, n_grids);
        if (!err.empty()) { set_global_error(err); return VPTC_ERR_IO; }
        return VPTC_OK;
    } catch (const std::exception& e) { set_global_error(e.what()); return VPTC_ERR_IO; }
}

int vptc_synth_volume(const char* kind, const int32_t dims[3], double voxel_size, uint32_t seed, const double* anchor_xyz, void** density,
                      size_t* density_bytes, void** temperature, size_t* temperature_bytes)
{
    if (!kind || !dims || !density || !density_bytes) { set_global_error("vptc_synth_volume: null argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    try {
        vpt_host::SynthResult r = vpt_host::synth_volume(kind, dims, voxel_size, seed, anchor_xyz);
        *density = dup_aligned(r.density); *density_bytes = r.density.size();
        if (temperature && temperature_bytes) {
            if (!r.temperature.empty()) { *temperature = dup_aligned(r.temperature); *temperature_bytes = r.temperature.size(); }
            else { *temperature = nullptr; *temperature_bytes = 0; }
        }
        return VPTC_OK;
    } catch (const std::exception& e) { set_global_error(e.what()); return VPTC_ERR_INVALID_ARGUMENT; }
}

int vptc_scene_from_json(const char* json_text, vptc_scene* out, char* volume_path_out, size_t volume_path_cap)
{
    if (!json_text || !out) { set_global_error("vptc_scene_from_json: null argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    std::string vol, err;
    if (!vpt_host::parse_scene_json(json_text, *out, vol, err)) { set_global_error(err); return VPTC_ERR_CONFIG; }
    if (volume_path_out && volume_path_cap) {
        if (vol.size() + 1 > volume_path_cap) { set_global_error("volume_path buffer too small"); return VPTC_ERR_INVALID_ARGUMENT; }
        std::memcpy(volume_path_out, vol.c_str(), vol.size() + 1);
    }
    return VPTC_OK;
}

int vptc_scene_from_json_file(const char* path, vptc_scene* out, char* volume_path_out, size_t volume_path_cap)
{
    if (!path || !out) { set_global_error("vptc_scene_from_json_file: null argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    std::string text, err;
    if (!vpt_host::read_text_file(path, text, err)) { set_global_error(err); return VPTC_ERR_IO; }
    const int rc = vptc_scene_from_json(text.c_str(), out, volume_path_out, volume_path_cap);
    if (rc != VPTC_OK) set_global_error(std::string("Failed to read configuration file \"") + path + "\": " + vptc_internal::global_error());
    return rc;
}

int vptc_write_png(const char* path, const uint8_t* rgb8, int64_t width, int64_t height)
{
    if (!path || !rgb8 || width <= 0 || height <= 0) { set_global_error("vptc_write_png: bad argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    std::string err;
    if (!vpt_host::write_png_rgb8(path, rgb8, width, height, err)) { set_global_error(err); return VPTC_ERR_IO; }
    return VPTC_OK;
}

// ---- film checkpoint ----
namespace {
struct FilmHeader { char magic[8]; uint32_t version; uint32_t next_wave; int64_t width, height; uint64_t fingerprint; };
// FNV-1a over every scene field that determines the film, field by field (the POD has padding); num_waves and
// num_workers are left out on purpose.
uint64_t scene_fingerprint(const vptc_scene& s)
{
    uint64_t h = 1469598103934665603ULL;
    auto mix = [&](const void* p, size_t n) { for (size_t i = 0; i < n; ++i) { h ^= static_cast<const unsigned char*>(p)[i]; h *= 1099511628211ULL; } };
    mix(&s.seed, sizeof s.seed); mix(s.output_size, sizeof s.output_size); mix(s.tile_size, sizeof s.tile_size);
    const vptc_camera_parameters& c = s.camera_parameters;
    mix(c.position, sizeof c.position); mix(c.look, sizeof c.look); mix(c.up, sizeof c.up); mix(&c.vfov_deg, 4); mix(&c.imaging_ratio, 4);
    const vptc_worker_parameters& w = s.worker_parameters;
    mix(&w.single_pixel_enabled, 4); mix(w.single_pixel_coord, sizeof w.single_pixel_coord); mix(&w.use_jitter, 4);
    mix(w.infinite_light.xyz, 12); mix(&w.infinite_light.multiplier, 4);
    mix(w.distant_light.xyz, 12); mix(&w.distant_light.multiplier, 4); mix(w.distant_light.inv_direction, 12); mix(&w.max_depth, 4);
    const vptc_volume_parameters& v = s.volume_parameters;
    mix(&v.henyey_greenstein_g, 4); mix(&v.le_scale, 4); mix(&v.sigma_a, 4); mix(&v.sigma_s, 4); mix(&v.temperature_offset, 4); mix(&v.temperature_scale, 4);
    return h;
}
const char kFilmMagic[8] = {'V', 'P', 'T', 'F', 'I', 'L', 'M', '1'};
} // namespace

int vptc_film_save(const char* path, const vptc_scene* scene, const float* film_xyzw, uint32_t next_wave)
{
    if (!path || !scene || !film_xyzw || scene->output_size[0] <= 0 || scene->output_size[1] <= 0) { set_global_error("vptc_film_save: bad argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    FilmHeader h{};
    std::memcpy(h.magic, kFilmMagic, 8);
    h.version = 1; h.next_wave = next_wave; h.width = scene->output_size[0]; h.height = scene->output_size[1]; h.fingerprint = scene_fingerprint(*scene);
    const size_t n = size_t(h.width) * size_t(h.height) * 4;
    const std::string tmp = std::string(path) + ".tmp";
    std::FILE* f = std::fopen(tmp.c_str(), "wb");
    if (!f) { set_global_error("vptc_film_save: cannot open " + tmp); return VPTC_ERR_IO; }
    const bool ok = std::fwrite(&h, sizeof h, 1, f) == 1 && std::fwrite(film_xyzw, sizeof(float), n, f) == n;
    const bool closed = std::fclose(f) == 0;
    if (!ok || !closed || std::rename(tmp.c_str(), path) != 0) { std::remove(tmp.c_str()); set_global_error(std::string("vptc_film_save: write failed for ") + path); return VPTC_ERR_IO; }
    return VPTC_OK;
}

// Volume::log_dda_trace (src/volume.cpp:194-225) for the un-jittered camera ray through the centre of pixel (px, py): host code on the
// caller's NanoVDB buffer (value masks and node statistics are not part of the baked device layout).
int vptc_dda_trace_csv(const void* density_grid, size_t density_bytes, const vptc_scene* scene, int64_t px, int64_t py, const char* csv_path, size_t* n_rows)
{
    if (n_rows) *n_rows = 0;
    if (!density_grid || !scene || !csv_path) { set_global_error("vptc_dda_trace_csv: null argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    if (scene->output_size[0] <= 0 || scene->output_size[1] <= 0 || scene->tile_size[0] <= 0 || scene->tile_size[1] <= 0 || px < 0 || py < 0 || px >= scene->output_size[0] ||
        py >= scene->output_size[1]) { set_global_error("vptc_dda_trace_csv: pixel outside the image"); return VPTC_ERR_INVALID_ARGUMENT; }
    vptd::SceneDev S;
    vpt_host::prepare_scene(*scene, false, S);
    // Camera::generate_ray (camera.hpp:14-23) with jitter 0: the kernels' arithmetic, on the host (-ffp-contract=off)
    auto sum3 = [](float a, float b, float c) { return a + (b + c); };
    const float rx = float(px) + 0.5f, ry = float(py) + 0.5f;
    float d[3];
    const float o[3] = {S.cam_pos[0], S.cam_pos[1], S.cam_pos[2]};
    for (int r = 0; r < 3; ++r) d[r] = sum3(S.cam_m[r * 3] * rx, S.cam_m[r * 3 + 1] * ry, S.cam_m[r * 3 + 2] * 0.0f) + S.cam_t[r];
    const float nrm = std::sqrt(sum3(d[0] * d[0], d[1] * d[1], d[2] * d[2]));
    for (int r = 0; r < 3; ++r) d[r] /= nrm;
    try {
        const std::string err = vpt_host::write_dda_trace(density_grid, density_bytes, o, d, csv_path, n_rows);
        if (!err.empty()) { set_global_error("vptc_dda_trace_csv: " + err); return VPTC_ERR_BAD_GRID; }
    } catch (const std::exception& e) { set_global_error(std::string("vptc_dda_trace_csv: ") + e.what()); return VPTC_ERR_IO; }
    return VPTC_OK;
}
// The same for an explicit world-space ray (the reference's log_dda_trace takes a vpt::Ray).
int vptc_dda_trace_ray_csv(const void* density_grid, size_t density_bytes, const float origin_xyz[3], const float direction_xyz[3], const char* csv_path, size_t* n_rows)
{
    if (n_rows) *n_rows = 0;
    if (!density_grid || !origin_xyz || !direction_xyz || !csv_path) { set_global_error("vptc_dda_trace_ray_csv: null argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    try {
        const std::string err = vpt_host::write_dda_trace(density_grid, density_bytes, origin_xyz, direction_xyz, csv_path, n_rows);
        if (!err.empty()) { set_global_error("vptc_dda_trace_ray_csv: " + err); return VPTC_ERR_BAD_GRID; }
    } catch (const std::exception& e) { set_global_error(std::string("vptc_dda_trace_ray_csv: ") + e.what()); return VPTC_ERR_IO; }
    return VPTC_OK;
}

int vptc_film_load(const char* path, const vptc_scene* scene, float* film_xyzw, uint32_t* next_wave)
{
    if (!path || !scene || !film_xyzw || !next_wave) { set_global_error("vptc_film_load: bad argument"); return VPTC_ERR_INVALID_ARGUMENT; }
    std::FILE* f = std::fopen(path, "rb");
    if (!f) { set_global_error(std::string("vptc_film_load: cannot open ") + path); return VPTC_ERR_IO; }
    FilmHeader h{};
    int rc = VPTC_OK;
    if (std::fread(&h, sizeof h, 1, f) != 1 || std::memcmp(h.magic, kFilmMagic, 8) != 0 || h.version != 1) { set_global_error(std::string("vptc_film_load: not a film checkpoint: ") + path); rc = VPTC_ERR_IO; }
    else if (h.width != scene->output_size[0] || h.height != scene->output_size[1] || h.fingerprint != scene_fingerprint(*scene)) {
        set_global_error(std::string("vptc_film_load: checkpoint was written for a different scene: ") + path); rc = VPTC_ERR_CONFIG;
    } else {
        const size_t n = size_t(h.width) * size_t(h.height) * 4;
        std::vector<float> buf(n); // the caller's film is only touched by a complete read
        if (std::fread(buf.data(), sizeof(float), n, f) != n) { set_global_error(std::string("vptc_film_load: truncated checkpoint: ") + path); rc = VPTC_ERR_IO; }
        else { std::memcpy(film_xyzw, buf.data(), n * sizeof(float)); *next_wave = h.next_wave; }
    }
    std::fclose(f);
    return rc;
}

} // extern "C"
