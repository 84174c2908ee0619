// oracle/ref/reference_capi.cpp — TEST INFRASTRUCTURE (CPU oracle, `oracle/_ref` build only). Not part of the product.
//
// C entry points over the REFERENCE'S OWN translation units, compiled unchanged from where they lie:
//   /root/reference/src/{worker,volume,majorant_transmittance_sampler,camera,precompute_blackbody,spectral,tile_provider,
//                        volume_grids,ray}.cpp   with   -I oracle/_ref/gen -I oracle/ref/eigen_shim -I /root/reference/include
// (oracle/Makefile target `refsrc` -> oracle/_ref/libvpt_reference.so). Eigen is replaced by oracle/ref/eigen_shim (Eigen is
// not installed here), vpt/random.hpp by a generated shadow that only hoists two specialisations (oracle/ref/gen_random_hpp.py);
// glaze (JSON), raylib (window) and spng (PNG) are not needed: this file fills vpt::Configuration's structs from the C POD and
// plays src/main.cpp:46-87 (TileProvider, film, Camera, one jthread per worker calling vpt::run) without the window.
// Purpose: pin oracle/vpt_restated*.hpp — tests/test_reference_compiled.py asserts restated == this, bit for bit — and serve as
// the "reference" CPU baseline of bench.py. Loaded with ctypes by tests/ and bench.py's CPU legs ONLY.
#include <vpt/camera.hpp>
#include <vpt/color.hpp>
#include <vpt/configuration.hpp>
#include <vpt/image.hpp>
#include <vpt/majorant_transmittance_sampler.hpp>
#include <vpt/random.hpp>
#include <vpt/spectral.hpp>
#include <vpt/tile_provider.hpp>
#include <vpt/volume.hpp>
#include <vpt/volume_grids.hpp>
#include <vpt/worker.hpp>

#include <memory>
#include <mutex>
#include <thread>

#include "vptc.h"

namespace {

using GridHandleT = vpt::VolumeGrids::GridHandleT;

GridHandleT handle_from_bytes(const void* p, size_t n)
{
    auto buf = nanovdb::HostBuffer::create(n);
    std::memcpy(buf.data(), p, n);
    return GridHandleT(std::move(buf));
}

struct Ctx {
    std::unique_ptr<vpt::VolumeGrids> grids;
};

Eigen::Vector3f v3(const float* p) { return Eigen::Vector3f(p[0], p[1], p[2]); }

vpt::CameraParameters camera_params(const vptc_camera_parameters& c)
{
    return vpt::CameraParameters{v3(c.position), v3(c.look), v3(c.up), c.vfov_deg, c.imaging_ratio};
}
vpt::WorkerParameters worker_params(const vptc_worker_parameters& w)
{
    vpt::WorkerParameters p;
    p.single_pixel.enabled = w.single_pixel_enabled != 0;
    p.single_pixel.coord = vpt::image_point_t{Eigen::Index(w.single_pixel_coord[0]), Eigen::Index(w.single_pixel_coord[1])};
    p.use_jitter = w.use_jitter != 0;
    p.infinite_light = vpt::InfiniteLightParameters{v3(w.infinite_light.xyz), w.infinite_light.multiplier};
    p.distant_light = vpt::DistantLightParameters{v3(w.distant_light.xyz), w.distant_light.multiplier, v3(w.distant_light.inv_direction)};
    p.max_depth = w.max_depth;
    return p;
}
vpt::VolumeParameters volume_params(const vptc_volume_parameters& v)
{
    return vpt::VolumeParameters{v.henyey_greenstein_g, v.le_scale, v.sigma_a, v.sigma_s, v.temperature_offset, v.temperature_scale};
}
vpt::image_size_t size2(const int64_t* s) { return vpt::image_size_t{Eigen::Index(s[0]), Eigen::Index(s[1])}; }

std::once_flag g_init;
void init_once()
{
    std::call_once(g_init, [] {
        vpt::init_blackbody_radiation_xyz(); // src/main.cpp:32
        std::clog.rdbuf(nullptr);            // vptINFO("Starting wave N") per wave (tile_provider.cpp:90): keep test output quiet
    });
}

} // namespace

extern "C" {

const char* vptr_about(void) { return "reference translation units compiled unchanged (Eigen stand-in, generated random.hpp shadow)"; }

int vptr_create(void** out, const void* density, size_t dbytes, const void* temperature, size_t tbytes)
{
    if (!out || !density || !dbytes) return -1;
    init_once();
    try {
        auto c = std::make_unique<Ctx>();
        if (temperature && tbytes) c->grids = std::make_unique<vpt::VolumeGrids>(handle_from_bytes(density, dbytes), handle_from_bytes(temperature, tbytes));
        else c->grids = std::make_unique<vpt::VolumeGrids>(handle_from_bytes(density, dbytes));
        *out = c.release();
    } catch (const std::exception&) {
        return -2;
    }
    return 0;
}
void vptr_destroy(void* ctx) { delete static_cast<Ctx*>(ctx); }

// src/main.cpp:41-87 without the window: Volume (applies fix_majorants_for_interpolation, volume.cpp:169 — idempotent, it reads voxel
// values only), TileProvider, zeroed film, Camera, num_threads jthreads each running vpt::run with its own RandomNumberGenerator(seed).
// Waves first_wave .. first_wave + n_waves - 1: TileProvider always starts at job 0, so the tokens of the earlier waves are taken and
// dropped unrendered first (a token's destructor marks its tile done for that wave, tile_provider.hpp:22-27) — public interface only.
// The reference's film starts at zero; it is ADDED into `film` afterwards.
int vptr_render(void* ctx, const vptc_scene* s, uint32_t first_wave, uint32_t n_waves, uint32_t num_threads, float* film, double* seconds)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c || !s || !film || first_wave < 1 || num_threads < 1) return -1;
    const vpt::WorkerParameters wp = worker_params(s->worker_parameters);
    vpt::Volume vol(*c->grids, volume_params(s->volume_parameters));
    const vpt::image_size_t out_size = size2(s->output_size), tile_size = size2(s->tile_size);
    vpt::TileProvider provider(out_size, first_wave + n_waves - 1, tile_size);
    const size_t ntiles = size_t((out_size.x() + tile_size.x() - 1) / tile_size.x()) * size_t((out_size.y() + tile_size.y() - 1) / tile_size.y());
    for (size_t k = 0; k < size_t(first_wave - 1) * ntiles; ++k) {
        auto tok = provider.next();
        if (!tok) return -3;
    }
    vpt::Image<float, 4> img(out_size);
    img.data().fill(decltype(img)::value_t::Zero());
    vpt::Camera camera(camera_params(s->camera_parameters), out_size);
    const auto t0 = std::chrono::steady_clock::now();
    {
        std::vector<std::jthread> threads;
        for (unsigned int i = 0; i < num_threads; ++i)
            threads.emplace_back([&]() {
                vpt::RandomNumberGenerator rng(s->seed);
                vpt::run(wp, vol, camera, provider, img, rng);
            });
    }
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    auto d = img.data();
    for (Eigen::Index i = 0; i < d.rows(); ++i)
        for (Eigen::Index j = 0; j < d.cols(); ++j)
            for (int k = 0; k < 4; ++k) film[(size_t(i) * size_t(d.cols()) + size_t(j)) * 4 + k] += d(i, j)[k];
    return 0;
}

// Patched leaf maxima in leaf-array order, after Volume::Volume ran fix_majorants_for_interpolation on the grid (volume.cpp:104-170).
int vptr_leaf_maxima(void* ctx, float* out, uint64_t cap, uint64_t* n_out)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c || !n_out) return -1;
    vpt::Volume vol(*c->grids, vpt::VolumeParameters{});
    auto& g = c->grids->density();
    const auto* leaf = g.tree().getFirstLeaf();
    const uint64_t n = g.tree().nodeCount<vpt::VolumeGrids::GridT::LeafNodeType>();
    *n_out = n;
    for (uint64_t i = 0; i < n && i < cap; ++i) out[i] = leaf[i].getMax();
    return 0;
}

// Volume::intersect + RayMajorantIterator::next (volume.cpp:38-88): merged majorant segments {t0, t1, d_maj} per world ray.
int vptr_trace_segments(void* ctx, const float* origins, const float* dirs, size_t n_rays, uint32_t max_segs, float* segs, uint32_t* n_segs)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c) return -1;
    vpt::Volume vol(*c->grids, vpt::VolumeParameters{});
    auto acc = c->grids->density().getAccessor();
    for (size_t r = 0; r < n_rays; ++r) {
        uint32_t n = 0;
        vpt::Ray ray(v3(origins + 3 * r), v3(dirs + 3 * r));
        if (auto it = vol.intersect(ray, acc)) {
            vpt::RayMajorantIterator iter = *it;
            while (auto seg = iter.next()) {
                if (n < max_segs) { float* o = segs + (r * max_segs + n) * 3; o[0] = seg->t0; o[1] = seg->t1; o[2] = seg->d_maj; }
                ++n;
            }
        }
        n_segs[r] = n;
    }
    return 0;
}

// MajorantTransmittanceSampler::next along world rays with a RandomNumberGenerator begun at job ids[r]: the tentative collisions
// {point xyz, sigma_maj, density} (majorant_transmittance_sampler.cpp:21-81); at most max_hits stored per ray.
int vptr_sample_collisions(void* ctx, uint32_t seed, const uint64_t* jids, const float* origins, const float* dirs, size_t n_rays, float sigma_t,
                           uint32_t max_hits, float* hits, uint32_t* n_hits)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c) return -1;
    vpt::Volume vol(*c->grids, vpt::VolumeParameters{});
    auto acc = c->grids->density().getAccessor();
    vpt::RandomNumberGenerator rng(seed);
    for (size_t r = 0; r < n_rays; ++r) {
        uint32_t n = 0;
        rng.begin_job(jids[r]);
        vpt::Ray ray(v3(origins + 3 * r), v3(dirs + 3 * r));
        if (auto it = vol.intersect(ray, acc)) {
            vpt::MajorantTransmittanceSampler sampler(*it, rng, c->grids->density(), acc, sigma_t);
            while (auto props = sampler.next()) {
                if (n < max_hits) { float* o = hits + (r * max_hits + n) * 5; o[0] = props->point.x(); o[1] = props->point.y(); o[2] = props->point.z(); o[3] = props->sigma_maj; o[4] = props->density; }
                ++n;
            }
        }
        n_hits[r] = n;
    }
    return 0;
}

// Volume::log_dda_trace (volume.cpp:194-225): writes "dda_trace.csv" into the CURRENT DIRECTORY (the reference hard-codes the name).
int vptr_log_dda_trace(void* ctx, const float* origin, const float* dir)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c) return -1;
    vpt::Volume vol(*c->grids, vpt::VolumeParameters{});
    auto acc = c->grids->density().getAccessor();
    vol.log_dda_trace(vpt::Ray(v3(origin), v3(dir)), acc);
    return 0;
}

// TileProvider::next / token::compute_rect for wave 1 (tile_provider.cpp:27-67,95-105): {x0, y0, w, h, jid} per tile, in hand-out order.
int vptr_tile_rects(const vptc_scene* s, int64_t* out5, size_t cap, size_t* n_out)
{
    init_once();
    vpt::TileProvider provider(size2(s->output_size), 1, size2(s->tile_size));
    size_t n = 0;
    while (auto tok = provider.next()) {
        if (n < cap) {
            const vpt::image_rect_t r = tok.compute_rect();
            int64_t* o = out5 + n * 5;
            o[0] = r.start.x(); o[1] = r.start.y(); o[2] = r.size.x(); o[3] = r.size.y(); o[4] = int64_t(tok.jid());
        }
        ++n;
    }
    *n_out = n;
    return 0;
}

void vptr_rng(uint32_t seed, uint64_t jid, size_t n, uint64_t* state0, uint32_t* out_u32, float* out_f32)
{
    if (state0) *state0 = vpt::hash(uint64_t(seed), size_t(jid));
    vpt::RandomNumberGenerator a(seed), b(seed);
    a.begin_job(jid); b.begin_job(jid);
    for (size_t i = 0; i < n; ++i) { if (out_u32) out_u32[i] = a.uniform<uint32_t>(); if (out_f32) out_f32[i] = b.uniform<float>(); }
}

void vptr_camera_rays(const vptc_scene* s, size_t n, const int64_t* px, const int64_t* py, const float* jx, const float* jy, float* origins, float* dirs)
{
    const vpt::Camera cam(camera_params(s->camera_parameters), size2(s->output_size));
    for (size_t i = 0; i < n; ++i) {
        const vpt::Ray r = cam.generate_ray(vpt::image_point_t{Eigen::Index(px[i]), Eigen::Index(py[i])}, Eigen::Vector2f{jx[i], jy[i]});
        for (int k = 0; k < 3; ++k) { origins[3 * i + k] = r.origin()[k]; dirs[3 * i + k] = r.direction()[k]; }
    }
}

void vptr_blackbody(const float* temps, size_t n, float* out_xyz)
{
    init_once();
    for (size_t i = 0; i < n; ++i) { const Eigen::Vector3f v = vpt::blackbody_radiation_xyz(temps[i]); for (int k = 0; k < 3; ++k) out_xyz[3 * i + k] = v[k]; }
}
// The 500-entry table (precompute_blackbody.cpp:12-22) is file-static; entry i is what an exact hit at (i-1)*100 K returns (:45-47),
// entries 0 and 1 belong to T <= 0 and are zero (planck_law returns 0 there, spectral.cpp:8).
void vptr_blackbody_table(float* out1500)
{
    init_once();
    for (int i = 0; i < 500; ++i) {
        const Eigen::Vector3f v = i < 2 ? Eigen::Vector3f::Zero() : vpt::blackbody_radiation_xyz(float(i - 1) * 100.0f);
        for (int k = 0; k < 3; ++k) out1500[3 * i + k] = v[k];
    }
}
void vptr_hg_sample(const float* w, const float* u, size_t n, float g, float* out)
{
    for (size_t i = 0; i < n; ++i) { const Eigen::Vector3f d = vpt::sample_henyey_greenstein(v3(w + 3 * i), Eigen::Vector2f{u[2 * i], u[2 * i + 1]}, g); for (int k = 0; k < 3; ++k) out[3 * i + k] = d[k]; }
}
void vptr_hg_eval(const float* cos_theta, size_t n, float g, float* out) { for (size_t i = 0; i < n; ++i) out[i] = vpt::henyey_greenstein(cos_theta[i], g); }

// film_to_image (src/main.cpp:12-24; main.cpp itself needs raylib.h and is not compiled): the same five statements over the reference's
// own xyz_to_linsrgb / linsrgb_to_srgb (include/vpt/color.hpp:8-30). A w = 0 pixel is 0/0 = NaN, whose cast to unsigned char is
// undefined in the reference; written as 0 here.
void vptr_film_to_image(const float* film, int64_t W, int64_t H, uint8_t* rgb)
{
    for (int64_t i = 0; i < W * H; ++i) {
        const Eigen::Vector4f xyzw(film[4 * i], film[4 * i + 1], film[4 * i + 2], film[4 * i + 3]);
        const Eigen::Vector3f xyz = xyzw.topRows<3>() / xyzw.w();
        const Eigen::Vector3f linsrgb = vpt::xyz_to_linsrgb(xyz);
        const Eigen::Vector3f srgb = vpt::linsrgb_to_srgb(linsrgb);
        const Eigen::Vector3f scaled = srgb.cwiseMax(0.0f).cwiseMin(1.0f) * 255.0f;
        for (int k = 0; k < 3; ++k) rgb[3 * i + k] = scaled[k] == scaled[k] ? static_cast<unsigned char>(scaled[k]) : 0;
    }
}

} // extern "C"
