// oracle/oracle_capi.cpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product.
//
// C entry points of the CPU oracle, loaded with ctypes by tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs ONLY. Built twice from this one file:
//   oracle/libvpt_oracle.so           self-contained (substrate_own.hpp), builds anywhere with g++
//   oracle/_ref/libvpt_oracle_ref.so  -DVPTO_USE_NANOVDB: on the reference's vendored NanoVDB/pcg headers
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>

#ifdef VPTO_USE_NANOVDB
#include "ref/substrate_nanovdb.hpp"
using S = vpto_nano::Substrate;
#else
#include "substrate_own.hpp"
using S = vpto_own::Substrate;
#endif
#include "vpt_restated.hpp"

using namespace vpto;

namespace {
struct AlignedBuf {
    void* p = nullptr; size_t n = 0;
    bool assign(const void* src, size_t bytes)
    {
        if (posix_memalign(&p, 64, bytes ? bytes : 64) != 0) return false;
        std::memcpy(p, src, bytes); n = bytes; return true;
    }
    ~AlignedBuf() { std::free(p); }
};
struct Ctx {
    AlignedBuf dens, temp;
    std::unique_ptr<S::Grid> gd, gt;
    bool patched = false;
};
Volume<S> make_volume(Ctx* c, const vptc_volume_parameters& vp)
{
    // fix_majorants_for_interpolation mutates the grid buffer; the reference applies it once (Volume::Volume).
    Volume<S> v(*c->gd, c->gt.get(), vp, /*apply_patch=*/!c->patched);
    c->patched = true;
    return v;
}
void copy_counters(const Counters& c, vptc_counters* o)
{
    if (!o) return;
    o->pixel_samples = c.pixel_samples; o->density_evals = c.density_evals; o->temperature_evals = c.temperature_evals;
    o->majorant_fetches = c.majorant_fetches; o->rng_draws = c.rng_draws; o->scatter_events = c.scatter_events; o->shadow_rays = c.shadow_rays;
}
} // namespace

extern "C" {

const char* vpto_substrate(void) { return S::name(); }

int vpto_create(void** out, const void* density, size_t dbytes, const void* temperature, size_t tbytes)
{
    if (!out || !density || dbytes < sizeof(nvdbf::GridData)) return -1;
    auto c = std::make_unique<Ctx>();
    if (!c->dens.assign(density, dbytes)) return -2;
    c->gd = std::make_unique<S::Grid>(c->dens.p);
    if (temperature && tbytes) {
        if (!c->temp.assign(temperature, tbytes)) return -2;
        c->gt = std::make_unique<S::Grid>(c->temp.p);
    }
    vptc_volume_parameters vp{};
    (void)make_volume(c.get(), vp); // apply fix_majorants_for_interpolation once, as Volume::Volume does
    *out = c.release();
    return 0;
}
void vpto_destroy(void* ctx) { delete static_cast<Ctx*>(ctx); }

// Patched leaf maxima in leaf-array order (for baker parity).
int vpto_leaf_maxima(void* ctx, float* out, uint64_t cap, uint64_t* n_out)
{
    auto* c = static_cast<Ctx*>(ctx);
    const auto* g = static_cast<const nvdbf::GridData*>(c->dens.p);
    const nvdbf::TreeData* t = nvdbf::tree_of(g);
    const auto* leaf = reinterpret_cast<const nvdbf::LeafData*>(reinterpret_cast<const char*>(t) + t->nodeOffset[0]);
    *n_out = t->nodeCount[0];
    for (uint64_t i = 0; i < t->nodeCount[0] && i < cap; ++i) out[i] = leaf[i].maximum;
    return 0;
}

int vpto_render(void* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, uint32_t num_threads, float* film,
                vptc_counters* counters, double* seconds)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c || !scene || !film || first_wave < 1 || num_threads < 1) return -1;
    Volume<S> vol = make_volume(c, scene->volume_parameters);
    Counters cnt;
    const double s = render<S>(*scene, vol, first_wave, n_waves, num_threads, film, &cnt);
    copy_counters(cnt, counters);
    if (seconds) *seconds = s;
    return 0;
}

// Bounded sample for the timed CPU baseline: every job_stride-th job of the wave range (see render()).
// With job_stride > 1 two threads may hit the same tile in different waves; the film is then only a timing
// by-product (callers pass a scratch film), the counters stay exact.
int vpto_render_sample(void* ctx, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, uint64_t job_stride, uint32_t num_threads,
                       float* film, vptc_counters* counters, double* seconds)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c || !scene || !film || first_wave < 1 || num_threads < 1 || job_stride < 1) return -1;
    Volume<S> vol = make_volume(c, scene->volume_parameters);
    Counters cnt;
    const double s = render<S>(*scene, vol, first_wave, n_waves, num_threads, film, &cnt, job_stride);
    copy_counters(cnt, counters);
    if (seconds) *seconds = s;
    return 0;
}

int vpto_trace_jobs(void* ctx, const vptc_scene* scene, const uint64_t* jids, size_t n_jobs, float* per_pixel_xyz, uint32_t* per_pixel_ndraws,
                    vptc_counters* counters)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c || !scene || !jids) return -1;
    Volume<S> vol = make_volume(c, scene->volume_parameters);
    const TileGrid tg(*scene);
    const Camera camera(scene->camera_parameters, tg.W, tg.H);
    const size_t slots = size_t(tg.tw) * size_t(tg.th);
    if (per_pixel_xyz) std::memset(per_pixel_xyz, 0, n_jobs * slots * 3 * sizeof(float));
    if (per_pixel_ndraws) std::memset(per_pixel_ndraws, 0, n_jobs * slots * sizeof(uint32_t));
    Counters cnt;
    Rng<S::Engine> rng(scene->seed);
    auto density_acc = vol.m_density.getAccessor();
    auto temp_acc = vol.m_temperature.getAccessor();
    S::Sampler temp_sampler(temp_acc);
    for (size_t j = 0; j < n_jobs; ++j)
        run_job<S>(*scene, vol, camera, tg, jids[j], rng, density_acc, vol.m_has_temp ? &temp_sampler : nullptr, nullptr,
                   per_pixel_xyz ? per_pixel_xyz + j * slots * 3 : nullptr, per_pixel_ndraws ? per_pixel_ndraws + j * slots : nullptr, &cnt);
    copy_counters(cnt, counters);
    return 0;
}

// Logger<true> + single_pixel (worker.cpp:15-49,113-119): the event log of the one camera path job `jid` traces through pixel
// (px, py) when every other pixel of its tile is skipped before drawing. 8 floats per event (see Counters::event).
int vpto_path_log(void* ctx, const vptc_scene* scene_in, uint64_t jid, int64_t px, int64_t py, float* events, size_t cap_events, size_t* n_events)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c || !scene_in || !n_events) return -1;
    vptc_scene scene = *scene_in;
    scene.worker_parameters.single_pixel_enabled = 1;
    scene.worker_parameters.single_pixel_coord[0] = px; scene.worker_parameters.single_pixel_coord[1] = py;
    Volume<S> vol = make_volume(c, scene.volume_parameters);
    const TileGrid tg(scene);
    const Camera camera(scene.camera_parameters, tg.W, tg.H);
    std::vector<float> log;
    Counters cnt; cnt.log = &log;
    Rng<S::Engine> rng(scene.seed);
    auto density_acc = vol.m_density.getAccessor();
    auto temp_acc = vol.m_temperature.getAccessor();
    S::Sampler temp_sampler(temp_acc);
    run_job<S>(scene, vol, camera, tg, jid, rng, density_acc, vol.m_has_temp ? &temp_sampler : nullptr, nullptr, nullptr, nullptr, &cnt);
    *n_events = log.size() / 8;
    if (events) std::memcpy(events, log.data(), std::min(log.size(), cap_events * 8) * sizeof(float));
    return 0;
}

int vpto_trace_segments(void* ctx, const float* origins, const float* dirs, size_t n_rays, uint32_t max_segs, float* segs, uint32_t* n_segs)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c) return -1;
    vptc_volume_parameters vp{};
    Volume<S> vol = make_volume(c, vp);
    auto acc = vol.m_density.getAccessor();
    Counters cnt;
    for (size_t r = 0; r < n_rays; ++r) {
        uint32_t n = 0;
        if (auto it = vol.intersect(to_v3(origins + 3 * r), to_v3(dirs + 3 * r), acc, &cnt)) {
            auto iter = *it;
            while (auto seg = iter.next()) {
                if (n < max_segs) { float* o = segs + (r * max_segs + n) * 3; o[0] = seg->t0; o[1] = seg->t1; o[2] = seg->d_maj; }
                ++n;
            }
        }
        n_segs[r] = n;
    }
    return 0;
}

int vpto_sample_grid(void* ctx, int grid, const float* xyz, size_t n, float* values, uint32_t* dims, float* majorants)
{
    auto* c = static_cast<Ctx*>(ctx);
    if (!c || (grid == 1 && !c->gt)) return -1;
    S::Grid& g = grid == 1 ? *c->gt : *c->gd;
    auto acc = g.getAccessor();
    S::Sampler sampler(acc);
    S::Ray dummy(S::Vec3f(0, 0, 0), S::Vec3f(1, 0, 0));
    for (size_t i = 0; i < n; ++i) {
        S::Vec3f p(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
        if (values) values[i] = sampler(p);
        const S::Coord ijk = p.floor();
        if (dims) dims[i] = std::max<uint32_t>(8, acc.getDim(ijk, dummy));
        if (majorants) {
            float v;
            if (S::probe_leaf_max(acc, ijk, v)) majorants[i] = v;
            else if (acc.probeValue(ijk, v)) majorants[i] = v;
            else majorants[i] = 0.0f;
        }
    }
    return 0;
}

void vpto_rng(uint32_t seed, uint64_t jid, size_t n, uint64_t* state0, uint32_t* out_u32, float* out_f32)
{
    if (state0) *state0 = hash_jid(seed, jid);
    Rng<S::Engine> a(seed), b(seed);
    a.begin_job(jid); b.begin_job(jid);
    for (size_t i = 0; i < n; ++i) { if (out_u32) out_u32[i] = a.uniform_u32(); if (out_f32) out_f32[i] = b.uniform(); }
}

void vpto_camera_rays(const vptc_scene* scene, size_t n, const int64_t* px, const int64_t* py, const float* jx, const float* jy, float* origins, float* dirs)
{
    const Camera cam(scene->camera_parameters, scene->output_size[0], scene->output_size[1]);
    for (size_t i = 0; i < n; ++i) {
        V3 o, d;
        cam.generate_ray(px[i], py[i], jx[i], jy[i], o, d);
        origins[3 * i] = o.x; origins[3 * i + 1] = o.y; origins[3 * i + 2] = o.z;
        dirs[3 * i] = d.x; dirs[3 * i + 1] = d.y; dirs[3 * i + 2] = d.z;
    }
}

void vpto_blackbody_table(float* out1500)
{
    const auto& t = bb::table().xyz;
    for (unsigned i = 0; i < bb::N_BREAKPOINTS; ++i) { out1500[3 * i] = t[i].x; out1500[3 * i + 1] = t[i].y; out1500[3 * i + 2] = t[i].z; }
}
void vpto_blackbody(const float* temps, size_t n, float* out_xyz)
{
    for (size_t i = 0; i < n; ++i) { V3 v = bb::blackbody_radiation_xyz(temps[i]); out_xyz[3 * i] = v.x; out_xyz[3 * i + 1] = v.y; out_xyz[3 * i + 2] = v.z; }
}
void vpto_hg_sample(const float* w, const float* u, size_t n, float g, float* out)
{
    for (size_t i = 0; i < n; ++i) { V3 d = sample_henyey_greenstein(to_v3(w + 3 * i), u[2 * i], u[2 * i + 1], g); out[3 * i] = d.x; out[3 * i + 1] = d.y; out[3 * i + 2] = d.z; }
}
void vpto_hg_eval(const float* cos_theta, size_t n, float g, float* out) { for (size_t i = 0; i < n; ++i) out[i] = henyey_greenstein(cos_theta[i], g); }
void vpto_film_to_image(const float* film, int64_t W, int64_t H, uint8_t* rgb) { film_to_image(film, W, H, rgb); }

} // extern "C"
