// oracle/vpt_restated_volume.inl — TEST INFRASTRUCTURE (CPU oracle), second half of vpt_restated.hpp:
// volume wrapper, majorant iterator, delta-tracking sampler, integrator, tile/wave scheduler.
#pragma once

namespace vpto {

inline V3 to_v3(const float* p) { return V3(p[0], p[1], p[2]); }

// ---------------------------------------------------------------- src/volume.cpp:11-98, include/vpt/volume.hpp:24-76
template <class S>
struct RayMajorantIterator {
    using RayT = typename S::Ray;
    using AccT = typename S::Accessor;
    struct Segment { float t0, t1, d_maj; };

    static uint32_t get_hdda_dim(const typename S::Coord& ijk, const AccT& acc, const RayT& ray)
    {
        return std::max<uint32_t>(8, acc.getDim(ijk, ray)); // volume.cpp:11-14
    }
    RayMajorantIterator(const RayT& ray, const typename S::Grid& density, const AccT& acc, Counters* cnt)
        : m_scale(1 / density.worldToIndexDirF(ray.dir()).length()), m_ray(ray),
          m_majorant(std::numeric_limits<float>::quiet_NaN()), m_acc(&acc),
          m_dda(ray, int(get_hdda_dim(ray.start().floor(), acc, ray))), m_cnt(cnt) {}

    void update_current_majorant() // volume.cpp:18-36
    {
        ++m_cnt->majorant_fetches;
        auto ijk = m_dda.voxel();
        float v;
        if (S::probe_leaf_max(*m_acc, ijk, v)) { m_majorant = v; return; }
        float value;
        if (m_acc->probeValue(ijk, value)) { m_majorant = value; return; }
        m_majorant = 0.0f;
    }
    std::optional<Segment> next() // volume.cpp:38-76
    {
        if (m_dda.time() >= m_dda.maxTime()) return std::nullopt;
        Segment ans;
        ans.t0 = m_dda.time();
        if (std::isnan(m_majorant)) update_current_majorant();
        do {
            ans.d_maj = m_majorant;
            if (!m_dda.step()) { ans.t1 = m_dda.maxTime(); return ans; }
            uint32_t new_dim = get_hdda_dim(m_ray(m_dda.time() + 1.0001f).floor(), *m_acc, m_ray);
            m_dda.update(m_ray, int(new_dim));
            update_current_majorant();
        } while (m_majorant == ans.d_maj);
        ans.t1 = m_dda.time();
        return ans;
    }
    const RayT& ray() const { return m_ray; }
    float idx_to_world_scale() const { return m_scale; }

    float m_scale;
    RayT m_ray;
    float m_majorant;
    const AccT* m_acc;
    typename S::HDDA m_dda;
    Counters* m_cnt;
};

template <class S>
struct Volume {
    // `density`/`temperature` are views of grid buffers owned by the caller; Volume's ctor mutates the
    // density buffer exactly as the reference does (volume.cpp:169): leaf maxima get the halo patch.
    Volume(typename S::Grid density, const typename S::Grid* temperature, const vptc_volume_parameters& p, bool apply_patch = true)
        : m_density(density), m_has_temp(temperature != nullptr), m_temperature(temperature ? *temperature : density), m_params(p)
    {
        if (apply_patch) S::fix_majorants_for_interpolation(m_density, 1);
    }
    std::optional<RayMajorantIterator<S>> intersect(const V3& origin, const V3& direction, const typename S::Accessor& acc, Counters* cnt) const
    { // volume.cpp:78-88
        typename S::Ray w_ray(typename S::Vec3f(origin.x, origin.y, origin.z), typename S::Vec3f(direction.x, direction.y, direction.z));
        typename S::Ray i_ray = w_ray.worldToIndexF(m_density);
        if (!i_ray.clip(m_density.indexBBox())) return std::nullopt;
        return RayMajorantIterator<S>(i_ray, m_density, acc, cnt);
    }
    typename S::Grid m_density;
    bool m_has_temp;
    typename S::Grid m_temperature;
    vptc_volume_parameters m_params;
};

// ---------------------------------------------------------------- src/majorant_transmittance_sampler.cpp:7-81
struct MediumProperties { V3 point; float sigma_maj; float density; };

template <class S>
struct MajorantTransmittanceSampler {
    MajorantTransmittanceSampler(const RayMajorantIterator<S>& it, Rng<typename S::Engine>& rng, const typename S::Grid& grid,
                                 const typename S::Accessor& acc, float sigma_t, Counters* cnt)
        : m_sigma_t(sigma_t), m_rng(rng), m_iterator(it), m_grid(grid), m_sampler(acc), m_cnt(cnt) {}

    std::optional<MediumProperties> next()
    {
        while (true) {
            if (!m_segment) {
                m_segment = m_iterator.next();
                if (!m_segment) return std::nullopt;
                if (m_segment->d_maj <= 0) { m_segment.reset(); continue; }
            }
            float sigma_maj = m_segment->d_maj * m_sigma_t;
            float dt_m = sample_exponential(m_rng.uniform(), sigma_maj);
            float t = m_segment->t0 + dt_m / m_iterator.idx_to_world_scale();
            if (t < m_segment->t1) {
                m_segment->t0 = t;
                // (m_T_maj bookkeeping of the reference is dead code: no caller reads it)
                typename S::Vec3f p_idx = m_iterator.ray()(t);
                typename S::Vec3f p_world = m_grid.indexToWorldF(p_idx);
                ++m_cnt->density_evals;
                float density = m_sampler(p_idx);
                if (density <= 0.0f) continue;
                return MediumProperties{V3(p_world[0], p_world[1], p_world[2]), sigma_maj, density};
            } else {
                m_segment.reset();
            }
        }
    }
    float m_sigma_t;
    Rng<typename S::Engine>& m_rng;
    RayMajorantIterator<S> m_iterator;
    const typename S::Grid& m_grid;
    typename S::Sampler m_sampler;
    std::optional<typename RayMajorantIterator<S>::Segment> m_segment;
    Counters* m_cnt;
};

// ---------------------------------------------------------------- src/worker.cpp:52-90
template <class S>
V3 sample_Ld(const vptc_worker_parameters& params, const Volume<S>& vol, Rng<typename S::Engine>& rng, const V3& pos, const V3& w,
             const typename S::Accessor& density_acc, Counters* cnt)
{
    V3 wi = normalized(to_v3(params.distant_light.inv_direction));
    V3 Li = to_v3(params.distant_light.xyz) * params.distant_light.multiplier;
    if (Li == V3()) return Li;
    float sigma_t = vol.m_params.sigma_a + vol.m_params.sigma_s;
    float T_ray = 1.0f;
    ++cnt->shadow_rays;
    if (auto maj_iter = vol.intersect(pos, wi, density_acc, cnt)) {
        MajorantTransmittanceSampler<S> sampler(*maj_iter, rng, vol.m_density, density_acc, sigma_t, cnt);
        while (auto props = sampler.next()) {
            float sigma_n = std::max(0.0f, props->sigma_maj - sigma_t * props->density);
            T_ray *= sigma_n / props->sigma_maj;
            if (T_ray <= 0.05f) {
                float q = 0.75f;
                if (rng.uniform() < q) T_ray = 0.0f;
                else T_ray /= 1 - q;
            }
            if (T_ray <= 0.0f) return V3();
        }
    }
    float p = henyey_greenstein(dot(w, wi), vol.m_params.henyey_greenstein_g);
    return p * T_ray * Li;
}

// One camera path: src/worker.cpp:121-200. Returns L (un-scaled radiance).
template <class S>
V3 trace_pixel(const vptc_worker_parameters& params, const Volume<S>& vol, const Camera& camera, Rng<typename S::Engine>& rng,
               int64_t px, int64_t py, const typename S::Accessor& density_acc, const typename S::Sampler* temp_sampler, Counters* cnt)
{
    float j0 = rng.uniform(), j1 = rng.uniform(); // braced-init: left to right
    const float jm = params.use_jitter ? 0.5f : 0.0f;
    V3 ro, rd;
    camera.generate_ray(px, py, j0 * jm, j1 * jm, ro, rd);
    cnt->event(0, ro.x, ro.y, ro.z, rd.x, rd.y, rd.z); // logger.new_ray(r)   worker.cpp:125

    V3 L;
    bool terminated = false;
    const vptc_volume_parameters& vp = vol.m_params;
    for (unsigned int depth = 0; depth < params.max_depth; ++depth) {
        bool scattered = false;
        auto intersection = vol.intersect(ro, rd, density_acc, cnt);
        if (!intersection) break;
        MajorantTransmittanceSampler<S> sampler(*intersection, rng, vol.m_density, density_acc, vp.sigma_a + vp.sigma_s, cnt);
        while (auto props = sampler.next()) {
            cnt->event(1, props->point.x, props->point.y, props->point.z, props->density); // logger.sampled_point   worker.cpp:146
            float p_a = (vp.sigma_a * props->density) / props->sigma_maj;
            float p_s = (vp.sigma_s * props->density) / props->sigma_maj;
            float p_n = std::max<float>(1.0f - p_a - p_s, 0.0f);
            if (temp_sampler) {
                typename S::Vec3f tc = vol.m_temperature.worldToIndexF(typename S::Vec3f(props->point.x, props->point.y, props->point.z));
                ++cnt->temperature_evals;
                float temp_adim = (*temp_sampler)(tc);
                float temp_K = temp_adim * vp.temperature_scale + vp.temperature_offset;
                L += p_a * vp.le_scale * bb::blackbody_radiation_xyz(temp_K);
            }
            ScatterEvent event = sample_discrete({{ScatterEvent::Null, p_n}, {ScatterEvent::Absorption, p_a}, {ScatterEvent::Scatter, p_s}}, rng.uniform());
            if (event == ScatterEvent::Null) {
                cnt->event(2); // logger.null()   worker.cpp:166
                continue;
            } else if (event == ScatterEvent::Scatter) {
                if (depth++ >= params.max_depth) { cnt->event(5); terminated = true; break; } // logger.scatter_terminated()   :170
                ++cnt->scatter_events;
                L += sample_Ld(params, vol, rng, props->point, rd, density_acc, cnt);
                float u0 = rng.uniform(), u1 = rng.uniform();
                V3 new_dir = sample_henyey_greenstein(rd, u0, u1, vp.henyey_greenstein_g);
                ro = props->point; rd = new_dir;
                cnt->event(3, ro.x, ro.y, ro.z, rd.x, rd.y, rd.z); // logger.scatter(r)   worker.cpp:181
                scattered = true;
                break;
            } else {
                cnt->event(4); // logger.absorbed()   worker.cpp:185
                terminated = true;
                break;
            }
        }
        if (!scattered) break;
    }
    if (!terminated) L += to_v3(params.infinite_light.xyz) * params.infinite_light.multiplier;
    ++cnt->pixel_samples;
    return L;
}

// ---------------------------------------------------------------- src/tile_provider.cpp:14-105
struct TileGrid {
    int64_t W, H, tw, th, ntx, nty;
    explicit TileGrid(const vptc_scene& s) : W(s.output_size[0]), H(s.output_size[1]), tw(s.tile_size[0]), th(s.tile_size[1])
    {
        ntx = W / tw + (W % tw != 0); nty = H / th + (H % th != 0);
    }
    uint64_t ntiles() const { return uint64_t(ntx) * uint64_t(nty); }
    void rect(uint64_t tile, int64_t& x0, int64_t& y0, int64_t& sx, int64_t& sy) const
    {
        x0 = int64_t(tile % uint64_t(ntx)) * tw; y0 = int64_t(tile / uint64_t(ntx)) * th;
        sx = std::min(W - x0, tw); sy = std::min(H - y0, th);
    }
};

// One job = one (tile, wave) token: src/worker.cpp:104-206. `film` may be null (trace mode); per-pixel
// outputs (tile_w*tile_h slots) are optional.
template <class S>
void run_job(const vptc_scene& scene, const Volume<S>& vol, const Camera& camera, const TileGrid& tg, uint64_t jid,
             Rng<typename S::Engine>& rng, const typename S::Accessor& density_acc, const typename S::Sampler* temp_sampler,
             float* film, float* per_pixel_xyz, uint32_t* per_pixel_ndraws, Counters* cnt)
{
    const uint64_t tile = jid % tg.ntiles();
    int64_t x0, y0, sx, sy;
    tg.rect(tile, x0, y0, sx, sy);
    rng.begin_job(jid);
    const uint64_t draws0 = rng.draws;
    const vptc_worker_parameters& wp = scene.worker_parameters;
    for (int64_t y = 0; y < sy; ++y) {
        for (int64_t x = 0; x < sx; ++x) {
            const int64_t px = x0 + x, py = y0 + y;
            if (wp.single_pixel_enabled && (wp.single_pixel_coord[0] != px || wp.single_pixel_coord[1] != py)) continue;
            V3 L = trace_pixel<S>(wp, vol, camera, rng, px, py, density_acc, temp_sampler, cnt);
            if (film) {
                float* f = film + (size_t(py) * size_t(tg.W) + size_t(px)) * 4;
                f[3] += 1.0f;
                V3 s = scene.camera_parameters.imaging_ratio * L;
                f[0] += s.x; f[1] += s.y; f[2] += s.z;
            }
            const size_t slot = size_t(y) * size_t(tg.tw) + size_t(x);
            if (per_pixel_xyz) { per_pixel_xyz[slot * 3 + 0] = L.x; per_pixel_xyz[slot * 3 + 1] = L.y; per_pixel_xyz[slot * 3 + 2] = L.z; }
            if (per_pixel_ndraws) per_pixel_ndraws[slot] = uint32_t(rng.draws - draws0);
        }
    }
    cnt->rng_draws += rng.draws - draws0;
}

// Multi-threaded render of waves [first_wave, first_wave+n_waves): the reference's worker loop
// (main.cpp:63-87) with TileProvider's ordering guarantee — a tile's waves are processed in order and
// never concurrently (tile_provider.cpp:40-60) — so the film is bit-deterministic.
// `job_stride` > 1 renders only every job_stride-th job of the range (a bounded, frame-covering SAMPLE of the
// workload for the timed CPU baseline); pick it coprime to ntiles so successive waves visit different tiles.
template <class S>
double render(const vptc_scene& scene, const Volume<S>& vol, uint32_t first_wave, uint32_t n_waves, unsigned num_threads, float* film,
              Counters* total, uint64_t job_stride = 1)
{
    const TileGrid tg(scene);
    const Camera camera(scene.camera_parameters, tg.W, tg.H);
    const uint64_t ntiles = tg.ntiles(), njobs_all = ntiles * n_waves, job0 = uint64_t(first_wave - 1) * ntiles;
    const uint64_t njobs = (njobs_all + job_stride - 1) / job_stride;
    std::atomic<uint64_t> job_idx{0};
    std::vector<std::atomic<uint32_t>> tile_wave(ntiles);
    for (auto& t : tile_wave) t.store(first_wave - 1);
    std::vector<Counters> cnts(num_threads);
    const auto t_start = std::chrono::steady_clock::now();
    auto work = [&](unsigned tid) {
        Rng<typename S::Engine> rng(scene.seed);
        auto density_acc = vol.m_density.getAccessor();
        auto temp_acc = vol.m_temperature.getAccessor();
        typename S::Sampler temp_sampler(temp_acc);
        for (;;) {
            const uint64_t j = job_idx.fetch_add(1, std::memory_order_relaxed);
            if (j >= njobs) break;
            const uint64_t jid = job0 + j * job_stride, tile = jid % ntiles;
            const uint32_t wave = uint32_t(1 + jid / ntiles);
            // per-tile wave ordering (tile_provider.cpp:40-60); with a job stride a tile may skip waves, so wait
            // only until no EARLIER wave of this tile is still in flight (the counter holds the last finished one)
            if (job_stride == 1) while (tile_wave[tile].load(std::memory_order_acquire) != wave - 1) std::this_thread::yield();
            run_job<S>(scene, vol, camera, tg, jid, rng, density_acc, vol.m_has_temp ? &temp_sampler : nullptr, film, nullptr, nullptr, &cnts[tid]);
            tile_wave[tile].store(wave, std::memory_order_release);
        }
    };
    std::vector<std::thread> threads;
    for (unsigned i = 0; i < num_threads; ++i) threads.emplace_back(work, i);
    for (auto& t : threads) t.join();
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_start).count();
    if (total) for (const auto& c : cnts) total->add(c);
    return secs;
}

// ---------------------------------------------------------------- src/main.cpp:12-24, include/vpt/color.hpp:8-30
inline void film_to_image(const float* film, int64_t W, int64_t H, uint8_t* rgb)
{
    auto srgb = [](float x) { return x <= 0.0031308f ? (12.92f * x) : (1.055f * std::pow(x, 1.0f / 2.4f) - 0.055f); };
    static const float m[3][3] = {{float(3.240479), float(-1.537150), float(-0.498535)},
                                  {float(-0.969256), float(1.875991), float(0.041556)},
                                  {float(0.055648), float(-0.204043), float(1.057311)}};
    for (int64_t i = 0; i < W * H; ++i) {
        const float* f = film + i * 4;
        V3 xyz = V3(f[0], f[1], f[2]) / f[3];
        float lin[3];
        for (int r = 0; r < 3; ++r) lin[r] = sum3(m[r][0] * xyz.x, m[r][1] * xyz.y, m[r][2] * xyz.z);
        for (int r = 0; r < 3; ++r) {
            float s = srgb(lin[r]);
            s = std::min(std::max(s, 0.0f), 1.0f) * 255.0f; // cwiseMax(0).cwiseMin(1) * 255; NaN -> max(NaN,0) per std::max = NaN... see note
            rgb[i * 3 + r] = (s == s) ? static_cast<unsigned char>(s) : 0; // NaN (w = 0) cast is implementation-defined in the reference; 0 here
        }
    }
}

} // namespace vpto
