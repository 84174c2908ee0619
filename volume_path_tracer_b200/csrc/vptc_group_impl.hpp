// vptc_group_impl.hpp — vptc_group_*: ONE host thread driving SEVERAL GPUs, the way the reference's main() drives all of its workers
// (src/main.cpp:63-87): the baked volume replicated in every device's HBM, the (tile, wave) job space partitioned statically, the films
// combined on the first device over NVLink peer memory. Included at the end of vptc_device.cu, inside extern "C".
//
//   VPTC_SHARD_WAVES  device g of G renders waves first + g, first + g + G, ... of the whole frame; the G partial films are SUMMED by
//                     vpt_film_reduce_kernel on the first device, which reads the other devices' films directly through peer memory.
//   VPTC_SHARD_TILES  device g renders every wave of the ROWS of tiles g, g + G, g + 2G, ... (src/tile_provider.cpp:95-105 numbers tiles
//                     row-major, so a row of tiles is one contiguous band of tile_h film rows; interleaving the rows shares the expensive
//                     part of the frame evenly — contiguous halves of the 4K cloud differed by 70 % in cost); the bands are GATHERED with
//                     peer copies — disjoint pixels, 1/G of the bytes each, no arithmetic.
// Job ids are unchanged by either partition (jid = (wave-1)*ntiles + tile), so the union of the devices' RNG streams is exactly the
// single-GPU render: sample counts are equal, radiance sums differ only by fp32 add order.
//
// Calls are asynchronous per device (the range ring of vptc_device.cu): vptc_group_render_async only enqueues; the device films keep
// accumulating until vptc_group_wait combines them, adds the result into the caller's host film and clears them.

struct vptc_group {
    std::vector<vptc_ctx*> ctx;
    std::vector<float*> film;       // one accumulating device film per device
    size_t film_floats = 0;
    float* gather = nullptr;        // on device 0: staging for the devices device 0 cannot peer-map
    float* pinned = nullptr;
    std::vector<char> peer;         // device 0 can read device g's memory directly
    bool dirty = false;             // something was rendered since the last wait
    int last_sharding = VPTC_SHARD_WAVES;
    int64_t W = 0, H = 0, th = 1, nty = 0;
    std::string error;
    float last_ms = 0.0f;
};

namespace {
int gfail(vptc_group* g, int code, const std::string& msg)
{
    if (g) g->error = msg;
    vptc_internal::set_global_error(msg);
    return code;
}
#define GCU(call)                                                                                                    \
    do {                                                                                                             \
        cudaError_t e_ = (call);                                                                                     \
        if (e_ != cudaSuccess) return gfail(g, VPTC_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));   \
    } while (0)

void group_free_films(vptc_group* g)
{
    for (size_t i = 0; i < g->film.size(); ++i) {
        if (!g->film[i]) continue;
        cudaSetDevice(g->ctx[i]->device);
        cudaFree(g->film[i]);
        g->film[i] = nullptr;
    }
    if (g->gather) { cudaSetDevice(g->ctx[0]->device); cudaFree(g->gather); g->gather = nullptr; }
    if (g->pinned) { cudaFreeHost(g->pinned); g->pinned = nullptr; }
    g->film_floats = 0;
}

int group_ensure_films(vptc_group* g, size_t n)
{
    if (g->film_floats == n) return VPTC_OK;
    group_free_films(g);
    for (size_t i = 0; i < g->ctx.size(); ++i) {
        GCU(cudaSetDevice(g->ctx[i]->device));
        GCU(cudaMalloc(&g->film[i], n * sizeof(float)));
        GCU(cudaMemsetAsync(g->film[i], 0, n * sizeof(float), g->ctx[i]->stream));
    }
    GCU(cudaSetDevice(g->ctx[0]->device));
    bool all_peer = true;
    for (size_t i = 1; i < g->ctx.size(); ++i) all_peer = all_peer && g->peer[i];
    if (!all_peer) GCU(cudaMalloc(&g->gather, n * sizeof(float)));
    GCU(cudaMallocHost(&g->pinned, n * sizeof(float)));
    g->film_floats = n;
    return VPTC_OK;
}
} // namespace

int vptc_group_create(vptc_group** out, const int* devices, int n_devices)
{
    if (!out) return gfail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_create: out is null");
    *out = nullptr;
    if (!devices || n_devices < 1 || n_devices > 64) return gfail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_create: need 1..64 device indices");
    for (int i = 0; i < n_devices; ++i)
        for (int j = 0; j < i; ++j)
            if (devices[i] == devices[j]) return gfail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_create: a device is listed twice");
    std::unique_ptr<vptc_group> g(new vptc_group);
    for (int i = 0; i < n_devices; ++i) {
        vptc_ctx* c = nullptr;
        const int rc = vptc_create(&c, devices[i]);
        if (rc != VPTC_OK) {
            for (vptc_ctx* x : g->ctx) vptc_destroy(x);
            return rc; // vptc_create left the message in the context-less error slot
        }
        g->ctx.push_back(c);
    }
    g->film.assign(size_t(n_devices), nullptr);
    g->peer.assign(size_t(n_devices), 0);
    // device 0 combines the films: let it read the other devices' memory through NVLink / PCIe peer mappings where the platform allows
    cudaSetDevice(devices[0]);
    for (int i = 1; i < n_devices; ++i) {
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, devices[0], devices[i]) == cudaSuccess && can) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[i], 0);
            if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) g->peer[size_t(i)] = 1;
            cudaGetLastError();
        }
    }
    *out = g.release();
    return VPTC_OK;
}

void vptc_group_destroy(vptc_group* g)
{
    if (!g) return;
    for (vptc_ctx* c : g->ctx) { cudaSetDevice(c->device); wait_all(c); cudaStreamSynchronize(c->stream); }
    group_free_films(g);
    for (vptc_ctx* c : g->ctx) vptc_destroy(c);
    delete g;
}

const char* vptc_group_last_error(const vptc_group* g) { return g ? g->error.c_str() : vptc_internal::global_error(); }
int vptc_group_size(const vptc_group* g) { return g ? int(g->ctx.size()) : 0; }
int vptc_group_peer_mapped(const vptc_group* g) // how many of the other devices the first one reads directly (the rest go through a staging copy)
{
    int n = 0;
    if (g) for (size_t i = 1; i < g->peer.size(); ++i) n += g->peer[i] ? 1 : 0;
    return n;
}
float vptc_group_last_render_ms(const vptc_group* g) { return g ? g->last_ms : 0.0f; }

int vptc_group_set_volume(vptc_group* g, const void* density_grid, size_t density_bytes, const void* temperature_grid, size_t temperature_bytes)
{
    if (!g) return gfail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_set_volume: group is null");
    if (!density_grid) return gfail(g, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_set_volume: the density grid is required (volume_grids.cpp:48-56)");
    vpt_host::BakedGrid b, t; // baked ONCE on the host, staged to every device
    bool has_t = false;
    if (int rc = bake_volume(g->ctx[0], density_grid, density_bytes, temperature_grid, temperature_bytes, b, t, has_t)) return gfail(g, rc, g->ctx[0]->error);
    std::vector<int> rcs(g->ctx.size(), VPTC_OK);
    std::vector<std::thread> th; // one staging thread per device: the H2D copies of the replicas run side by side
    for (size_t i = 0; i < g->ctx.size(); ++i) th.emplace_back([&, i] { rcs[i] = stage_volume(g->ctx[i], b, has_t ? &t : nullptr); });
    for (auto& x : th) x.join();
    for (size_t i = 0; i < g->ctx.size(); ++i)
        if (rcs[i] != VPTC_OK) return gfail(g, rcs[i], "device " + std::to_string(g->ctx[i]->device) + ": " + g->ctx[i]->error);
    return VPTC_OK;
}

int vptc_group_render_async(vptc_group* g, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, int sharding)
{
    if (!g) return gfail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_render_async: group is null");
    if (sharding != VPTC_SHARD_WAVES && sharding != VPTC_SHARD_TILES) return gfail(g, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_render_async: sharding must be VPTC_SHARD_WAVES or VPTC_SHARD_TILES");
    if (int rc = check_scene(g->ctx[0], scene)) return gfail(g, rc, g->ctx[0]->error);
    if (first_wave < 1) return gfail(g, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_render_async: waves are 1-based (tile_provider.cpp:30)");
    const size_t n = size_t(scene->output_size[0]) * size_t(scene->output_size[1]) * 4;
    const int64_t th = scene->tile_size[1], ntx = (scene->output_size[0] + scene->tile_size[0] - 1) / scene->tile_size[0], nty = (scene->output_size[1] + th - 1) / th;
    if (g->dirty && (g->film_floats != n || g->last_sharding != sharding || g->W != scene->output_size[0] || g->th != th))
        return gfail(g, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_render_async: frame size, tile height or sharding changed before vptc_group_wait collected the previous calls");
    if (int rc = group_ensure_films(g, n)) return rc;
    const uint32_t G = uint32_t(g->ctx.size());
    g->W = scene->output_size[0]; g->H = scene->output_size[1]; g->th = th; g->last_sharding = sharding;
    g->nty = nty;
    for (uint32_t d = 0; d < G; ++d) {
        vptc_job_range jr{};
        if (sharding == VPTC_SHARD_WAVES) { // waves first + d, first + d + G, ...
            if (n_waves <= d) continue;
            jr.first_wave = first_wave + d; jr.n_waves = (n_waves - d + G - 1) / G; jr.wave_stride = G;
        } else {
            if (int64_t(d) >= nty) continue;
            const int64_t rows = (nty - int64_t(d) + int64_t(G) - 1) / int64_t(G); // rows of tiles d, d + G, ...
            jr.first_wave = first_wave; jr.n_waves = n_waves; jr.wave_stride = 1;
            jr.tile_begin = uint64_t(int64_t(d) * ntx); jr.tile_count = uint64_t(rows * ntx); jr.tile_row_stride = G;
        }
        if (int rc = enqueue_async(g->ctx[d], scene, &jr, g->film[d], nullptr)) return gfail(g, rc, "device " + std::to_string(g->ctx[d]->device) + ": " + g->ctx[d]->error);
    }
    g->dirty = true;
    return VPTC_OK;
}

int vptc_group_wait(vptc_group* g, float* film_xyzw)
{
    if (!g) return gfail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_wait: group is null");
    if (!g->dirty) return VPTC_OK;
    if (!film_xyzw) return gfail(g, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_wait: film is null");
    const size_t G = g->ctx.size(), n = g->film_floats;
    g->last_ms = 0.0f;
    for (size_t d = 0; d < G; ++d) {
        GCU(cudaSetDevice(g->ctx[d]->device));
        if (int rc = wait_all(g->ctx[d])) return gfail(g, rc, "device " + std::to_string(g->ctx[d]->device) + ": " + g->ctx[d]->error);
        GCU(cudaStreamSynchronize(g->ctx[d]->stream));
        g->last_ms = std::max(g->last_ms, g->ctx[d]->last_ms); // the devices run side by side: the slowest one is the render time
    }
    vptc_ctx* c0 = g->ctx[0];
    GCU(cudaSetDevice(c0->device));
    if (g->last_sharding == VPTC_SHARD_WAVES) {
        // SUM: device 0 reads the peer-mapped films in place; the others are staged through `gather` one at a time
        std::vector<const float*> direct;
        for (size_t d = 1; d < G; ++d) if (g->peer[d]) direct.push_back(g->film[d]);
        for (size_t i = 0; i < direct.size(); i += FILM_REDUCE_MAX_SRC) {
            const int k = int(std::min<size_t>(FILM_REDUCE_MAX_SRC, direct.size() - i));
            GCU(launch_film_reduce(g->film[0], direct.data() + i, k, n / 4, c0->stream));
        }
        for (size_t d = 1; d < G; ++d) {
            if (g->peer[d]) continue;
            GCU(cudaMemcpyPeerAsync(g->gather, c0->device, g->film[d], g->ctx[d]->device, n * sizeof(float), c0->stream));
            const float* src = g->gather;
            GCU(launch_film_reduce(g->film[0], &src, 1, n / 4, c0->stream));
        }
    } else {
        // GATHER: every band (one row of tiles = tile_h film rows) is copied from the device that rendered it into device 0's film
        // (peer copy over NVLink); nothing is added
        for (int64_t row = 0; row < g->nty; ++row) {
            const size_t d = size_t(row % int64_t(G));
            if (d == 0) continue;
            const size_t a = size_t(row * g->th) * size_t(g->W) * 4, b = size_t(std::min<int64_t>(g->H, (row + 1) * g->th)) * size_t(g->W) * 4;
            GCU(cudaMemcpyPeerAsync(g->film[0] + a, c0->device, g->film[d] + a, g->ctx[d]->device, (b - a) * sizeof(float), c0->stream));
        }
    }
    GCU(cudaMemcpyAsync(g->pinned, g->film[0], n * sizeof(float), cudaMemcpyDeviceToHost, c0->stream));
    GCU(cudaStreamSynchronize(c0->stream));
    for (size_t d = 0; d < G; ++d) { // clear the device films for the next series while the host adds
        GCU(cudaSetDevice(g->ctx[d]->device));
        GCU(cudaMemsetAsync(g->film[d], 0, n * sizeof(float), g->ctx[d]->stream));
    }
    host_add(film_xyzw, g->pinned, n);
    g->dirty = false;
    return VPTC_OK;
}

int vptc_group_render(vptc_group* g, const vptc_scene* scene, uint32_t first_wave, uint32_t n_waves, int sharding, float* film_xyzw)
{
    if (!g) return gfail(nullptr, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_render: group is null");
    if (!film_xyzw) return gfail(g, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_render: film is null");
    if (g->dirty) return gfail(g, VPTC_ERR_INVALID_ARGUMENT, "vptc_group_render: asynchronous calls are pending; collect them with vptc_group_wait first");
    if (int rc = vptc_group_render_async(g, scene, first_wave, n_waves, sharding)) return rc;
    return vptc_group_wait(g, film_xyzw);
}

vptc_ctx* vptc_group_context(vptc_group* g, int index) // the context of one device (parity hooks, counters); owned by the group
{
    return g && index >= 0 && size_t(index) < g->ctx.size() ? g->ctx[size_t(index)] : nullptr;
}
#undef GCU
