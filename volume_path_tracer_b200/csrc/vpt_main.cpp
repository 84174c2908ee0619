// vpt_main.cpp — the headless `vpt scene.json out.png` CLI on top of the C ABI (include/vptc.h).
//
// Keeps the reference's command line (src/main.cpp:26-30), its all-keys-mandatory JSON scene schema
// (src/configuration.cpp:8-22), `volume_path` resolved against the scene file's directory (main.cpp:34,40),
// the density-required / temperature-optional grid contract (src/volume_grids.cpp:36-67) and its
// "[FATAL]: ... ; exit(1)" error convention (include/vpt/logging.hpp:16). The jthread spawn + raylib
// preview loop (main.cpp:63-134) is replaced by vptc_render calls; there is no window on this box, so the
// image is saved when the last wave has finished instead of when the window closes (main.cpp:136-143).
// Optional environment: VPT_DEVICE (CUDA device index, default 0), VPT_WAVES_PER_CALL (progress granularity, default 64).
// Progressive semantics of the preview window, headless:
//   VPT_SAVE_EVERY=K   rewrite the output image every K waves (the window's periodic redraw, main.cpp:107-131)
//   VPT_CHECKPOINT=F   keep the float4 film + next wave in F (written whenever the image is, and at the end); when F
//                      exists and matches the scene the render RESUMES from it — num_waves may have been raised since
//   VPT_TRACE_PIXEL=x,y  also write ./majorant_trace.csv for that pixel's centre ray (Volume::log_majorant_trace, volume.cpp:176-192:
//                      the ray_visualizer's CSV for scripts/plot_majorant_trace.py), from the device's own segment iterator, and ./log.csv,
//                      the path event log of worker.cpp's Logger<true> in single_pixel mode (scripts/plot_raytrace.py), for wave 1
//   SIGINT / SIGTERM   = closing the window: finish the waves in flight, save image (+ checkpoint), exit 0
//                      (TileProvider::stop_at_next_wave, tile_provider.cpp:107-110; main.cpp:133-143)
#include <algorithm>
#include <chrono>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <filesystem>
#include <iostream>
#include <string>
#include <vector>

#include "vptc.h"

#define vptFATAL(x) do { std::clog << "[FATAL]: " << x << "\n"; std::exit(1); } while (false)
#define vptWARN(x) do { std::clog << "[WARN]: " << x << "\n"; } while (false)
#define vptINFO(x) do { std::clog << "[INFO]: " << x << "\n"; } while (false)

static volatile std::sig_atomic_t g_stop = 0;
static void on_signal(int) { g_stop = 1; }

int main(int argc, char* argv[])
{
    if (argc != 3) vptFATAL("Usage: " << argv[0] << " config_path output_path");

    std::error_code ec;
    const std::filesystem::path config_path = std::filesystem::canonical(argv[1], ec);
    if (ec) vptFATAL("Failed to read configuration file \"" << argv[1] << "\": " << ec.message());
    const std::filesystem::path output_path(argv[2]);

    vptc_scene cfg;
    char volume_rel[4096];
    if (vptc_scene_from_json_file(config_path.c_str(), &cfg, volume_rel, sizeof(volume_rel)) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));

    // VolumeGrids::read_from_file
    const std::filesystem::path volume_path = config_path.parent_path() / volume_rel;
    void *density = nullptr, *temperature = nullptr;
    size_t density_bytes = 0, temperature_bytes = 0;
    if (vptc_nvdb_read_grid(volume_path.c_str(), "density", &density, &density_bytes) != VPTC_OK) {
        vptWARN("NanoVDB failed to read grid \"density\" from file " << volume_path << ": " << vptc_last_error(nullptr));
        vptFATAL("NanoVDB file " << volume_path << " does not contain the \"density\" grid.");
    }
    if (vptc_nvdb_read_grid(volume_path.c_str(), "temperature", &temperature, &temperature_bytes) != VPTC_OK) {
        vptWARN("NanoVDB failed to read grid \"temperature\" from file " << volume_path << ": " << vptc_last_error(nullptr));
        temperature = nullptr; temperature_bytes = 0;
    }

    if (const char* tp = std::getenv("VPT_TRACE_PIXEL")) { // dda_trace.csv needs the NanoVDB buffer itself (masks, node statistics): before it is freed
        long long tx = -1, ty = -1;
        size_t rows = 0;
        if (std::sscanf(tp, "%lld,%lld", &tx, &ty) != 2) vptFATAL("VPT_TRACE_PIXEL must be \"x,y\"");
        if (vptc_dda_trace_csv(density, density_bytes, &cfg, tx, ty, "dda_trace.csv", &rows) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));
        vptINFO("Wrote dda_trace.csv for pixel (" << tx << ", " << ty << "): " << rows << " voxels");
    }
    // Devices. VPT_DEVICES=0,1,2,... renders on several GPUs behind one vptc_group (the reference drives all its workers from main(),
    // src/main.cpp:63-87): volume replicated, job space sharded by waves (VPT_SHARD=waves, default) or by stripes of tile rows
    // (VPT_SHARD=tiles), films combined on the first device. VPT_DEVICE=n (or nothing) renders on one GPU.
    std::vector<int> devices;
    if (const char* ds = std::getenv("VPT_DEVICES")) {
        for (const char* p = ds; *p;) {
            char* end = nullptr;
            const long v = std::strtol(p, &end, 10);
            if (end == p) vptFATAL("VPT_DEVICES must be a comma-separated list of device indices");
            devices.push_back(int(v));
            p = *end == ',' ? end + 1 : end;
            if (*end && *end != ',') vptFATAL("VPT_DEVICES must be a comma-separated list of device indices");
        }
    }
    const char* shard_env = std::getenv("VPT_SHARD");
    const int sharding = shard_env && std::string(shard_env) == "tiles" ? VPTC_SHARD_TILES : VPTC_SHARD_WAVES;
    if (shard_env && std::string(shard_env) != "tiles" && std::string(shard_env) != "waves") vptFATAL("VPT_SHARD must be \"waves\" or \"tiles\"");
    const char* dev_env = std::getenv("VPT_DEVICE");
    vptc_group* group = nullptr;
    vptc_ctx* ctx = nullptr;
    if (devices.size() > 1) {
        if (vptc_group_create(&group, devices.data(), int(devices.size())) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));
        if (vptc_group_set_volume(group, density, density_bytes, temperature, temperature_bytes) != VPTC_OK) vptFATAL(vptc_group_last_error(group));
        ctx = vptc_group_context(group, 0); // trace hooks and the resolve run on the first device
        vptINFO("Rendering on " << devices.size() << " devices (" << (sharding == VPTC_SHARD_TILES ? "tile stripes, gathered" : "waves, summed") << "; "
                << vptc_group_peer_mapped(group) << " of " << devices.size() - 1 << " peers mapped directly)");
    } else {
        if (vptc_create(&ctx, devices.size() == 1 ? devices[0] : (dev_env ? std::atoi(dev_env) : 0)) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));
        if (vptc_set_volume(ctx, density, density_bytes, temperature, temperature_bytes) != VPTC_OK) vptFATAL(vptc_last_error(ctx));
    }
    vptc_free(density); vptc_free(temperature); // the library baked its own copy
    if (const char* tp = std::getenv("VPT_TRACE_PIXEL")) {
        long long tx = -1, ty = -1;
        if (std::sscanf(tp, "%lld,%lld", &tx, &ty) != 2) vptFATAL("VPT_TRACE_PIXEL must be \"x,y\"");
        if (vptc_trace_pixel_csv(ctx, &cfg, tx, ty, "majorant_trace.csv") != VPTC_OK) vptFATAL(vptc_last_error(ctx));
        vptINFO("Wrote majorant_trace.csv for pixel (" << tx << ", " << ty << ")");
        // Logger<true> + single_pixel (worker.cpp:15-49,113-119): log.csv of the path wave 1 traces through that pixel alone
        const int64_t ntx = cfg.output_size[0] / cfg.tile_size[0] + (cfg.output_size[0] % cfg.tile_size[0] != 0);
        const uint64_t jid = uint64_t(ty / cfg.tile_size[1]) * uint64_t(ntx) + uint64_t(tx / cfg.tile_size[0]);
        std::vector<float> ev(size_t(1) << 20);
        size_t n_ev = 0;
        if (vptc_trace_path_log(ctx, &cfg, jid, tx, ty, ev.data(), ev.size() / 8, &n_ev) != VPTC_OK) vptFATAL(vptc_last_error(ctx));
        if (FILE* f = std::fopen("log.csv", "w")) {
            static const char* names[] = {"new_ray", "sampled_point", "null", "scatter", "absorbed", "scatter_terminated"};
            for (size_t i = 0; i < std::min(n_ev, ev.size() / 8); ++i) {
                const float* e = &ev[i * 8];
                const int t = std::min(5, std::max(0, int(e[0])));
                if (t == 0 || t == 3) std::fprintf(f, "%s,%.9g,%.9g,%.9g,%.9g,%.9g,%.9g\n", names[t], e[1], e[2], e[3], e[4], e[5], e[6]);
                else if (t == 1) std::fprintf(f, "%s,%.9g,%.9g,%.9g,%.9g\n", names[t], e[1], e[2], e[3], e[4]);
                else std::fprintf(f, "%s\n", names[t]);
            }
            std::fclose(f);
            vptINFO("Wrote log.csv (" << n_ev << " path events)");
        } else vptFATAL("cannot write log.csv");
    }

    const int64_t W = cfg.output_size[0], H = cfg.output_size[1];
    std::vector<float> film(size_t(W) * size_t(H) * 4, 0.0f);
    const char* wpc_env = std::getenv("VPT_WAVES_PER_CALL");
    const uint32_t per_call = std::max(1, wpc_env ? std::atoi(wpc_env) : 64);

    const char* every_env = std::getenv("VPT_SAVE_EVERY");
    const uint32_t save_every = every_env ? uint32_t(std::max(0, std::atoi(every_env))) : 0u;
    const char* ckpt = std::getenv("VPT_CHECKPOINT");
    uint32_t first_wave = 1;
    if (ckpt && std::filesystem::exists(ckpt)) {
        uint32_t next = 1;
        if (vptc_film_load(ckpt, &cfg, film.data(), &next) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));
        first_wave = next;
        vptINFO("Resuming from checkpoint " << ckpt << " at wave " << first_wave);
    }
    std::signal(SIGINT, on_signal);
    std::signal(SIGTERM, on_signal);

    std::vector<uint8_t> img(size_t(W) * size_t(H) * 3);
    auto save = [&](uint32_t next_wave) {
        if (vptc_film_to_image(ctx, film.data(), W, H, img.data()) != VPTC_OK) vptFATAL(vptc_last_error(ctx));
        if (vptc_write_png(output_path.c_str(), img.data(), W, H) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));
        if (ckpt && vptc_film_save(ckpt, &cfg, film.data(), next_wave) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));
    };

    // The render loop. Calls are ASYNCHRONOUS (vptc_render_async): call k+1 is handed to the GPU before call k's film is brought home, and the
    // running kernel carries on into it, so the device never drains between calls (the reference's workers never stop between waves either,
    // tile_provider.cpp:27-67). When vptc_render_async(k+1) returns, `film` holds every wave of the calls up to k: that is the state the
    // periodic save writes (PNG + checkpoint with next_wave = first wave of call k+1) — no pipeline flush for a save.
    const auto start = std::chrono::high_resolution_clock::now();
    double device_ms = 0.0;
    uint32_t wave = first_wave;      // first wave not yet handed to the GPU
    uint32_t in_film = first_wave;   // first wave not yet accumulated in `film`
    while (wave <= cfg.num_waves && !g_stop) {
        uint32_t n = std::min(per_call, cfg.num_waves - wave + 1);
        if (save_every) n = std::min(n, save_every - (wave - 1) % save_every); // calls end on multiples of K
        vptINFO("Starting wave " << wave << "/" << cfg.num_waves);
        if (group) { // the devices' films keep accumulating on the devices; they are combined where the film is needed (a save, the end)
            if (vptc_group_render_async(group, &cfg, wave, n, sharding) != VPTC_OK) vptFATAL(vptc_group_last_error(group));
            wave += n;
            if (save_every && (wave - 1) % save_every == 0 && wave <= cfg.num_waves) {
                if (vptc_group_wait(group, film.data()) != VPTC_OK) vptFATAL(vptc_group_last_error(group));
                device_ms += vptc_group_last_render_ms(group);
                save(wave);
            }
            continue;
        }
        if (vptc_render_async(ctx, &cfg, wave, n, film.data()) != VPTC_OK) vptFATAL(vptc_last_error(ctx));
        in_film = wave; // the previous calls are in `film` now; this one is on the device
        wave += n;
        if (save_every && in_film > first_wave && (in_film - 1) % save_every == 0) save(in_film);
    }
    if (group) {
        if (vptc_group_wait(group, film.data()) != VPTC_OK) vptFATAL(vptc_group_last_error(group));
        device_ms += vptc_group_last_render_ms(group);
    } else {
        if (vptc_wait(ctx) != VPTC_OK) vptFATAL(vptc_last_error(ctx));
        device_ms = vptc_last_render_ms(ctx);
    }
    if (g_stop && wave <= cfg.num_waves) vptINFO("Stopped at wave " << wave - 1 << "/" << cfg.num_waves);
    const auto elapsed = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - start);
    const double rendered = double(W) * double(H) * double(wave - first_wave);
    vptINFO("Rendering complete in " << elapsed.count() << " ms (" << device_ms << " ms on the device, "
            << (device_ms > 0.0 ? rendered / (device_ms * 1e-3) / 1e6 : 0.0) << " M pixel-samples/s)");
    save(wave);
    if (group) vptc_group_destroy(group);
    else vptc_destroy(ctx);
    return 0;
}
