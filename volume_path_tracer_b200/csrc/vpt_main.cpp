// vpt_main.cpp — the headless `vpt scene.json out.png` CLI on top of the C ABI (include/vptc.h).
//
// Keeps the reference's command line (src/main.cpp:26-30), its all-keys-mandatory JSON scene schema
// (src/configuration.cpp:8-22), `volume_path` resolved against the scene file's directory (main.cpp:34,40),
// the density-required / temperature-optional grid contract (src/volume_grids.cpp:36-67) and its
// "[FATAL]: ... ; exit(1)" error convention (include/vpt/logging.hpp:16). The jthread spawn + raylib
// preview loop (main.cpp:63-134) is replaced by vptc_render calls; there is no window on this box, so the
// image is saved when the last wave has finished instead of when the window closes (main.cpp:136-143).
// Optional environment: VPT_DEVICE (CUDA device index, default 0), VPT_WAVES_PER_CALL (progress granularity).
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <filesystem>
#include <iostream>
#include <string>
#include <vector>

#include "vptc.h"

#define vptFATAL(x) do { std::clog << "[FATAL]: " << x << "\n"; std::exit(1); } while (false)
#define vptWARN(x) do { std::clog << "[WARN]: " << x << "\n"; } while (false)
#define vptINFO(x) do { std::clog << "[INFO]: " << x << "\n"; } while (false)

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

    const char* dev_env = std::getenv("VPT_DEVICE");
    vptc_ctx* ctx = nullptr;
    if (vptc_create(&ctx, dev_env ? std::atoi(dev_env) : 0) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));
    if (vptc_set_volume(ctx, density, density_bytes, temperature, temperature_bytes) != VPTC_OK) vptFATAL(vptc_last_error(ctx));
    vptc_free(density); vptc_free(temperature); // the library baked its own copy

    const int64_t W = cfg.output_size[0], H = cfg.output_size[1];
    std::vector<float> film(size_t(W) * size_t(H) * 4, 0.0f);
    const char* wpc_env = std::getenv("VPT_WAVES_PER_CALL");
    const uint32_t per_call = std::max(1, wpc_env ? std::atoi(wpc_env) : 64);

    const auto start = std::chrono::high_resolution_clock::now();
    double device_ms = 0.0;
    for (uint32_t wave = 1; wave <= cfg.num_waves; wave += per_call) {
        const uint32_t n = std::min(per_call, cfg.num_waves - wave + 1);
        vptINFO("Starting wave " << wave << "/" << cfg.num_waves);
        if (vptc_render(ctx, &cfg, wave, n, film.data()) != VPTC_OK) vptFATAL(vptc_last_error(ctx));
        device_ms += vptc_last_render_ms(ctx);
    }
    const auto elapsed = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::high_resolution_clock::now() - start);
    vptINFO("Rendering complete in " << elapsed.count() << " ms (" << device_ms << " ms on the device, "
            << double(W) * double(H) * double(cfg.num_waves) / (device_ms * 1e-3) / 1e6 << " M pixel-samples/s)");

    std::vector<uint8_t> img(size_t(W) * size_t(H) * 3);
    if (vptc_film_to_image(ctx, film.data(), W, H, img.data()) != VPTC_OK) vptFATAL(vptc_last_error(ctx));
    if (vptc_write_png(output_path.c_str(), img.data(), W, H) != VPTC_OK) vptFATAL(vptc_last_error(nullptr));
    vptc_destroy(ctx);
    return 0;
}
