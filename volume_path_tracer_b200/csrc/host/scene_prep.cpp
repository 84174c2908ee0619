// scene_prep.cpp — host-side, once-per-render constants the kernels read (device/layout.hpp SceneDev):
// the camera's raster->world affine (src/camera.cpp:5-57), the light constants (src/worker.cpp:54-55,199),
// the tile grid (src/tile_provider.cpp:14-31) and the blackbody table (src/precompute_blackbody.cpp:7-22,
// src/spectral.cpp:7-20, include/vpt/spectral.hpp:62-75). Computed with the same fp32 operation order the
// reference's Eigen expressions produce (3-term reductions associate a + (b + c)); compile with
// -ffp-contract=off like the reference's no-FMA Release build.
#include "scene_prep.hpp"

#include <cmath>
#include <numbers>

namespace vpt_host {
namespace {
struct V { float x, y, z; };
inline float sum3(float a, float b, float c) { return a + (b + c); }
inline V sub(V a, V b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V norm(V a) { const float n = std::sqrt(sum3(a.x * a.x, a.y * a.y, a.z * a.z)); return {a.x / n, a.y / n, a.z / n}; }
inline V cross(V a, V b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }

const float kCIE[471][3] = {
#include "cie1931_xyz_1nm.inc"
};
} // namespace

const float* cie_table() { return &kCIE[0][0]; }
void lambda5_table(float out[471])
{
    // planck_law's `float lambda5 = std::pow(lambda_m, 5)` (spectral.cpp:15): the (float, int) overload evaluates in double
    for (unsigned int k = 0; k < 471; ++k) out[k] = float(std::pow(static_cast<float>(360 + k) * 1e-9f, 5));
}

void prepare_scene(const vptc_scene& s, bool has_temperature, vptd::SceneDev& d)
{
    const int64_t W = s.output_size[0], H = s.output_size[1];
    const vptc_camera_parameters& p = s.camera_parameters;
    // ---- Camera::Camera ----
    const float ar = static_cast<float>(W) / static_cast<float>(H);
    const float vfov_rad = float(std::numbers::pi) * p.vfov_deg / 180.0f;
    const V pos{p.position[0], p.position[1], p.position[2]}, look{p.look[0], p.look[1], p.look[2]}, up{p.up[0], p.up[1], p.up[2]};
    const V dir = norm(sub(look, pos));
    const V left = cross(norm(up), dir); // not re-normalised, as in the reference
    const V new_up = cross(dir, left);
    const float c2w[3][3] = {{left.x, new_up.x, dir.x}, {left.y, new_up.y, dir.y}, {left.z, new_up.z, dir.z}};
    const float tan_vfov2 = std::tan(vfov_rad / 2);
    // screen_to_camera: diag(ar*tan, tan, 0), translation (0,0,1); raster_to_screen: diag(-2/W, -2/H, 0), translation (1,1,0)
    const float s2c_d[3] = {ar * tan_vfov2, tan_vfov2, 0.0f};
    const float r2s_d[3] = {-(1.0f / (static_cast<float>(W) / 2.0f)), -(1.0f / (static_cast<float>(H) / 2.0f)), 0.0f};
    const float r2s_t[3] = {1.0f, 1.0f, 0.0f};
    float s2w_m[3][3], s2w_t[3];
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
            float term[3];
            for (int k = 0; k < 3; ++k) term[k] = c2w[r][k] * (k == c ? s2c_d[c] : 0.0f);
            s2w_m[r][c] = sum3(term[0], term[1], term[2]);
        }
        s2w_t[r] = sum3(c2w[r][0] * 0.0f, c2w[r][1] * 0.0f, c2w[r][2] * 1.0f) + 0.0f;
    }
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) {
            float term[3];
            for (int k = 0; k < 3; ++k) term[k] = s2w_m[r][k] * (k == c ? r2s_d[c] : 0.0f);
            d.cam_m[r * 3 + c] = sum3(term[0], term[1], term[2]);
        }
        d.cam_t[r] = sum3(s2w_m[r][0] * r2s_t[0], s2w_m[r][1] * r2s_t[1], s2w_m[r][2] * r2s_t[2]) + s2w_t[r];
    }
    d.cam_pos[0] = p.position[0]; d.cam_pos[1] = p.position[1]; d.cam_pos[2] = p.position[2];
    d.imaging_ratio = p.imaging_ratio;
    // ---- lights ----
    const vptc_worker_parameters& w = s.worker_parameters;
    for (int i = 0; i < 3; ++i) {
        d.inf_light[i] = w.infinite_light.xyz[i] * w.infinite_light.multiplier;
        d.dist_Li[i] = w.distant_light.xyz[i] * w.distant_light.multiplier;
    }
    const V wi = norm(V{w.distant_light.inv_direction[0], w.distant_light.inv_direction[1], w.distant_light.inv_direction[2]});
    d.dist_wi[0] = wi.x; d.dist_wi[1] = wi.y; d.dist_wi[2] = wi.z;
    d.dist_light_on = !(d.dist_Li[0] == 0.0f && d.dist_Li[1] == 0.0f && d.dist_Li[2] == 0.0f);
    // ---- volume parameters ----
    const vptc_volume_parameters& v = s.volume_parameters;
    d.sigma_a = v.sigma_a; d.sigma_s = v.sigma_s; d.sigma_t = v.sigma_a + v.sigma_s;
    d.g = v.henyey_greenstein_g; d.le_scale = v.le_scale;
    d.temperature_offset = v.temperature_offset; d.temperature_scale = v.temperature_scale;
    d.max_depth = w.max_depth;
    d.use_jitter = w.use_jitter ? 1 : 0;
    d.has_temperature = has_temperature ? 1 : 0;
    d.single_pixel_enabled = w.single_pixel_enabled ? 1 : 0;
    d.single_pixel_x = w.single_pixel_coord[0]; d.single_pixel_y = w.single_pixel_coord[1];
    // ---- TileProvider ----
    d.W = W; d.H = H; d.tw = s.tile_size[0]; d.th = s.tile_size[1];
    d.ntx = W / d.tw + (W % d.tw != 0); d.nty = H / d.th + (H % d.th != 0);
    d.ntiles = uint64_t(d.ntx) * uint64_t(d.nty);
    d.seed = s.seed;
}

// ---- init_blackbody_radiation_xyz ----
static float planck_law(float lambda_m, float temperature_k)
{
    if (temperature_k <= 0.0f) return 0.0f;
    constexpr float c = 299792458.f;
    constexpr float h = 6.62606957e-34f;
    constexpr float kb = 1.3806488e-23f;
    constexpr float num = 2 * h * c * c;
    float lambda5 = std::pow(lambda_m, 5); // (float, int) overload: computed in double
    float exp = std::exp((h * c) / (lambda_m * kb * temperature_k));
    float den = lambda5 * (exp - 1);
    return num / den;
}

void compute_blackbody_table(float out[1500])
{
    const float Y_integral = VPT_CIE_Y_INTEGRAL;
    for (int i = 0; i < 500; ++i) {
        const float T = (i - 1) * (50000.0f / 500); // idx_to_temp
        for (int ch = 0; ch < 3; ++ch) {
            float integral = 0.0f;
            for (unsigned int k = 0; k < 471; ++k) integral += kCIE[k][ch] * planck_law(static_cast<float>(360 + k) * 1e-9f, T);
            out[i * 3 + ch] = integral / Y_integral;
        }
    }
}

} // namespace vpt_host
