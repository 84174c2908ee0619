// scene_prep.hpp — host-precomputed constants for the kernels (see scene_prep.cpp).
#pragma once
#include "../device/layout.hpp"
#include "vptc.h"

namespace vpt_host {
void prepare_scene(const vptc_scene& scene, bool has_temperature, vptd::SceneDev& out);
void compute_blackbody_table(float out[1500]);
const float* cie_table(); // 471 x {X, Y, Z}
void lambda5_table(float out[471]); // lambda_m^5 per wavelength as the host's libm rounds it
} // namespace vpt_host
