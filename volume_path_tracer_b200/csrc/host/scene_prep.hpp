// scene_prep.hpp — host-precomputed constants for the kernels (see scene_prep.cpp).
#pragma once
#include "../device/layout.hpp"
#include "vptc.h"

namespace vpt_host {
void prepare_scene(const vptc_scene& scene, bool has_temperature, vptd::SceneDev& out);
void compute_blackbody_table(float out[1500]);
const float* cie_table(); // 471 x {X, Y, Z}
} // namespace vpt_host
