// json_scene.hpp — reader for the reference's scene JSON schema (see json_scene.cpp).
#pragma once
#include <string>

#include "vptc.h"

namespace vpt_host {
bool read_text_file(const std::string& path, std::string& out, std::string& err);
// All keys mandatory, unknown keys rejected (glaze error_on_missing_keys / error_on_unknown_keys,
// src/configuration.cpp:13-15). Returns false and fills err on failure.
bool parse_scene_json(const std::string& text, vptc_scene& out, std::string& volume_path, std::string& err);
} // namespace vpt_host
