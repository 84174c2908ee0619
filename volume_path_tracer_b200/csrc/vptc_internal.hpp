// vptc_internal.hpp — shared between the host and device halves of the C ABI.
#pragma once
#include <string>

namespace vptc_internal {
void set_global_error(const std::string& msg); // context-less error of the calling thread
const char* global_error();
} // namespace vptc_internal
