// Shared helpers of the cloud_b200 C++ runtime library (libcloud_b200_rt.so).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace cb2rt {
// printf-style, thread-local; read back through cb2rt_last_error().
void set_error(const char* fmt, ...);
}  // namespace cb2rt

extern "C" const char* cb2rt_last_error(void);
