#include "rt_common.h"

#include <stdarg.h>
#include <stdio.h>

namespace cb2rt {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }
}  // namespace cb2rt

extern "C" const char* cb2rt_last_error(void) { return cb2rt::last_error(); }
