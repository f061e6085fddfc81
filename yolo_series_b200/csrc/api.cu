// Error plumbing + ABI version for libyolo_b200.so.
#include <stdarg.h>

#include "common.cuh"

namespace yb {
static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace yb

extern "C" const char* yb_last_error(void) { return yb::g_err; }
extern "C" int yb_abi_version(void) { return YB_ABI_VERSION; }
