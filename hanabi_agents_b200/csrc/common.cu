#include "common.cuh"
#include <string.h>

namespace hb {
static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace hb

extern "C" {
const char* hb_last_error(void) { return hb::g_err; }
int hb_version(void) { return 100; }
const char* hb_build_arch(void) { return "sm_100a"; }
}
