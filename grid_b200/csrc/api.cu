// Error state, version and launch counter of libgridb200.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace g2 {
static thread_local char t_error[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_error, sizeof(t_error), fmt, ap);
  va_end(ap);
}
}  // namespace g2

extern "C" const char* gridb200_last_error(void) { return g2::t_error; }
extern "C" const char* gridb200_version(void) { return "gridb200 0.1 (sm_100a)"; }
extern "C" uint64_t gridb200_launch_count(void) { return g2::g_launches.load(); }
