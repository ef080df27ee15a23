// Library-level entry points: version, error string, launch accounting.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace b200 {

static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

}  // namespace b200

extern "C" int b200_version(void) { return 100; /* 0.1.0 */ }

extern "C" const char* b200_last_error(void) { return b200::t_err; }

extern "C" int b200_compiled_arch(void) {
#ifdef B200_ARCH
  return B200_ARCH;
#else
  return 100;
#endif
}

extern "C" uint64_t b200_launch_count(void) {
  return b200::g_launches.load(std::memory_order_relaxed);
}
