// Error string + version entry points of the C-ABI.
#include <stdarg.h>
#include <string.h>

#include "at_common.cuh"

#include <atomic>

namespace {
thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};
}

namespace abt {
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
}  // namespace abt

extern "C" int64_t at_launch_count(void) { return (int64_t)g_launches.load(); }
extern "C" int at_version(void) { return AT_VERSION; }
extern "C" const char* at_last_error_string(void) { return g_err; }
extern "C" int at_has_tcgen05(void) {
#ifdef AT_WITH_TCGEN05
  return 1;
#else
  return 0;
#endif
}
