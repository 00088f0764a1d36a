// Error reporting, version and launch accounting for libsb200.so.
#include <atomic>
#include <stdarg.h>
#include <stdio.h>

#include "common.cuh"

namespace sb {
static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof g_err, fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
}  // namespace sb

extern "C" const char* sb_last_error(void) { return sb::g_err; }
extern "C" int32_t sb_version(void) {
  SB_RANGE("sb_version"); return 100; }
extern "C" int64_t sb_launch_count(void) { return sb::g_launches.load(std::memory_order_relaxed); }
