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

#include <map>
#include <mutex>
#include <tuple>

namespace sb {
cudaError_t func_attr_cached_impl(const void* fn, cudaFuncAttribute attr, int value) {
  static std::mutex mu;
  static std::map<std::tuple<int, const void*, int>, int> seen;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(mu);
  auto key = std::make_tuple(dev, fn, (int)attr);
  auto it = seen.find(key);
  if (it != seen.end() && it->second == value) return cudaSuccess;
  e = cudaFuncSetAttribute(fn, attr, value);
  if (e == cudaSuccess) seen[key] = value;
  return e;
}
}  // namespace sb
