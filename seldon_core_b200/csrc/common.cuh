// Shared host/device helpers for libsb200.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>

#include "../../include/sb200.h"

namespace sb {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);
// cudaFuncSetAttribute, remembered per (device, kernel, attribute): the runtime call costs a few microseconds and the
// launchers used to make it in front of every launch - a visible part of the host time of a small request
cudaError_t func_attr_cached_impl(const void* fn, cudaFuncAttribute attr, int value);
template <class T>
inline cudaError_t func_attr_cached(T* fn, cudaFuncAttribute attr, int value) {
  return func_attr_cached_impl(reinterpret_cast<const void*>(fn), attr, value);
}

#define SB_CUDA(call)                                                                              \
  do {                                                                                             \
    cudaError_t e__ = (call);                                                                      \
    if (e__ != cudaSuccess) {                                                                      \
      sb::set_error("%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));           \
      return SB_ERR_CUDA;                                                                          \
    }                                                                                              \
  } while (0)

#define SB_REQUIRE(cond, code, ...)                                                                \
  do {                                                                                             \
    if (!(cond)) {                                                                                 \
      sb::set_error(__VA_ARGS__);                                                                  \
      return (code);                                                                               \
    }                                                                                              \
  } while (0)

#define SB_LAUNCH_CHECK()                                                                          \
  do {                                                                                             \
    sb::count_launch();                                                                            \
    SB_CUDA(cudaGetLastError());                                                                   \
  } while (0)

// NVTX ranges around the C-ABI entry points: compiled in with -DSB_NVTX (build.py: environment SB_NVTX=1), absent otherwise
#ifdef SB_NVTX
}  // namespace sb
#include <nvtx3/nvToolsExt.h>
namespace sb {
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};
#define SB_RANGE(name) sb::NvtxRange sb_nvtx_range__(name)
#else
#define SB_RANGE(name) do { } while (0)
#endif

constexpr int kNumSMs = 148;  // B200

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline uint64_t align_up(uint64_t a, uint64_t b) { return (a + b - 1) / b * b; }

// streaming (read-once) loads that do not allocate in L1
__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream_f4(float4* p, float4 v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ double2 ldg_stream_d2(const double2* p) {
  double2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  return r;
}

}  // namespace sb
