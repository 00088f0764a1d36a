// RANDOM_ABTEST, the engine's built-in two-way router, batched: T route() calls per launch.
//
// Replaces engine/src/main/java/io/seldon/engine/predictors/RandomABTestUnit.java:29-59 of the reference:
// `Random rand = new Random(1337)` (L31), per call `comparator = rand.nextFloat()` (L45) and branch 0 iff
// `comparator <= ratioA` (L52), else branch 1.  java.util.Random is the 48-bit linear congruential generator of its
// specification: seed0 = (seed ^ 0x5DEECE66D) mod 2^48, s <- (s * 0x5DEECE66D + 0xB) mod 2^48, next(bits) = s >> (48 -
// bits), nextFloat() = next(24) / 2^24 (exact in binary32).  The recurrence is affine, so the state before call t is
// A_t s0 + C_t with (A_t, C_t) the t-fold composition of the step: every thread jumps to the start of its own run of
// calls by square-and-multiply on affine maps and walks it from there - the stream is consumed in parallel and is
// still, call for call, the one the Java object produces.
#include "common.cuh"

namespace {

constexpr unsigned long long kMul = 0x5DEECE66DULL, kAdd = 0xBULL, kMask = (1ULL << 48) - 1;
constexpr int kCallsPerThread = 16;

// (a, c) o (b, d): apply (b, d) first, then (a, c)   -> x -> a (b x + d) + c
__device__ __forceinline__ void affine_pow(unsigned long long n, unsigned long long& A, unsigned long long& C) {
  unsigned long long a = kMul, c = kAdd;  // the map being squared
  A = 1ULL; C = 0ULL;                      // identity
  while (n) {
    if (n & 1ULL) { A = (a * A) & kMask; C = (a * C + c) & kMask; }
    c = (a * c + c) & kMask;
    a = (a * a) & kMask;
    n >>= 1;
  }
}

__global__ void __launch_bounds__(256)
abtest_route_kernel(unsigned long long* __restrict__ state, float ratio_a, long long T, int* __restrict__ decisions) {
  const unsigned long long s0 = state[0];
  const long long first = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * kCallsPerThread;
  if (first == 0) {  // the state after T calls, for the next launch (slot 1: every thread still reads slot 0)
    unsigned long long A, C;
    affine_pow((unsigned long long)T, A, C);
    state[1] = (A * s0 + C) & kMask;
  }
  if (first >= T) return;
  unsigned long long A, C;
  affine_pow((unsigned long long)first, A, C);
  unsigned long long s = (A * s0 + C) & kMask;
  const long long last = first + kCallsPerThread < T ? first + kCallsPerThread : T;
  for (long long t = first; t < last; ++t) {
    s = (s * kMul + kAdd) & kMask;
    const float comparator = (float)(unsigned)(s >> 24) * (1.0f / 16777216.0f);  // next(24) / (float)(1 << 24)
    decisions[t] = comparator <= ratio_a ? 0 : 1;
  }
}

}  // namespace

extern "C" uint64_t sb_abtest_initial_state(int64_t seed) { return ((uint64_t)seed ^ kMul) & kMask; }

extern "C" int32_t sb_abtest_route(uint64_t* d_state2, float ratio_a, int64_t T, int32_t* d_decisions, uintptr_t stream) {
  SB_RANGE("sb_abtest_route");
  SB_REQUIRE(d_state2 && T >= 0, SB_ERR_INVALID, "sb_abtest_route: bad arguments");
  SB_REQUIRE(ratio_a == ratio_a, SB_ERR_INVALID, "Parameter 'ratioA' is missing.");
  if (T == 0) return SB_OK;
  SB_REQUIRE(d_decisions, SB_ERR_INVALID, "sb_abtest_route: null decisions");
  cudaStream_t st = (cudaStream_t)stream;
  const long long threads = (T + kCallsPerThread - 1) / kCallsPerThread;
  abtest_route_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>((unsigned long long*)d_state2, ratio_a,
                                                                         (long long)T, d_decisions);
  SB_LAUNCH_CHECK();
  SB_CUDA(cudaMemcpyAsync(d_state2, d_state2 + 1, sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
  return SB_OK;
}
