// Combiner aggregate (elementwise mean over child outputs) and the standardise transformer.
//
// Replaces examples/combiners/mnist_combiner/MnistCombiner.py:8-11 generalised to N children per
// engine/.../AverageCombinerUnit.java:35-79 (sum in child order, divide by N), and CoreVAE.py:128.
// Both are pure streaming kernels: 4*(N+1) B per element for the mean, 8 B per element for standardise;
// the roofline is HBM.  Loads bypass L1 (read once), 16-byte vectors, grid = a multiple of the SM count.
#include "common.cuh"

namespace {

constexpr int kMaxChildren = 16;
struct ChildPtrs {
  const float* p[kMaxChildren];
};

template <int N, int UNROLL>
__global__ void __launch_bounds__(256) combiner_mean_vec4(ChildPtrs in, float* __restrict__ out, int64_t n_vec,
                                                          float inv_n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (UNROLL - 1) * stride < n_vec; i += UNROLL * stride) {
    float4 acc[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc[u] = sb::ldg_stream_f4(reinterpret_cast<const float4*>(in.p[0]) + i + u * stride);
#pragma unroll
    for (int j = 1; j < N; ++j) {
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        float4 v = sb::ldg_stream_f4(reinterpret_cast<const float4*>(in.p[j]) + i + u * stride);
        acc[u].x += v.x; acc[u].y += v.y; acc[u].z += v.z; acc[u].w += v.w;
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      float4 r = make_float4(acc[u].x * inv_n, acc[u].y * inv_n, acc[u].z * inv_n, acc[u].w * inv_n);
      sb::stg_stream_f4(reinterpret_cast<float4*>(out) + i + u * stride, r);
    }
  }
  for (; i < n_vec; i += stride) {
    float4 acc = sb::ldg_stream_f4(reinterpret_cast<const float4*>(in.p[0]) + i);
#pragma unroll
    for (int j = 1; j < N; ++j) {
      float4 v = sb::ldg_stream_f4(reinterpret_cast<const float4*>(in.p[j]) + i);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    sb::stg_stream_f4(reinterpret_cast<float4*>(out) + i, make_float4(acc.x * inv_n, acc.y * inv_n, acc.z * inv_n, acc.w * inv_n));
  }
}

// any N, any alignment: scalar elements [first, n)
__global__ void combiner_mean_scalar(ChildPtrs in, int n_children, float* __restrict__ out, int64_t first, int64_t n,
                                     float inv_n, int exact_div) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = first + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    float acc = in.p[0][i];
    for (int j = 1; j < n_children; ++j) acc += in.p[j][i];
    out[i] = exact_div ? acc / (float)n_children : acc * inv_n;
  }
}

template <int N>
void launch_vec(const ChildPtrs& ptrs, float* out, int64_t n_vec, float inv_n, cudaStream_t st) {
  constexpr int UNROLL = (N <= 4) ? 4 : 2;
  int64_t want = sb::ceil_div(n_vec, 256 * UNROLL);
  int grid = (int)(want < (int64_t)sb::kNumSMs * 8 ? (want > 0 ? want : 1) : (int64_t)sb::kNumSMs * 8);
  combiner_mean_vec4<N, UNROLL><<<grid, 256, 0, st>>>(ptrs, out, n_vec, inv_n);
}

__global__ void __launch_bounds__(256) standardise_kernel(const float* __restrict__ in, int64_t rows, int cols,
                                                          int64_t ld_in, const float* __restrict__ mu,
                                                          const float* __restrict__ inv_sigma, float* __restrict__ out,
                                                          int64_t ld_out) {
  const int64_t total = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    int64_t r = i / cols;
    int c = (int)(i - r * cols);
    out[r * ld_out + c] = (in[r * ld_in + c] - __ldg(mu + c)) * __ldg(inv_sigma + c);
  }
}

// contiguous rows, cols % 4 == 0: the matrix as a flat array of float4; the column of a vector advances by a fixed step
// per grid stride, so no division in the loop; four independent vectors in flight per thread
__global__ void __launch_bounds__(256) standardise_vec4(const float4* __restrict__ in, int64_t n_vec, int q,
                                                         const float4* __restrict__ mu, const float4* __restrict__ inv_sigma,
                                                         float4* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c = (int)(i % q);
  const int step = (int)(stride % q);
  constexpr int U = 4;
  for (; i < n_vec; i += U * stride) {
    float4 v[U];
    int cc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      cc[u] = c;
      c += step;
      if (c >= q) c -= q;
      const int64_t k = i + u * stride;
      if (k < n_vec) v[u] = sb::ldg_stream_f4(in + k);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t k = i + u * stride;
      if (k >= n_vec) break;
      const float4 m = __ldg(mu + cc[u]), s = __ldg(inv_sigma + cc[u]);
      sb::stg_stream_f4(out + k, make_float4((v[u].x - m.x) * s.x, (v[u].y - m.y) * s.y, (v[u].z - m.z) * s.z,
                                             (v[u].w - m.w) * s.w));
    }
  }
}

}  // namespace

extern "C" int32_t sb_combiner_mean(const float* const* h_inputs, int32_t n_children, int64_t n_elems, float* d_out,
                                    uintptr_t stream) {
  SB_RANGE("sb_combiner_mean");
  SB_REQUIRE(h_inputs != nullptr && n_children >= 1, SB_ERR_INVALID, "Combiner received no inputs");
  SB_REQUIRE(n_children <= kMaxChildren, SB_ERR_INVALID, "sb_combiner_mean: at most %d children", kMaxChildren);
  SB_REQUIRE(n_elems >= 0 && (d_out != nullptr || n_elems == 0), SB_ERR_INVALID, "sb_combiner_mean: bad output");
  if (n_elems == 0) return SB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  ChildPtrs ptrs{};
  bool aligned = ((uintptr_t)d_out & 15) == 0;
  for (int j = 0; j < n_children; ++j) {
    SB_REQUIRE(h_inputs[j] != nullptr, SB_ERR_INVALID, "sb_combiner_mean: child %d is null", j);
    ptrs.p[j] = h_inputs[j];
    aligned = aligned && (((uintptr_t)h_inputs[j] & 15) == 0);
  }
  const float inv_n = 1.0f / (float)n_children;
  const bool pow2 = (n_children & (n_children - 1)) == 0;  // multiplying by 1/N is exact only then
  int64_t n_vec = (aligned && pow2) ? n_elems / 4 : 0;
  if (n_vec > 0) {
    switch (n_children) {
      case 1: launch_vec<1>(ptrs, d_out, n_vec, inv_n, st); break;
      case 2: launch_vec<2>(ptrs, d_out, n_vec, inv_n, st); break;
      case 4: launch_vec<4>(ptrs, d_out, n_vec, inv_n, st); break;
      case 8: launch_vec<8>(ptrs, d_out, n_vec, inv_n, st); break;
      case 16: launch_vec<16>(ptrs, d_out, n_vec, inv_n, st); break;
      default: n_vec = 0; break;
    }
    if (n_vec > 0) SB_LAUNCH_CHECK();
  }
  if (n_vec * 4 < n_elems) {
    int64_t rest = n_elems - n_vec * 4;
    int64_t want = sb::ceil_div(rest, 256);
    int grid = (int)(want < (int64_t)sb::kNumSMs * 16 ? want : (int64_t)sb::kNumSMs * 16);
    combiner_mean_scalar<<<grid, 256, 0, st>>>(ptrs, n_children, d_out, n_vec * 4, n_elems, inv_n, pow2 ? 0 : 1);
    SB_LAUNCH_CHECK();
  }
  return SB_OK;
}

extern "C" int32_t sb_standardise(const float* d_in, int64_t rows, int64_t cols, int64_t ld_in, const float* d_mu,
                                  const float* d_inv_sigma, float* d_out, int64_t ld_out, uintptr_t stream) {
  SB_RANGE("sb_standardise");
  SB_REQUIRE(rows >= 0 && cols >= 1 && ld_in >= cols && ld_out >= cols, SB_ERR_INVALID, "sb_standardise: bad shape");
  SB_REQUIRE(d_mu && d_inv_sigma, SB_ERR_INVALID, "sb_standardise: mu/inv_sigma required");
  if (rows == 0) return SB_OK;
  SB_REQUIRE(d_in && d_out, SB_ERR_INVALID, "sb_standardise: null matrix");
  const bool vec = cols % 4 == 0 && ld_in == cols && ld_out == cols &&
                   (((uintptr_t)d_in | (uintptr_t)d_out | (uintptr_t)d_mu | (uintptr_t)d_inv_sigma) & 15) == 0;
  if (vec) {
    const int64_t n_vec = rows * cols / 4;
    int64_t w4 = sb::ceil_div(n_vec, 256 * 4);
    int g4 = (int)(w4 < (int64_t)sb::kNumSMs * 8 ? (w4 > 0 ? w4 : 1) : (int64_t)sb::kNumSMs * 8);
    standardise_vec4<<<g4, 256, 0, (cudaStream_t)stream>>>((const float4*)d_in, n_vec, (int)(cols / 4), (const float4*)d_mu,
                                                          (const float4*)d_inv_sigma, (float4*)d_out);
    SB_LAUNCH_CHECK();
    return SB_OK;
  }
  int64_t want = sb::ceil_div(rows * cols, 256 * 4);
  int grid = (int)(want < (int64_t)sb::kNumSMs * 8 ? (want > 0 ? want : 1) : (int64_t)sb::kNumSMs * 8);
  standardise_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_in, rows, (int)cols, ld_in, d_mu, d_inv_sigma, d_out,
                                                            ld_out);
  SB_LAUNCH_CHECK();
  return SB_OK;
}
