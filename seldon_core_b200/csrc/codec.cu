// SeldonMessage numeric payload codec on the device: Tensor (packed doubles) and ndarray (ListValue) <-> fp32.
//
// Replaces python/seldon_core/utils.py:161-176 (decode) and 235-304 (encode) of the reference.  Wire layout
// (proto3):
//   Tensor    : ... 12 <varint 8*count> <count x little-endian f64>      -> the values are one contiguous block
//   ListValue : per row  0a <varint L1> 32 <varint L2> ( 0a 09 11 <f64> ) x cols   (fixed stride for fixed cols)
// All four kernels are streaming: 12 B/value for the tensor forms, 11 + 4 (+ header) B/value for ListValue.
#include "common.cuh"

namespace {

__host__ __device__ inline int varint_len(uint64_t v) {
  int n = 1;
  while (v >= 0x80) { v >>= 7; ++n; }
  return n;
}
__host__ __device__ inline void listvalue_layout(int64_t cols, int64_t* row_bytes, int* hdr, uint64_t* l1, uint64_t* l2) {
  *l2 = 11ull * (uint64_t)cols;
  *l1 = 1 + varint_len(*l2) + *l2;
  *hdr = 1 + varint_len(*l1) + 1 + varint_len(*l2);
  *row_bytes = *hdr + (int64_t)*l2;
}

// 8 bytes at an arbitrary byte address through two aligned 8-byte loads (byte loads next to the buffer end, so
// that nothing past `end` is ever touched)
__device__ __forceinline__ uint64_t load_u64_unaligned(const uint8_t* p, const uint8_t* end) {
  const uintptr_t a = (uintptr_t)p;
  if (p + 16 > end) {
    uint64_t v = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) v |= (uint64_t)p[k] << (8 * k);
    return v;
  }
  const uint64_t* base = (const uint64_t*)(a & ~(uintptr_t)7);
  const int sh = (int)(a & 7) * 8;
  uint64_t w0 = __ldg(base);
  if (sh == 0) return w0;
  uint64_t w1 = __ldg(base + 1);
  return (w0 >> sh) | (w1 << (64 - sh));
}

template <bool STD>
__global__ void __launch_bounds__(256)
unpack_tensor_kernel(const uint8_t* __restrict__ vals, int64_t rows, int64_t cols, const float* __restrict__ mu,
                     const float* __restrict__ inv_sigma, float* __restrict__ out, int64_t ld, int aligned) {
  const int64_t total = rows * cols;
  const uint8_t* end = vals + 8 * total;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    double v;
    if (aligned) {
      v = __ldcs(reinterpret_cast<const double*>(vals) + i);
    } else {
      v = __longlong_as_double((long long)load_u64_unaligned(vals + 8 * i, end));
    }
    int64_t r = i, c = 0;
    if (ld != cols || STD) { r = i / cols; c = i - r * cols; }
    if (STD) v = (v - (double)__ldg(mu + c)) * (double)__ldg(inv_sigma + c);
    if (ld != cols) out[r * ld + c] = (float)v; else out[i] = (float)v;
  }
}

__global__ void __launch_bounds__(256)
unpack_listvalue_kernel(const uint8_t* __restrict__ wire, int64_t rows, int64_t cols, int64_t row_bytes, int hdr,
                        uint64_t l1, uint64_t l2, float* __restrict__ out, int64_t ld, int32_t* __restrict__ status) {
  const int64_t total = rows * cols;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const uint8_t* end = wire + rows * row_bytes;
  int bad = 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / cols, c = i - r * cols;
    const uint8_t* rec = wire + r * row_bytes + hdr + 11 * c;
    // the 3 tag bytes and the value: bytes rec[0..10]
    const uint64_t lo8 = load_u64_unaligned(rec, end);      // rec[0..7]
    const uint64_t hi8 = load_u64_unaligned(rec + 3, end);  // rec[3..10] = the f64
    if ((lo8 & 0xffffffull) != 0x11090aull) bad = 1;
    out[r * ld + c] = (float)__longlong_as_double((long long)hi8);
    if (c == 0) {  // row header: 0a <varint L1> 32 <varint L2>
      const uint8_t* h = wire + r * row_bytes;
      int pos = 0;
      if (h[pos++] != 0x0a) bad = 1;
      uint64_t v = 0; int shf = 0;
      for (;;) { uint8_t b = h[pos++]; v |= (uint64_t)(b & 0x7f) << shf; shf += 7; if (!(b & 0x80) || pos >= hdr) break; }
      if (v != l1) bad = 1;
      if (pos < hdr && h[pos++] != 0x32) bad = 1;
      v = 0; shf = 0;
      while (pos < hdr) { uint8_t b = h[pos++]; v |= (uint64_t)(b & 0x7f) << shf; shf += 7; if (!(b & 0x80)) break; }
      if (v != l2 || pos != hdr) bad = 1;
    }
  }
  if (bad) atomicExch(status, 1);
}

template <typename TIn>
__global__ void __launch_bounds__(256)
pack_tensor_kernel(const TIn* __restrict__ in, int64_t n, uint8_t* __restrict__ out, int aligned) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const double v = (double)in[i];
    if (aligned) {
      __stcs(reinterpret_cast<double*>(out) + i, v);
    } else {
      uint64_t b = (uint64_t)__double_as_longlong(v);
#pragma unroll
      for (int k = 0; k < 8; ++k) out[8 * i + k] = (uint8_t)(b >> (8 * k));
    }
  }
}

// one thread per value: 11 bytes `0a 09 11 <f64>`; thread c == 0 also writes the row header
template <typename TIn>
__global__ void __launch_bounds__(256)
pack_listvalue_kernel(const TIn* __restrict__ in, int64_t rows, int64_t cols, int64_t row_bytes, int hdr, uint64_t l1,
                      uint64_t l2, uint8_t* __restrict__ out) {
  const int64_t ncol = cols > 0 ? cols : 1;
  const int64_t total = rows * ncol;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int64_t r = i / ncol, c = i - r * ncol;
    uint8_t* rec = cols > 0 ? out + r * row_bytes + hdr + 11 * c : out + 11 * i;
    const uint64_t b = (uint64_t)__double_as_longlong((double)in[i]);
    rec[0] = 0x0a; rec[1] = 0x09; rec[2] = 0x11;
#pragma unroll
    for (int k = 0; k < 8; ++k) rec[3 + k] = (uint8_t)(b >> (8 * k));
    if (cols > 0 && c == 0) {
      uint8_t* h = out + r * row_bytes;
      int pos = 0;
      h[pos++] = 0x0a;
      uint64_t v = l1;
      while (v >= 0x80) { h[pos++] = (uint8_t)(v | 0x80); v >>= 7; }
      h[pos++] = (uint8_t)v;
      h[pos++] = 0x32;
      v = l2;
      while (v >= 0x80) { h[pos++] = (uint8_t)(v | 0x80); v >>= 7; }
      h[pos++] = (uint8_t)v;
    }
  }
}

int grid_for(int64_t total, int per_thread) {
  int64_t want = sb::ceil_div(total, 256 * (int64_t)per_thread);
  if (want < 1) want = 1;
  return (int)(want < (int64_t)sb::kNumSMs * 8 ? want : (int64_t)sb::kNumSMs * 8);
}

}  // namespace

extern "C" uint64_t sb_listvalue_nbytes(int64_t rows, int64_t cols) {
  if (rows <= 0) return 0;
  if (cols <= 0) return 11ull * (uint64_t)rows;
  int64_t rb; int hdr; uint64_t l1, l2;
  listvalue_layout(cols, &rb, &hdr, &l1, &l2);
  return (uint64_t)rb * (uint64_t)rows;
}

extern "C" uint64_t sb_tensor_values_offset(int64_t rows, int64_t cols) {
  const int shape_bytes = varint_len((uint64_t)rows) + varint_len((uint64_t)cols);
  return 1 + varint_len(shape_bytes) + shape_bytes + 1 + varint_len(8ull * (uint64_t)rows * (uint64_t)cols);
}

extern "C" int32_t sb_unpack_tensor_f64(const uint8_t* d_wire, uint64_t values_off, int64_t rows, int64_t cols,
                                        const float* d_mu, const float* d_inv_sigma, float* d_out, int64_t ld,
                                        uintptr_t stream) {
  SB_REQUIRE(rows >= 0 && cols >= 1 && ld >= cols, SB_ERR_INVALID, "sb_unpack_tensor_f64: bad shape");
  SB_REQUIRE((d_mu == nullptr) == (d_inv_sigma == nullptr), SB_ERR_INVALID, "sb_unpack_tensor_f64: mu and inv_sigma go together");
  if (rows == 0) return SB_OK;
  SB_REQUIRE(d_wire && d_out, SB_ERR_INVALID, "sb_unpack_tensor_f64: null buffer");
  const uint8_t* vals = d_wire + values_off;
  const int aligned = ((uintptr_t)vals & 7) == 0;
  const int grid = grid_for(rows * cols, 4);
  cudaStream_t st = (cudaStream_t)stream;
  if (d_mu)
    unpack_tensor_kernel<true><<<grid, 256, 0, st>>>(vals, rows, cols, d_mu, d_inv_sigma, d_out, ld, aligned);
  else
    unpack_tensor_kernel<false><<<grid, 256, 0, st>>>(vals, rows, cols, d_mu, d_inv_sigma, d_out, ld, aligned);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_unpack_listvalue(const uint8_t* d_wire, uint64_t nbytes, int64_t rows, int64_t cols, float* d_out,
                                       int64_t ld, int32_t* d_status, uintptr_t stream) {
  SB_REQUIRE(rows >= 0 && cols >= 1 && ld >= cols, SB_ERR_INVALID, "sb_unpack_listvalue: bad shape");
  SB_REQUIRE(d_status != nullptr, SB_ERR_INVALID, "sb_unpack_listvalue: status word required");
  SB_REQUIRE(nbytes == sb_listvalue_nbytes(rows, cols), SB_ERR_INVALID,
             "sb_unpack_listvalue: %llu bytes is not a %lld x %lld numeric ListValue (expected %llu)",
             (unsigned long long)nbytes, (long long)rows, (long long)cols,
             (unsigned long long)sb_listvalue_nbytes(rows, cols));
  cudaStream_t st = (cudaStream_t)stream;
  SB_CUDA(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
  if (rows == 0) return SB_OK;
  SB_REQUIRE(d_wire && d_out, SB_ERR_INVALID, "sb_unpack_listvalue: null buffer");
  int64_t rb; int hdr; uint64_t l1, l2;
  listvalue_layout(cols, &rb, &hdr, &l1, &l2);
  unpack_listvalue_kernel<<<grid_for(rows * cols, 4), 256, 0, st>>>(d_wire, rows, cols, rb, hdr, l1, l2, d_out, ld, d_status);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_pack_tensor_f64(const void* d_in, int32_t in_dtype, int64_t n, uint8_t* d_wire_values,
                                      uintptr_t stream) {
  SB_REQUIRE(n >= 0 && (in_dtype == SB_F32 || in_dtype == SB_I32), SB_ERR_INVALID, "sb_pack_tensor_f64: bad arguments");
  if (n == 0) return SB_OK;
  SB_REQUIRE(d_in && d_wire_values, SB_ERR_INVALID, "sb_pack_tensor_f64: null buffer");
  const int aligned = ((uintptr_t)d_wire_values & 7) == 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (in_dtype == SB_F32)
    pack_tensor_kernel<float><<<grid_for(n, 4), 256, 0, st>>>((const float*)d_in, n, d_wire_values, aligned);
  else
    pack_tensor_kernel<int32_t><<<grid_for(n, 4), 256, 0, st>>>((const int32_t*)d_in, n, d_wire_values, aligned);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_pack_listvalue(const void* d_in, int32_t in_dtype, int64_t rows, int64_t cols, uint8_t* d_wire,
                                     uintptr_t stream) {
  SB_REQUIRE(rows >= 0 && cols >= 0 && (in_dtype == SB_F32 || in_dtype == SB_I32), SB_ERR_INVALID,
             "sb_pack_listvalue: bad arguments");
  if (rows == 0) return SB_OK;
  SB_REQUIRE(d_in && d_wire, SB_ERR_INVALID, "sb_pack_listvalue: null buffer");
  int64_t rb = 11; int hdr = 0; uint64_t l1 = 0, l2 = 0;
  if (cols > 0) listvalue_layout(cols, &rb, &hdr, &l1, &l2);
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for(rows * (cols > 0 ? cols : 1), 4);
  if (in_dtype == SB_F32)
    pack_listvalue_kernel<float><<<grid, 256, 0, st>>>((const float*)d_in, rows, cols, rb, hdr, l1, l2, d_wire);
  else
    pack_listvalue_kernel<int32_t><<<grid, 256, 0, st>>>((const int32_t*)d_in, rows, cols, rb, hdr, l1, l2, d_wire);
  SB_LAUNCH_CHECK();
  return SB_OK;
}
