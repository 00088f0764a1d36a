// SeldonMessage numeric payload codec on the device: Tensor (packed doubles) and ndarray (ListValue) <-> fp32.
//
// Replaces python/seldon_core/utils.py:161-176 (decode) and 235-304 (encode) of the reference.  Wire layout
// (proto3):
//   Tensor    : ... 12 <varint 8*count> <count x little-endian f64>      -> the values are one contiguous block
//   ListValue : per row  0a <varint L1> 32 <varint L2> ( 0a 09 11 <f64> ) x cols   (fixed stride for fixed cols)
// All four kernels are streaming: 12 B/value for the tensor forms, 11 + 4 (+ header) B/value for ListValue.
#include <stdlib.h>

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
  constexpr int U = 4;  // independent loads in flight per thread, each warp access stays contiguous
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += U * stride) {
    double v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int64_t i = i0 + j * stride;
      v[j] = 0.0;
      if (i < total)
        v[j] = aligned ? __ldcs(reinterpret_cast<const double*>(vals) + i)
                       : __longlong_as_double((long long)load_u64_unaligned(vals + 8 * i, end));
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int64_t i = i0 + j * stride;
      if (i >= total) break;
      int64_t r = i, c = 0;
      if (ld != cols || STD) { r = i / cols; c = i - r * cols; }
      double w = v[j];
      if (STD) w = (w - (double)__ldg(mu + c)) * (double)__ldg(inv_sigma + c);
      if (ld != cols) out[r * ld + c] = (float)w; else __stcs(out + i, (float)w);
    }
  }
}

// 16-byte aligned, contiguous case: four values per thread (32 wire bytes <-> one float4), two groups in flight;
// the column of a group advances by a fixed step per grid stride (no division in the loop)
template <bool STD>
__global__ void __launch_bounds__(256)
unpack_tensor_vec4(const uint4* __restrict__ vals, int64_t n_grp, int q, const float4* __restrict__ mu,
                   const float4* __restrict__ inv_sigma, float4* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int c = STD ? (int)(i % q) : 0;
  const int step = STD ? (int)(stride % q) : 0;
  constexpr int U = 2;
  for (; i < n_grp; i += U * stride) {
    uint4 a[U], b[U];
    int cc[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      cc[u] = c;
      c += step;
      if (c >= q) c -= q;
      const int64_t k = i + u * stride;
      if (k < n_grp) { a[u] = __ldcs(vals + 2 * k); b[u] = __ldcs(vals + 2 * k + 1); }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t k = i + u * stride;
      if (k >= n_grp) break;
      double v0 = __hiloint2double((int)a[u].y, (int)a[u].x), v1 = __hiloint2double((int)a[u].w, (int)a[u].z);
      double v2 = __hiloint2double((int)b[u].y, (int)b[u].x), v3 = __hiloint2double((int)b[u].w, (int)b[u].z);
      if (STD) {
        const float4 m = __ldg(mu + cc[u]), s = __ldg(inv_sigma + cc[u]);
        v0 = (v0 - (double)m.x) * (double)s.x; v1 = (v1 - (double)m.y) * (double)s.y;
        v2 = (v2 - (double)m.z) * (double)s.z; v3 = (v3 - (double)m.w) * (double)s.w;
      }
      sb::stg_stream_f4(out + k, make_float4((float)v0, (float)v1, (float)v2, (float)v3));
    }
  }
}

// two values per thread: an 8-byte load and one 16-byte store, both contiguous across the warp; four pairs in flight
template <typename TIn>
__global__ void __launch_bounds__(256)
pack_tensor_vec2(const uint2* __restrict__ in, int64_t n_pair, uint4* __restrict__ out) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  constexpr int U = 4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pair; i += U * stride) {
    uint2 x[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i + u * stride < n_pair) x[u] = __ldcs(in + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int64_t k = i + u * stride;
      if (k >= n_pair) break;
      TIn t0, t1;
      memcpy(&t0, &x[u].x, 4);
      memcpy(&t1, &x[u].y, 4);
      const double d0 = (double)t0, d1 = (double)t1;
      __stcs(out + k, make_uint4((uint32_t)__double2loint(d0), (uint32_t)__double2hiint(d0), (uint32_t)__double2loint(d1),
                                 (uint32_t)__double2hiint(d1)));
    }
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
  constexpr int U = 4;
  for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += U * stride) {
    TIn x[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int64_t i = i0 + j * stride;
      x[j] = i < n ? __ldcs(in + i) : TIn(0);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int64_t i = i0 + j * stride;
      if (i >= n) break;
      const double v = (double)x[j];
      if (aligned) {
        __stcs(reinterpret_cast<double*>(out) + i, v);
      } else {
        uint64_t b = (uint64_t)__double_as_longlong(v);
#pragma unroll
        for (int k = 0; k < 8; ++k) out[8 * i + k] = (uint8_t)(b >> (8 * k));
      }
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

// ---- shared-memory staged ListValue kernels -------------------------------------------------------------------
// A CTA moves a chunk of R rows (R a multiple of 16, so the chunk is a whole number of 16-byte words when the payload
// base is 16-byte aligned) between global and shared memory with coalesced 16-byte accesses; the 11-byte records are
// then picked out of / assembled in shared memory with aligned 32-bit words and funnel shifts.
__device__ __forceinline__ uint32_t smem_u32_at(const uint32_t* w, int byte_off) {  // 4 bytes at any byte offset
  const int k = byte_off >> 2, sh = (byte_off & 3) * 8;
  return __funnelshift_r(w[k], w[k + 1], sh);
}

__global__ void __launch_bounds__(256)
unpack_listvalue_staged(const uint8_t* __restrict__ wire, int64_t rows, int cols, int row_bytes, int hdr, uint32_t l1,
                        uint32_t l2, int R, float* __restrict__ out, int64_t ld, int32_t* __restrict__ status) {
  extern __shared__ __align__(16) uint32_t sm_words[];
  const int64_t nchunks = (rows + R - 1) / R;
  int bad = 0;
  for (int64_t ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
    const int64_t r0 = ch * R;
    const int nr = (int)min((int64_t)R, rows - r0);
    const int nbytes = nr * row_bytes;
    const uint8_t* src = wire + r0 * (int64_t)row_bytes;  // 16-byte aligned (R % 16 == 0, aligned base)
    __syncthreads();
    const int nvec = nbytes >> 4;
    for (int v = threadIdx.x; v < nvec; v += blockDim.x)
      reinterpret_cast<uint4*>(sm_words)[v] = __ldcs(reinterpret_cast<const uint4*>(src) + v);
    for (int b = (nvec << 4) + threadIdx.x; b < nbytes; b += blockDim.x)
      reinterpret_cast<uint8_t*>(sm_words)[b] = src[b];
    if (threadIdx.x < 2) sm_words[(nbytes >> 2) + 1 + threadIdx.x] = 0;  // words touched by the last funnel shifts
    __syncthreads();
    const int total = nr * cols;
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
      const int r = idx / cols, c = idx - r * cols;
      const int o = r * row_bytes + hdr + 11 * c;
      const uint32_t tag = smem_u32_at(sm_words, o) & 0xffffffu;
      const uint32_t lo = smem_u32_at(sm_words, o + 3), hi = smem_u32_at(sm_words, o + 7);
      if (tag != 0x11090au) bad = 1;
      out[(r0 + r) * ld + c] = (float)__hiloint2double((int)hi, (int)lo);
      if (c == 0) {  // row header: 0a <varint L1> 32 <varint L2>
        const uint8_t* h = reinterpret_cast<const uint8_t*>(sm_words) + r * row_bytes;
        int pos = 0;
        if (h[pos++] != 0x0a) bad = 1;
        uint32_t v = 0; int shf = 0;
        while (pos < hdr) { const uint8_t b = h[pos++]; v |= (uint32_t)(b & 0x7f) << shf; shf += 7; if (!(b & 0x80)) break; }
        if (v != l1) bad = 1;
        if (pos < hdr && h[pos++] != 0x32) bad = 1;
        v = 0; shf = 0;
        while (pos < hdr) { const uint8_t b = h[pos++]; v |= (uint32_t)(b & 0x7f) << shf; shf += 7; if (!(b & 0x80)) break; }
        if (v != l2 || pos != hdr) bad = 1;
      }
    }
  }
  if (bad) atomicExch(status, 1);
}

template <typename TIn>
__global__ void __launch_bounds__(256)
pack_listvalue_staged(const TIn* __restrict__ in, int64_t rows, int cols, int row_bytes, int hdr, uint32_t l1,
                      uint32_t l2, int R, uint8_t* __restrict__ out) {
  extern __shared__ __align__(16) uint32_t sm_words[];
  const int ncol = cols > 0 ? cols : 1;
  const int64_t nchunks = (rows + R - 1) / R;
  for (int64_t ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
    const int64_t r0 = ch * R;
    const int nr = (int)min((int64_t)R, rows - r0);
    const int nbytes = nr * row_bytes;
    __syncthreads();
    for (int v = threadIdx.x; v < ((nbytes + 15) >> 4) + 1; v += blockDim.x)
      reinterpret_cast<uint4*>(sm_words)[v] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    const int total = nr * ncol;
    for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
      const int r = idx / ncol, c = idx - r * ncol;
      const int o = r * row_bytes + hdr + 11 * c;
      const double dv = (double)in[(r0 + r) * (int64_t)ncol + c];
      const uint32_t lo = (uint32_t)__double2loint(dv), hi = (uint32_t)__double2hiint(dv);
      // record bytes 0a 09 11 v0..v7 as three little-endian words, shifted to the byte offset inside the word
      const uint32_t w0 = 0x0011090au | (lo << 24), w1 = (lo >> 8) | (hi << 24), w2 = hi >> 8;
      const int k = o >> 2, sh = (o & 3) * 8;
      atomicOr(sm_words + k, w0 << sh);
      atomicOr(sm_words + k + 1, __funnelshift_l(w0, w1, sh));
      atomicOr(sm_words + k + 2, __funnelshift_l(w1, w2, sh));
      if (sh >= 16) atomicOr(sm_words + k + 3, __funnelshift_l(w2, 0u, sh));
      if (cols > 0 && c == 0) {
        uint8_t hb[12];
        int pos = 0;
        hb[pos++] = 0x0a;
        uint32_t v = l1;
        while (v >= 0x80) { hb[pos++] = (uint8_t)(v | 0x80); v >>= 7; }
        hb[pos++] = (uint8_t)v;
        hb[pos++] = 0x32;
        v = l2;
        while (v >= 0x80) { hb[pos++] = (uint8_t)(v | 0x80); v >>= 7; }
        hb[pos++] = (uint8_t)v;
        const int ob = r * row_bytes;
        for (int b = 0; b < pos; ++b) atomicOr(sm_words + ((ob + b) >> 2), (uint32_t)hb[b] << (((ob + b) & 3) * 8));
      }
    }
    __syncthreads();
    uint8_t* dst = out + r0 * (int64_t)row_bytes;
    const int nvec = nbytes >> 4;
    for (int v = threadIdx.x; v < nvec; v += blockDim.x)
      __stcs(reinterpret_cast<uint4*>(dst) + v, reinterpret_cast<const uint4*>(sm_words)[v]);
    for (int b = (nvec << 4) + threadIdx.x; b < nbytes; b += blockDim.x) dst[b] = reinterpret_cast<const uint8_t*>(sm_words)[b];
  }
}

// ---- cols % 4 == 0: four consecutive records (44 bytes = 11 words) per thread, no division, no per-byte work ----
// the parse of one staged chunk of nr rows (cols % 4 == 0): returns nonzero if a tag or row header is wrong
__device__ __forceinline__ int unpack4_chunk(const uint32_t* sm_words, int nr, int q, int row_bytes, int hdr, uint32_t l1,
                                             uint32_t l2, float* __restrict__ out, int64_t r0, int64_t ld, bool vec_out) {
  int bad = 0;
  // (row, run) pairs: consecutive threads take consecutive runs of a row
  const int total = nr * q;
  int r = threadIdx.x / q, u = threadIdx.x - r * q;
  const int dr = blockDim.x / q, du = blockDim.x - dr * q;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int o = r * row_bytes + hdr + 44 * u;  // byte offset of the first of 4 records
    const int k = o >> 2, sh = (o & 3) * 8;
    uint32_t w[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) w[i] = sm_words[k + i];
    uint32_t a[11];  // the 44 bytes, word aligned
#pragma unroll
    for (int i = 0; i < 11; ++i) a[i] = __funnelshift_r(w[i], w[i + 1], sh);
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      // record e occupies bytes 11e .. 11e+10 of a[]: tag at 11e, value at 11e+3
      const int tb = 11 * e, vb = tb + 3;
      const uint32_t tag = __funnelshift_r(a[tb >> 2], a[(tb >> 2) + 1 < 11 ? (tb >> 2) + 1 : 10], (tb & 3) * 8) & 0xffffffu;
      const uint32_t lo = __funnelshift_r(a[vb >> 2], a[(vb >> 2) + 1], (vb & 3) * 8);
      const uint32_t hi = (vb >> 2) + 2 < 11 ? __funnelshift_r(a[(vb >> 2) + 1], a[(vb >> 2) + 2], (vb & 3) * 8)
                                             : (a[(vb >> 2) + 1] >> ((vb & 3) * 8));
      if (tag != 0x11090au) bad = 1;
      v[e] = (float)__hiloint2double((int)hi, (int)lo);
    }
    float* dst = out + (r0 + r) * ld + 4 * u;
    if (vec_out) *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    else { dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2]; dst[3] = v[3]; }
    if (u == 0) {
      const uint8_t* h = reinterpret_cast<const uint8_t*>(sm_words) + r * row_bytes;
      int pos = 0;
      if (h[pos++] != 0x0a) bad = 1;
      uint32_t vv = 0; int shf = 0;
      while (pos < hdr) { const uint8_t b = h[pos++]; vv |= (uint32_t)(b & 0x7f) << shf; shf += 7; if (!(b & 0x80)) break; }
      if (vv != l1) bad = 1;
      if (pos < hdr && h[pos++] != 0x32) bad = 1;
      vv = 0; shf = 0;
      while (pos < hdr) { const uint8_t b = h[pos++]; vv |= (uint32_t)(b & 0x7f) << shf; shf += 7; if (!(b & 0x80)) break; }
      if (vv != l2 || pos != hdr) bad = 1;
    }
    r += dr; u += du;
    if (u >= q) { u -= q; ++r; }
  }
  return bad;
}


__global__ void __launch_bounds__(256)
unpack_listvalue_staged4(const uint8_t* __restrict__ wire, int64_t rows, int cols, int row_bytes, int hdr, uint32_t l1,
                         uint32_t l2, int R, float* __restrict__ out, int64_t ld, int32_t* __restrict__ status) {
  extern __shared__ __align__(16) uint32_t sm_words[];
  const int64_t nchunks = (rows + R - 1) / R;
  const int q = cols >> 2;  // 4-record runs per row
  const bool vec_out = (ld & 3) == 0 && (((uintptr_t)out) & 15) == 0;
  int bad = 0;
  for (int64_t ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
    const int64_t r0 = ch * R;
    const int nr = (int)min((int64_t)R, rows - r0);
    const int nbytes = nr * row_bytes;
    const uint8_t* src = wire + r0 * (int64_t)row_bytes;
    __syncthreads();
    const int nvec = nbytes >> 4;
    for (int v = threadIdx.x; v < nvec; v += blockDim.x)
      reinterpret_cast<uint4*>(sm_words)[v] = __ldcs(reinterpret_cast<const uint4*>(src) + v);
    for (int b = (nvec << 4) + threadIdx.x; b < nbytes; b += blockDim.x)
      reinterpret_cast<uint8_t*>(sm_words)[b] = src[b];
    __syncthreads();
    bad |= unpack4_chunk(sm_words, nr, q, row_bytes, hdr, l1, l2, out, r0, ld, vec_out);
  }
  if (bad) atomicExch(status, 1);
}

template <typename TIn>
struct Vec4 { TIn v[4]; };
// four consecutive inputs (16-byte aligned: cols % 4 == 0 and an aligned base, else element loads)
template <typename TIn>
__device__ __forceinline__ Vec4<TIn> load4(const TIn* src, bool aligned) {
  Vec4<TIn> o;
  if (aligned) {
    const uint4 w = __ldcs(reinterpret_cast<const uint4*>(src));
    const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) memcpy(&o.v[e], &ww[e], 4);
  } else {
#pragma unroll
    for (int e = 0; e < 4; ++e) o.v[e] = src[e];
  }
  return o;
}
template <typename TIn>
__device__ __forceinline__ void pack4_chunk(uint32_t* sm_words, const TIn* __restrict__ in, int64_t r0, int nr, int cols,
                                            int q, int row_bytes, int hdr, uint32_t l1, uint32_t l2);

template <typename TIn>
__global__ void __launch_bounds__(256)
pack_listvalue_staged4(const TIn* __restrict__ in, int64_t rows, int cols, int row_bytes, int hdr, uint32_t l1,
                       uint32_t l2, int R, uint8_t* __restrict__ out) {
  extern __shared__ __align__(16) uint32_t sm_words[];
  const int64_t nchunks = (rows + R - 1) / R;
  const int q = cols >> 2;
  for (int64_t ch = blockIdx.x; ch < nchunks; ch += gridDim.x) {
    const int64_t r0 = ch * R;
    const int nr = (int)min((int64_t)R, rows - r0);
    const int nbytes = nr * row_bytes;
    __syncthreads();
    // only the words shared between neighbours are OR-ed into, so only the header area / row edges need zeroing;
    // zero everything (cheap) to keep it simple
    for (int v = threadIdx.x; v < ((nbytes + 15) >> 4) + 1; v += blockDim.x)
      reinterpret_cast<uint4*>(sm_words)[v] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    pack4_chunk(sm_words, in, r0, nr, cols, q, row_bytes, hdr, l1, l2);
    __syncthreads();
    uint8_t* dst = out + r0 * (int64_t)row_bytes;
    const int nvec = nbytes >> 4;
    for (int v = threadIdx.x; v < nvec; v += blockDim.x)
      __stcs(reinterpret_cast<uint4*>(dst) + v, reinterpret_cast<const uint4*>(sm_words)[v]);
    for (int b = (nvec << 4) + threadIdx.x; b < nbytes; b += blockDim.x) dst[b] = reinterpret_cast<const uint8_t*>(sm_words)[b];
  }
}

// ---- ring-buffered variants: the chunk moves between global and shared memory by bulk copies (cp.async.bulk), so a
// CTA has up to kLvStages - 1 chunks in flight while it parses / assembles another.
constexpr int kLvStages = 4;
// two slots: a third one (or fetching the inputs a chunk ahead) costs resident CTAs and was measured slower
constexpr int kLvPackSlots = 2;

__device__ __forceinline__ void lv_mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  while (!done)
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done) : "r"(bar), "r"(parity) : "memory");
}

__global__ void __launch_bounds__(256)
unpack_listvalue_ring4(const uint8_t* __restrict__ wire, int64_t rows, int cols, int row_bytes, int hdr, uint32_t l1,
                       uint32_t l2, int R, int slot_bytes, float* __restrict__ out, int64_t ld,
                       int32_t* __restrict__ status) {
  extern __shared__ __align__(128) uint8_t lv_ring[];  // [kLvStages][slot_bytes]
  __shared__ uint64_t full[kLvStages];
  const int64_t nchunks = (rows + R - 1) / R;
  const int q = cols >> 2;
  const bool vec_out = (ld & 3) == 0 && (((uintptr_t)out) & 15) == 0;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kLvStages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&full[s])));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  // this CTA's chunks: blockIdx.x + j * gridDim.x
  const int64_t mine = blockIdx.x < nchunks ? (nchunks - 1 - blockIdx.x) / gridDim.x + 1 : 0;
  auto issue = [&](int64_t j) {  // thread 0 only
    const int64_t ch = blockIdx.x + j * (int64_t)gridDim.x;
    const int64_t r0 = ch * R;
    const int nr = (int)min((int64_t)R, rows - r0);
    const uint32_t bytes = (uint32_t)(nr * row_bytes) & ~15u;  // the last chunk's tail (< 16 B) is read directly
    const uint32_t slot = (uint32_t)(j % kLvStages);
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&full[slot]);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(lv_ring + (size_t)slot * slot_bytes);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic reads of the slot are done
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    if (bytes)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst), "l"(wire + r0 * (int64_t)row_bytes), "r"(bytes), "r"(bar) : "memory");
  };
  if (threadIdx.x == 0)
    for (int64_t j = 0; j < kLvStages - 1 && j < mine; ++j) issue(j);
  int bad = 0;
  for (int64_t i = 0; i < mine; ++i) {
    // every thread has passed the barrier that ended iteration i - 1, so the slot of chunk i - 1 is free again
    if (threadIdx.x == 0 && i + kLvStages - 1 < mine) issue(i + kLvStages - 1);
    const int64_t ch = blockIdx.x + i * (int64_t)gridDim.x;
    const int64_t r0 = ch * R;
    const int nr = (int)min((int64_t)R, rows - r0);
    const int nbytes = nr * row_bytes;
    uint8_t* slot = lv_ring + (size_t)(i % kLvStages) * slot_bytes;
    lv_mbar_wait((uint32_t)__cvta_generic_to_shared(&full[i % kLvStages]), (uint32_t)((i / kLvStages) & 1));
    if ((nbytes & 15) != 0) {  // last chunk only
      const uint8_t* src = wire + r0 * (int64_t)row_bytes;
      for (int b = (nbytes & ~15) + threadIdx.x; b < nbytes; b += blockDim.x) slot[b] = src[b];
      __syncthreads();
    }
    bad |= unpack4_chunk(reinterpret_cast<const uint32_t*>(slot), nr, q, row_bytes, hdr, l1, l2, out, r0, ld, vec_out);
    __syncthreads();
  }
  if (bad) atomicExch(status, 1);
}

// the assembly of one chunk of nr rows in shared memory (cols % 4 == 0); the slot must have been zeroed
template <typename TIn>
__device__ __forceinline__ void pack4_chunk(uint32_t* sm_words, const TIn* __restrict__ in, int64_t r0, int nr, int cols,
                                            int q, int row_bytes, int hdr, uint32_t l1, uint32_t l2) {
  const bool aligned = (((uintptr_t)in) & 15) == 0;
  const int total = nr * q;
  int r = threadIdx.x / q, u = threadIdx.x - r * q;
  const int dr = blockDim.x / q, du = blockDim.x - dr * q;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const TIn* src = in + (r0 + r) * (int64_t)cols + 4 * u;
    const Vec4<TIn> x = load4(src, aligned);
    uint32_t a[11];
    uint32_t lo[4], hi[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const double dv = (double)x.v[e];
      lo[e] = (uint32_t)__double2loint(dv);
      hi[e] = (uint32_t)__double2hiint(dv);
    }
    // bytes: [0a 09 11 v0..v7] x 4, little endian words
    a[0] = 0x0011090au | (lo[0] << 24);
    a[1] = (lo[0] >> 8) | (hi[0] << 24);
    a[2] = (hi[0] >> 8) | 0x0a000000u;
    a[3] = 0x00001109u | (lo[1] << 16);
    a[4] = (lo[1] >> 16) | (hi[1] << 16);
    a[5] = (hi[1] >> 16) | 0x090a0000u;
    a[6] = 0x00000011u | (lo[2] << 8);
    a[7] = (lo[2] >> 24) | (hi[2] << 8);
    a[8] = (hi[2] >> 24) | 0x11090a00u;
    a[9] = lo[3];
    a[10] = hi[3];
    const int o = r * row_bytes + hdr + 44 * u;
    const int k = o >> 2, sh = (o & 3) * 8;
    if (sh == 0) {
#pragma unroll
      for (int i = 0; i < 11; ++i) sm_words[k + i] = a[i];
    } else {
      atomicOr(sm_words + k, a[0] << sh);  // first and last words are shared with the neighbouring runs / header
#pragma unroll
      for (int i = 1; i < 11; ++i) sm_words[k + i] = __funnelshift_l(a[i - 1], a[i], sh);
      atomicOr(sm_words + k + 11, a[10] >> (32 - sh));
    }
    if (u == 0) {
      uint8_t hb[12];
      int pos = 0;
      hb[pos++] = 0x0a;
      uint32_t v = l1;
      while (v >= 0x80) { hb[pos++] = (uint8_t)(v | 0x80); v >>= 7; }
      hb[pos++] = (uint8_t)v;
      hb[pos++] = 0x32;
      v = l2;
      while (v >= 0x80) { hb[pos++] = (uint8_t)(v | 0x80); v >>= 7; }
      hb[pos++] = (uint8_t)v;
      const int ob = r * row_bytes;
      for (int b = 0; b < pos; ++b) atomicOr(sm_words + ((ob + b) >> 2), (uint32_t)hb[b] << (((ob + b) & 3) * 8));
    }
    r += dr; u += du;
    if (u >= q) { u -= q; ++r; }
  }
}

template <typename TIn>
__global__ void __launch_bounds__(256)
pack_listvalue_ring4(const TIn* __restrict__ in, int64_t rows, int cols, int row_bytes, int hdr, uint32_t l1, uint32_t l2,
                     int R, int slot_bytes, uint8_t* __restrict__ out) {
  extern __shared__ __align__(128) uint8_t lv_ring[];  // [kLvPackSlots][slot_bytes]
  const int64_t nchunks = (rows + R - 1) / R;
  const int q = cols >> 2;
  int64_t i = 0;
  for (int64_t ch = blockIdx.x; ch < nchunks; ch += gridDim.x, ++i) {
    const int64_t r0 = ch * R;
    const int nr = (int)min((int64_t)R, rows - r0);
    const int nbytes = nr * row_bytes;
    uint8_t* slot = lv_ring + (size_t)(i % kLvPackSlots) * slot_bytes;
    // the bulk store that last read this slot (kLvPackSlots chunks ago) must have finished reading it
    if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kLvPackSlots - 1) : "memory");
    __syncthreads();
    for (int v = threadIdx.x; v < ((nbytes + 15) >> 4) + 1; v += blockDim.x)
      reinterpret_cast<uint4*>(slot)[v] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    pack4_chunk(reinterpret_cast<uint32_t*>(slot), in, r0, nr, cols, q, row_bytes, hdr, l1, l2);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic stores -> visible to the bulk copy engine
    __syncthreads();
    uint8_t* dst = out + r0 * (int64_t)row_bytes;
    const uint32_t bytes = (uint32_t)nbytes & ~15u;
    if (threadIdx.x == 0) {
      if (bytes)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                     ::"l"(dst), "r"((uint32_t)__cvta_generic_to_shared(slot)), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    for (int b = (int)bytes + threadIdx.x; b < nbytes; b += blockDim.x) dst[b] = slot[b];  // last chunk's tail
  }
  if (threadIdx.x == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// rows per staged chunk: a multiple of 16 with the chunk near 32 KB; 0 = row too long for shared memory
int staged_rows(int64_t row_bytes) {
  if (row_bytes * 16 + 64 > 200 * 1024) return 0;
  static const int64_t target = getenv("SB200_LV_CHUNK") ? atoll(getenv("SB200_LV_CHUNK")) : 32 * 1024;
  int64_t m = target / (row_bytes * 16);
  if (m < 1) m = 1;
  return (int)(16 * m);
}

// ring kernels: rows per chunk (a multiple of 16) and bytes per ring slot; 0 rows = do not use them
int ring_rows(int64_t row_bytes, int stages) {
  static const int enabled = getenv("SB200_LV_RING") ? atoi(getenv("SB200_LV_RING")) : 1;
  static const int64_t target = getenv("SB200_LV_RING_CHUNK") ? atoll(getenv("SB200_LV_RING_CHUNK")) : 12 * 1024;
  if (!enabled) return 0;
  int64_t m = target / (row_bytes * 16);
  if (m < 1) m = 1;
  if ((m * 16 * row_bytes + 192) * stages > 200 * 1024) return 0;
  return (int)(16 * m);
}
int ring_slot_bytes(int64_t row_bytes, int R) { return (int)(((int64_t)R * row_bytes + 64 + 127) & ~(int64_t)127); }

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

extern "C" int32_t sb_unpack_tensor_f64(const uint8_t* d_wire, uint64_t wire_nbytes, uint64_t values_off, int64_t rows,
                                        int64_t cols,
                                        const float* d_mu, const float* d_inv_sigma, float* d_out, int64_t ld,
                                        uintptr_t stream) {
  SB_RANGE("sb_unpack_tensor_f64");
  SB_REQUIRE(rows >= 0 && cols >= 1 && ld >= cols, SB_ERR_INVALID, "sb_unpack_tensor_f64: bad shape");
  SB_REQUIRE((d_mu == nullptr) == (d_inv_sigma == nullptr), SB_ERR_INVALID, "sb_unpack_tensor_f64: mu and inv_sigma go together");
  SB_REQUIRE(values_off <= wire_nbytes && (uint64_t)rows <= (wire_nbytes - values_off) / 8 / (uint64_t)cols, SB_ERR_INVALID,
             "sb_unpack_tensor_f64: shape %lld x %lld needs %llu bytes of values at offset %llu, the payload has %llu",
             (long long)rows, (long long)cols, (unsigned long long)(8ull * rows * cols), (unsigned long long)values_off,
             (unsigned long long)wire_nbytes);
  if (rows == 0) return SB_OK;
  SB_REQUIRE(d_wire && d_out, SB_ERR_INVALID, "sb_unpack_tensor_f64: null buffer");
  const uint8_t* vals = d_wire + values_off;
  const int aligned = ((uintptr_t)vals & 7) == 0;
  const int grid = grid_for(rows * cols, 4);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t total = rows * cols;
  const bool vec = ld == cols && total % 4 == 0 && (((uintptr_t)vals | (uintptr_t)d_out) & 15) == 0 &&
                   (!d_mu || (cols % 4 == 0 && (((uintptr_t)d_mu | (uintptr_t)d_inv_sigma) & 15) == 0));
  if (vec) {
    const int64_t n_grp = total / 4;
    const int g4 = grid_for(n_grp, 2);
    if (d_mu)
      unpack_tensor_vec4<true><<<g4, 256, 0, st>>>((const uint4*)vals, n_grp, (int)(cols / 4), (const float4*)d_mu,
                                                  (const float4*)d_inv_sigma, (float4*)d_out);
    else
      unpack_tensor_vec4<false><<<g4, 256, 0, st>>>((const uint4*)vals, n_grp, 1, nullptr, nullptr, (float4*)d_out);
    SB_LAUNCH_CHECK();
    return SB_OK;
  }
  if (d_mu)
    unpack_tensor_kernel<true><<<grid, 256, 0, st>>>(vals, rows, cols, d_mu, d_inv_sigma, d_out, ld, aligned);
  else
    unpack_tensor_kernel<false><<<grid, 256, 0, st>>>(vals, rows, cols, d_mu, d_inv_sigma, d_out, ld, aligned);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_unpack_listvalue(const uint8_t* d_wire, uint64_t nbytes, int64_t rows, int64_t cols, float* d_out,
                                       int64_t ld, int32_t* d_status, uintptr_t stream) {
  SB_RANGE("sb_unpack_listvalue");
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
  const int R = staged_rows(rb);
  if (R > 0 && ((uintptr_t)d_wire & 15) == 0 && cols < (1 << 20)) {
    const size_t smem = (size_t)R * rb + 64;
    SB_CUDA(sb::func_attr_cached(unpack_listvalue_staged, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    int64_t nch = sb::ceil_div(rows, R);
    int grid = (int)(nch < (int64_t)sb::kNumSMs * 12 ? nch : (int64_t)sb::kNumSMs * 12);
    const int RR = cols % 4 == 0 ? ring_rows(rb, kLvStages) : 0;
    if (RR > 0) {
      const int sb = ring_slot_bytes(rb, RR);
      const size_t rsm = (size_t)kLvStages * sb;
      SB_CUDA(sb::func_attr_cached(unpack_listvalue_ring4, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      const int64_t nch2 = sb::ceil_div(rows, RR);
      const int per_sm = (int)(200 * 1024 / (rsm + 1024)) < 8 ? (int)(200 * 1024 / (rsm + 1024)) : 8;
      const int64_t gmax = (int64_t)sb::kNumSMs * (per_sm > 0 ? per_sm : 1);
      unpack_listvalue_ring4<<<(int)(nch2 < gmax ? nch2 : gmax), 256, rsm, st>>>(d_wire, rows, (int)cols, (int)rb, hdr,
                                                                              (uint32_t)l1, (uint32_t)l2, RR, sb, d_out,
                                                                              ld, d_status);
    } else if (cols % 4 == 0) {
      SB_CUDA(sb::func_attr_cached(unpack_listvalue_staged4, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      unpack_listvalue_staged4<<<grid, 256, smem, st>>>(d_wire, rows, (int)cols, (int)rb, hdr, (uint32_t)l1, (uint32_t)l2,
                                                       R, d_out, ld, d_status);
    } else {
      unpack_listvalue_staged<<<grid, 256, smem, st>>>(d_wire, rows, (int)cols, (int)rb, hdr, (uint32_t)l1, (uint32_t)l2,
                                                      R, d_out, ld, d_status);
    }
  } else {
    unpack_listvalue_kernel<<<grid_for(rows * cols, 4), 256, 0, st>>>(d_wire, rows, cols, rb, hdr, l1, l2, d_out, ld, d_status);
  }
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_pack_tensor_f64(const void* d_in, int32_t in_dtype, int64_t n, uint8_t* d_wire_values,
                                      uintptr_t stream) {
  SB_RANGE("sb_pack_tensor_f64");
  SB_REQUIRE(n >= 0 && (in_dtype == SB_F32 || in_dtype == SB_I32), SB_ERR_INVALID, "sb_pack_tensor_f64: bad arguments");
  if (n == 0) return SB_OK;
  SB_REQUIRE(d_in && d_wire_values, SB_ERR_INVALID, "sb_pack_tensor_f64: null buffer");
  const int aligned = ((uintptr_t)d_wire_values & 7) == 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (n % 2 == 0 && ((uintptr_t)d_in & 7) == 0 && ((uintptr_t)d_wire_values & 15) == 0) {
    const int64_t n_pair = n / 2;
    if (in_dtype == SB_F32)
      pack_tensor_vec2<float><<<grid_for(n_pair, 4), 256, 0, st>>>((const uint2*)d_in, n_pair, (uint4*)d_wire_values);
    else
      pack_tensor_vec2<int32_t><<<grid_for(n_pair, 4), 256, 0, st>>>((const uint2*)d_in, n_pair, (uint4*)d_wire_values);
    SB_LAUNCH_CHECK();
    return SB_OK;
  }
  if (in_dtype == SB_F32)
    pack_tensor_kernel<float><<<grid_for(n, 4), 256, 0, st>>>((const float*)d_in, n, d_wire_values, aligned);
  else
    pack_tensor_kernel<int32_t><<<grid_for(n, 4), 256, 0, st>>>((const int32_t*)d_in, n, d_wire_values, aligned);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_pack_listvalue(const void* d_in, int32_t in_dtype, int64_t rows, int64_t cols, uint8_t* d_wire,
                                     uintptr_t stream) {
  SB_RANGE("sb_pack_listvalue");
  SB_REQUIRE(rows >= 0 && cols >= 0 && (in_dtype == SB_F32 || in_dtype == SB_I32), SB_ERR_INVALID,
             "sb_pack_listvalue: bad arguments");
  if (rows == 0) return SB_OK;
  SB_REQUIRE(d_in && d_wire, SB_ERR_INVALID, "sb_pack_listvalue: null buffer");
  int64_t rb = 11; int hdr = 0; uint64_t l1 = 0, l2 = 0;
  if (cols > 0) listvalue_layout(cols, &rb, &hdr, &l1, &l2);
  cudaStream_t st = (cudaStream_t)stream;
  const int grid = grid_for(rows * (cols > 0 ? cols : 1), 4);
  const int R = staged_rows(rb);
  if (R > 0 && ((uintptr_t)d_wire & 15) == 0 && cols < (1 << 20)) {
    const size_t smem = (size_t)R * rb + 64;
    int64_t nch = sb::ceil_div(rows, R);
    int g2 = (int)(nch < (int64_t)sb::kNumSMs * 12 ? nch : (int64_t)sb::kNumSMs * 12);
    const int RR = (cols > 0 && cols % 4 == 0) ? ring_rows(rb, kLvPackSlots) : 0;
    if (RR > 0) {
      const int sb = ring_slot_bytes(rb, RR);
      const size_t rsm = (size_t)kLvPackSlots * sb;
      const int64_t nch2 = sb::ceil_div(rows, RR);
      const int per_sm = (int)(200 * 1024 / (rsm + 1024)) < 8 ? (int)(200 * 1024 / (rsm + 1024)) : 8;
      const int64_t gmax = (int64_t)sb::kNumSMs * (per_sm > 0 ? per_sm : 1);
      const int g3 = (int)(nch2 < gmax ? nch2 : gmax);
      if (in_dtype == SB_F32) {
        SB_CUDA(sb::func_attr_cached(pack_listvalue_ring4<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        pack_listvalue_ring4<float><<<g3, 256, rsm, st>>>((const float*)d_in, rows, (int)cols, (int)rb, hdr, (uint32_t)l1,
                                                         (uint32_t)l2, RR, sb, d_wire);
      } else {
        SB_CUDA(sb::func_attr_cached(pack_listvalue_ring4<int32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        pack_listvalue_ring4<int32_t><<<g3, 256, rsm, st>>>((const int32_t*)d_in, rows, (int)cols, (int)rb, hdr,
                                                           (uint32_t)l1, (uint32_t)l2, RR, sb, d_wire);
      }
    } else if (cols > 0 && cols % 4 == 0 && in_dtype == SB_F32) {
      SB_CUDA(sb::func_attr_cached(pack_listvalue_staged4<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      pack_listvalue_staged4<float><<<g2, 256, smem, st>>>((const float*)d_in, rows, (int)cols, (int)rb, hdr, (uint32_t)l1,
                                                          (uint32_t)l2, R, d_wire);
    } else if (cols > 0 && cols % 4 == 0) {
      SB_CUDA(sb::func_attr_cached(pack_listvalue_staged4<int32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      pack_listvalue_staged4<int32_t><<<g2, 256, smem, st>>>((const int32_t*)d_in, rows, (int)cols, (int)rb, hdr,
                                                            (uint32_t)l1, (uint32_t)l2, R, d_wire);
    } else if (in_dtype == SB_F32) {
      SB_CUDA(sb::func_attr_cached(pack_listvalue_staged<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      pack_listvalue_staged<float><<<g2, 256, smem, st>>>((const float*)d_in, rows, (int)cols, (int)rb, hdr, (uint32_t)l1,
                                                         (uint32_t)l2, R, d_wire);
    } else {
      SB_CUDA(sb::func_attr_cached(pack_listvalue_staged<int32_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      pack_listvalue_staged<int32_t><<<g2, 256, smem, st>>>((const int32_t*)d_in, rows, (int)cols, (int)rb, hdr,
                                                           (uint32_t)l1, (uint32_t)l2, R, d_wire);
    }
  } else if (in_dtype == SB_F32) {
    pack_listvalue_kernel<float><<<grid, 256, 0, st>>>((const float*)d_in, rows, cols, rb, hdr, l1, l2, d_wire);
  } else {
    pack_listvalue_kernel<int32_t><<<grid, 256, 0, st>>>((const int32_t*)d_in, rows, cols, rb, hdr, l1, l2, d_wire);
  }
  SB_LAUNCH_CHECK();
  return SB_OK;
}
