// Mahalanobis pass 1: clip + additive statistics, and the covariance update.
//
// Replaces CoreMahalanobis.py:84-100 of the reference (clip, running mean, the (nb,p,p) outer-product tensor
// and its sum) through the Gram identity of SURVEY App. A.2: with a = the start-of-call mean,
//     s = sum_i (xc_i - a),   G = sum_i (xc_i - a)(xc_i - a)^T
// are additive over rows, so every sub-segment of the batch produces one partial (s, G) and
//     sum_i B_i = G - s s^T/nb + n nb/(n+nb) dd^T,  d = s/nb          (L98-99)
//     C1 = (n-1)/den C0 + sum_i B_i / den,  den = n + max(1, nb-1)    (L100)
//     mean1 = a + nb/(n+nb) d                                         (L93 at i = nb-1)
//
// Kernels
//   pass1_gram64_tf32 : d = 64.  One warp per sub-segment streams 8 rows (2 KB) per step straight from HBM
//                       into mma.sync m16n8k8 TF32 fragments (A and B fragments are the same registers because
//                       the Gram is X^T X), 3xTF32 split for fp32-level products, fp32 accumulation, upper
//                       triangle only (20 of 32 tiles).  Algorithmic traffic 4*d B/row; HBM-bound at d = 64.
//   pass1_generic     : any d <= 512, exact fp32 FMA, 4x4 register tiles out of shared memory.
//   reduce_partials   : fp64 sum over sub-segments -> this rank's statistics record.
//   finalize_cov      : combines the records of all ranks in rank order, forms C1, mean1, clip1.
#include "maha.cuh"

namespace sb {

namespace {
SubLayout layout_for(int64_t rows, int64_t target, int64_t minlen) {
  SubLayout L{};
  L.rows = rows;
  L.shortlen = kShortLen;
  L.nshort = rows >= 8 * kShortLen ? 2 : 0;
  int64_t rest = rows - (int64_t)L.nshort * L.shortlen;
  if (rest <= 0) {
    L.sublen = minlen;
    L.nsub = L.nshort;
    return L;
  }
  int64_t m = ceil_div(rest, target * kMaxSubLen);
  int64_t len = ceil_div(rest, target * m);
  len = ceil_div(len, 16) * 16;
  if (len < minlen) len = minlen;
  L.sublen = len;
  L.nsub = L.nshort + (int32_t)ceil_div(rest, len);
  return L;
}
int64_t max_subs_for(int64_t max_rows, int64_t target, int64_t minlen) {
  // layout_for never produces more than target*m + nshort sub-segments, m as computed for max_rows
  int64_t m = ceil_div(max_rows > 0 ? max_rows : 1, target * kMaxSubLen);
  int64_t a = target * m + 4;
  int64_t b = ceil_div(max_rows, minlen) + 4;
  return a < b ? a : b;
}
}  // namespace
SubLayout make_layout(int64_t rows) { return layout_for(rows, kTargetChains, kMinSubLen); }
int64_t max_subsegments(int64_t max_rows) { return max_subs_for(max_rows, kTargetChains, kMinSubLen); }
SubLayout make_layout2(int64_t rows) { return layout_for(rows, kTargetChains2, kMinSubLen2); }
// Pass 1 of the split path: its partials are only ever summed (pass 2 takes its per-sub-segment statistics from the
// projection kernel), so it needs no more sub-segments than keep the persistent CTAs busy and an fp32 accumulation
// short (kMaxSubLen rows): a multiple of the SM count.  At d = 256 a partial is 263 KB - 2368 of them per batch were
// 623 MB written and read back, a third of the input of a 2 Mi-row shard.
SubLayout make_layout1(int64_t rows) { return layout_for(rows, kNumSMs, kMinSubLen); }
int64_t max_subsegments2(int64_t max_rows) { return max_subs_for(max_rows, kTargetChains2, kMinSubLen2); }

namespace {

// ---------------------------------------------------------------------------------------- d = 64, tensor cores
// round to nearest (ties away) to the TF32 mantissa with integer arithmetic: the value cvt.rna.tf32.f32 gives for finite
// inputs, in two instructions (ptxas expands the cvt into about six on sm_100a)
__device__ __forceinline__ uint32_t to_tf32(float x) { return (__float_as_uint(x) + 0x1000u) & 0xffffe000u; }
__device__ __forceinline__ void mma_m16n8k8_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                 uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Logical feature L = g + 8*jj (g = lane/4, jj = 0..7) is physical column phys64(L): each lane loads two float4
// per row (columns 4g..4g+3 and 32+4g..32+4g+3), so a warp instruction covers 128 contiguous bytes of each of 4 rows.
__device__ __forceinline__ int phys64(int L) {
  int g = L & 7, jj = L >> 3;
  return jj < 4 ? 4 * g + jj : 32 + 4 * g + (jj - 4);
}

template <bool CLIP, int TERMS>
__global__ void __launch_bounds__(256, 1)
pass1_gram64_tf32(const float* __restrict__ X, int64_t ld, SubLayout lay, const float* __restrict__ mean_f,
                  const float* __restrict__ lo_f, const float* __restrict__ hi_f, float* __restrict__ partials) {
  constexpr int D = 64, PS = D + D * D;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int warps_total = gridDim.x * (blockDim.x >> 5);
  const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);

  // per-lane constants for its 8 physical columns
  float m[8], lo[8], hi[8];
#pragma unroll
  for (int jj = 0; jj < 8; ++jj) {
    int col = (jj < 4 ? 4 * g + jj : 32 + 4 * g + (jj - 4));
    m[jj] = mean_f[col];
    if (CLIP) { lo[jj] = lo_f[col]; hi[jj] = hi_f[col]; }
  }

  for (int sub = gw; sub < lay.nsub; sub += warps_total) {
    const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
    float acc[4][8][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;
    float ssum[8];
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) ssum[jj] = 0.f;

    float4 nxt[4];
    auto load_chunk = [&](int64_t r, float4 (&v)[4]) {
      const int64_t ra = r + t, rb = r + t + 4;
      const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4* pa = reinterpret_cast<const float4*>(X + ra * ld);
      const float4* pb = reinterpret_cast<const float4*>(X + rb * ld);
      v[0] = ra < r1 ? ldg_stream_f4(pa + g) : z;
      v[1] = ra < r1 ? ldg_stream_f4(pa + 8 + g) : z;
      v[2] = rb < r1 ? ldg_stream_f4(pb + g) : z;
      v[3] = rb < r1 ? ldg_stream_f4(pb + 8 + g) : z;
    };
    if (r0 < r1) load_chunk(r0, nxt);
    for (int64_t r = r0; r < r1; r += 8) {
      float x[2][8];
      x[0][0] = nxt[0].x; x[0][1] = nxt[0].y; x[0][2] = nxt[0].z; x[0][3] = nxt[0].w;
      x[0][4] = nxt[1].x; x[0][5] = nxt[1].y; x[0][6] = nxt[1].z; x[0][7] = nxt[1].w;
      x[1][0] = nxt[2].x; x[1][1] = nxt[2].y; x[1][2] = nxt[2].z; x[1][3] = nxt[2].w;
      x[1][4] = nxt[3].x; x[1][5] = nxt[3].y; x[1][6] = nxt[3].z; x[1][7] = nxt[3].w;
      const bool va = (r + t) < r1, vb = (r + t + 4) < r1;
      if (r + 8 < r1) load_chunk(r + 8, nxt);

      uint32_t big[2][8], sml[2][8];
#pragma unroll
      for (int rr = 0; rr < 2; ++rr) {
        const bool valid = rr == 0 ? va : vb;
#pragma unroll
        for (int jj = 0; jj < 8; ++jj) {
          float v = x[rr][jj];
          if (CLIP) v = fminf(fmaxf(v, lo[jj]), hi[jj]);
          v = valid ? v - m[jj] : 0.f;
          ssum[jj] += v;
          big[rr][jj] = to_tf32(v);
          if (TERMS == 3) sml[rr][jj] = to_tf32(v - __uint_as_float(big[rr][jj]));
        }
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (j >= 2 * i) {
            if (TERMS == 3) {
              mma_m16n8k8_tf32(acc[i][j], sml[0][2 * i], sml[0][2 * i + 1], sml[1][2 * i], sml[1][2 * i + 1],
                               big[0][j], big[1][j]);
              mma_m16n8k8_tf32(acc[i][j], big[0][2 * i], big[0][2 * i + 1], big[1][2 * i], big[1][2 * i + 1],
                               sml[0][j], sml[1][j]);
            }
            mma_m16n8k8_tf32(acc[i][j], big[0][2 * i], big[0][2 * i + 1], big[1][2 * i], big[1][2 * i + 1],
                             big[0][j], big[1][j]);
          }
        }
      }
    }

    // write the partial: s, then G (logical upper triangle mirrored to both halves)
    float* part = partials + (size_t)sub * PS;
#pragma unroll
    for (int jj = 0; jj < 8; ++jj) {
      float v = ssum[jj];
      v += __shfl_xor_sync(0xffffffffu, v, 1);
      v += __shfl_xor_sync(0xffffffffu, v, 2);
      if (t == 0) part[jj < 4 ? 4 * g + jj : 32 + 4 * g + (jj - 4)] = v;
    }
    float* G = part + D;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (j >= 2 * i) {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int Lm = 16 * i + g + ((e & 2) ? 8 : 0);
            const int Ln = 8 * j + 2 * t + (e & 1);
            if (Lm <= Ln) {
              const int pm = phys64(Lm), pn = phys64(Ln);
              G[pm * D + pn] = acc[i][j][e];
              G[pn * D + pm] = acc[i][j][e];
            }
          }
        }
      }
    }
  }
}

// ---------------------------------------------------------------------- d = 128 / 256, tensor cores, blocked
// The D x D Gram is cut into 64 x 64 blocks; only blocks on or above the diagonal are computed.  A warp owns one
// "unit": a diagonal block (20 of its 32 mma tiles, as in pass1_gram64_tf32) or half of an off-diagonal block
// (2 x 8 tiles).  NB*NB units per sub-segment (NB = D/64), 8 units per 256-thread CTA, so D = 256 uses two CTAs per
// sub-segment; all warps of a sub-segment stream the same rows, the repeats hit L1/L2.  At these widths the kernel
// is bound by the (legacy mma.sync) tensor pipe, not by HBM: 2*D*D*(NB+1)/(2*NB) flop per row per TF32 term.
template <int NB, bool CLIP, int TERMS>
__global__ void __launch_bounds__(NB == 2 ? 128 : 256, 1)
pass1_gram_blocked_tf32(const float* __restrict__ X, int64_t ld, SubLayout lay, const float* __restrict__ mean_f,
                        const float* __restrict__ lo_f, const float* __restrict__ hi_f, float* __restrict__ partials) {
  constexpr int D = 64 * NB, PS = D + D * D;
  constexpr int UNITS = NB * NB, CTAS_PER_SUB = (UNITS + 7) / 8;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int unit = (blockIdx.x % CTAS_PER_SUB) * 8 + warp;
  if (unit >= UNITS) return;
  // decode the unit
  int bi, bj, hh = 0;
  const bool diag = unit < NB;
  if (diag) { bi = bj = unit; }
  else {
    int o = unit - NB;
    hh = o & 1;
    int p = o >> 1;  // p-th pair (bi < bj) in row-major order
    bi = 0;
    while (p >= NB - 1 - bi) { p -= NB - 1 - bi; ++bi; }
    bj = bi + 1 + p;
  }
  const int ca = 64 * bi, cb = 64 * bj;
  const int subs_stride = gridDim.x / CTAS_PER_SUB;

  for (int sub = blockIdx.x / CTAS_PER_SUB; sub < lay.nsub; sub += subs_stride) {
    const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
    float* part = partials + (size_t)sub * PS;
    float* G = part + D;
    if (diag) {
      // ---- diagonal block: identical to the d = 64 kernel on columns ca .. ca+63
      float m[8], lo[8], hi[8];
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        const int col = ca + (jj < 4 ? 4 * g + jj : 32 + 4 * g + (jj - 4));
        m[jj] = mean_f[col];
        if (CLIP) { lo[jj] = lo_f[col]; hi[jj] = hi_f[col]; }
      }
      float acc[4][8][4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;
      float ssum[8];
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) ssum[jj] = 0.f;
      for (int64_t r = r0; r < r1; r += 8) {
        const int64_t ra = r + t, rb = r + t + 4;
        const bool va = ra < r1, vb = rb < r1;
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const float4* pa = reinterpret_cast<const float4*>(X + ra * ld + ca);
        const float4* pb = reinterpret_cast<const float4*>(X + rb * ld + ca);
        const float4 v0 = va ? __ldg(pa + g) : z4, v1 = va ? __ldg(pa + 8 + g) : z4;
        const float4 v2 = vb ? __ldg(pb + g) : z4, v3 = vb ? __ldg(pb + 8 + g) : z4;
        const float x[2][8] = {{v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w},
                               {v2.x, v2.y, v2.z, v2.w, v3.x, v3.y, v3.z, v3.w}};
        uint32_t big[2][8], sml[2][8];
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const bool valid = rr == 0 ? va : vb;
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) {
            float v = x[rr][jj];
            if (CLIP) v = fminf(fmaxf(v, lo[jj]), hi[jj]);
            v = valid ? v - m[jj] : 0.f;
            ssum[jj] += v;
            big[rr][jj] = to_tf32(v);
            if (TERMS == 3) sml[rr][jj] = to_tf32(v - __uint_as_float(big[rr][jj]));
          }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j >= 2 * i) {
              if (TERMS == 3) {
                mma_m16n8k8_tf32(acc[i][j], sml[0][2 * i], sml[0][2 * i + 1], sml[1][2 * i], sml[1][2 * i + 1], big[0][j], big[1][j]);
                mma_m16n8k8_tf32(acc[i][j], big[0][2 * i], big[0][2 * i + 1], big[1][2 * i], big[1][2 * i + 1], sml[0][j], sml[1][j]);
              }
              mma_m16n8k8_tf32(acc[i][j], big[0][2 * i], big[0][2 * i + 1], big[1][2 * i], big[1][2 * i + 1], big[0][j], big[1][j]);
            }
      }
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        float v = ssum[jj];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        if (t == 0) part[ca + (jj < 4 ? 4 * g + jj : 32 + 4 * g + (jj - 4))] = v;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j >= 2 * i) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int Lm = 16 * i + g + ((e & 2) ? 8 : 0), Ln = 8 * j + 2 * t + (e & 1);
              if (Lm <= Ln) {
                const int pm = ca + phys64(Lm), pn = ca + phys64(Ln);
                G[(size_t)pm * D + pn] = acc[i][j][e];
                G[(size_t)pn * D + pm] = acc[i][j][e];
              }
            }
          }
    } else {
      // ---- half of an off-diagonal block: m-tiles {2hh, 2hh+1} of block bi against all 8 n-tiles of block bj
      float ma[4], loa[4], hia[4], mb[8], lob[8], hib[8];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int col = ca + 32 * hh + 4 * g + q;
        ma[q] = mean_f[col];
        if (CLIP) { loa[q] = lo_f[col]; hia[q] = hi_f[col]; }
      }
#pragma unroll
      for (int jj = 0; jj < 8; ++jj) {
        const int col = cb + (jj < 4 ? 4 * g + jj : 32 + 4 * g + (jj - 4));
        mb[jj] = mean_f[col];
        if (CLIP) { lob[jj] = lo_f[col]; hib[jj] = hi_f[col]; }
      }
      float acc[2][8][4];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[i][j][e] = 0.f;
      for (int64_t r = r0; r < r1; r += 8) {
        const int64_t ra = r + t, rb = r + t + 4;
        const bool va = ra < r1, vb = rb < r1;
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const float* rowa = X + ra * ld;
        const float* rowb = X + rb * ld;
        const float4 a0 = va ? __ldg(reinterpret_cast<const float4*>(rowa + ca + 32 * hh) + g) : z4;
        const float4 a1 = vb ? __ldg(reinterpret_cast<const float4*>(rowb + ca + 32 * hh) + g) : z4;
        const float4 b0 = va ? __ldg(reinterpret_cast<const float4*>(rowa + cb) + g) : z4;
        const float4 b1 = va ? __ldg(reinterpret_cast<const float4*>(rowa + cb) + 8 + g) : z4;
        const float4 b2 = vb ? __ldg(reinterpret_cast<const float4*>(rowb + cb) + g) : z4;
        const float4 b3 = vb ? __ldg(reinterpret_cast<const float4*>(rowb + cb) + 8 + g) : z4;
        const float xa[2][4] = {{a0.x, a0.y, a0.z, a0.w}, {a1.x, a1.y, a1.z, a1.w}};
        const float xb[2][8] = {{b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w},
                                {b2.x, b2.y, b2.z, b2.w, b3.x, b3.y, b3.z, b3.w}};
        uint32_t bigA[2][4], smlA[2][4], bigB[2][8], smlB[2][8];
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const bool valid = rr == 0 ? va : vb;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            float v = xa[rr][q];
            if (CLIP) v = fminf(fmaxf(v, loa[q]), hia[q]);
            v = valid ? v - ma[q] : 0.f;
            bigA[rr][q] = to_tf32(v);
            if (TERMS == 3) smlA[rr][q] = to_tf32(v - __uint_as_float(bigA[rr][q]));
          }
#pragma unroll
          for (int jj = 0; jj < 8; ++jj) {
            float v = xb[rr][jj];
            if (CLIP) v = fminf(fmaxf(v, lob[jj]), hib[jj]);
            v = valid ? v - mb[jj] : 0.f;
            bigB[rr][jj] = to_tf32(v);
            if (TERMS == 3) smlB[rr][jj] = to_tf32(v - __uint_as_float(bigB[rr][jj]));
          }
        }
#pragma unroll
        for (int il = 0; il < 2; ++il)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (TERMS == 3) {
              mma_m16n8k8_tf32(acc[il][j], smlA[0][2 * il], smlA[0][2 * il + 1], smlA[1][2 * il], smlA[1][2 * il + 1], bigB[0][j], bigB[1][j]);
              mma_m16n8k8_tf32(acc[il][j], bigA[0][2 * il], bigA[0][2 * il + 1], bigA[1][2 * il], bigA[1][2 * il + 1], smlB[0][j], smlB[1][j]);
            }
            mma_m16n8k8_tf32(acc[il][j], bigA[0][2 * il], bigA[0][2 * il + 1], bigA[1][2 * il], bigA[1][2 * il + 1], bigB[0][j], bigB[1][j]);
          }
      }
#pragma unroll
      for (int il = 0; il < 2; ++il)
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int Lm = 16 * (2 * hh + il) + g + ((e & 2) ? 8 : 0), Ln = 8 * j + 2 * t + (e & 1);
            const int pm = ca + phys64(Lm), pn = cb + phys64(Ln);
            G[(size_t)pm * D + pn] = acc[il][j][e];
            G[(size_t)pn * D + pm] = acc[il][j][e];
          }
    }
  }
}

// --------------------------------------------------------------------------------------------- generic, fp32
constexpr int kGenRows = 16;

template <bool CLIP>
__global__ void __launch_bounds__(256)
pass1_generic(const float* __restrict__ X, int64_t ld, int d, SubLayout lay, const float* __restrict__ mean_f,
              const float* __restrict__ lo_f, const float* __restrict__ hi_f, float* __restrict__ partials) {
  extern __shared__ float xs[];  // [kGenRows][dp]
  const int dp = (d + 3) & ~3;
  const int T4 = dp >> 2;
  const size_t PS = (size_t)d + (size_t)d * d;
  for (int sub = blockIdx.x; sub < lay.nsub; sub += gridDim.x) {
    const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
    float* part = partials + (size_t)sub * PS;
    bool first = true;
    for (int64_t r = r0; r < r1 || first; r += kGenRows) {
      __syncthreads();
      for (int idx = threadIdx.x; idx < kGenRows * dp; idx += blockDim.x) {
        int rr = idx / dp, f = idx - rr * dp;
        float v = 0.f;
        if (f < d && r + rr < r1) {
          v = X[(r + rr) * ld + f];
          if (CLIP) v = fminf(fmaxf(v, lo_f[f]), hi_f[f]);
          v -= mean_f[f];
        }
        xs[idx] = v;
      }
      __syncthreads();
      for (int f = threadIdx.x; f < d; f += blockDim.x) {
        float a = 0.f;
#pragma unroll
        for (int rr = 0; rr < kGenRows; ++rr) a += xs[rr * dp + f];
        part[f] = first ? a : part[f] + a;
      }
      for (int tile = threadIdx.x; tile < T4 * T4; tile += blockDim.x) {
        const int ta = tile / T4, tb = tile - ta * T4;
        if (ta > tb) continue;
        float acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
#pragma unroll 4
        for (int rr = 0; rr < kGenRows; ++rr) {
          const float4 va = *reinterpret_cast<const float4*>(xs + rr * dp + 4 * ta);
          const float4 vb = *reinterpret_cast<const float4*>(xs + rr * dp + 4 * tb);
          const float a4[4] = {va.x, va.y, va.z, va.w}, b4[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
          for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = fmaf(a4[a], b4[b], acc[a][b]);
        }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const int fa = 4 * ta + a, fb = 4 * tb + b;
            if (fa < d && fb < d && fa <= fb) {
              float* p = part + d + (size_t)fa * d + fb;
              const float v = first ? acc[a][b] : *p + acc[a][b];
              *p = v;
              part[d + (size_t)fb * d + fa] = v;  // keep G fully symmetric in memory
            }
          }
      }
      first = false;
    }
  }
}

// stats[0] = rows, stats[1..d] = s, stats[1+d ..] = G (full, symmetric).  Partials hold G's upper triangle.
__global__ void __launch_bounds__(256)
reduce_partials(const float* __restrict__ partials, int nsub, int d, int64_t rows, double* __restrict__ stats) {
  const size_t PS = (size_t)d + (size_t)d * d;
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e == 0) stats[0] = (double)rows;
  if (e >= PS) return;
  size_t src = e;
  if (e >= (size_t)d) {
    size_t a = (e - d) / d, b = (e - d) - a * d;
    if (a > b) src = d + b * d + a;
  }
  double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
  int s = 0;
  for (; s + 3 < nsub; s += 4) {
    acc0 += (double)partials[(size_t)s * PS + src];
    acc1 += (double)partials[(size_t)(s + 1) * PS + src];
    acc2 += (double)partials[(size_t)(s + 2) * PS + src];
    acc3 += (double)partials[(size_t)(s + 3) * PS + src];
  }
  for (; s < nsub; ++s) acc0 += (double)partials[(size_t)s * PS + src];
  stats[1 + e] = (acc0 + acc1) + (acc2 + acc3);
}

// Two-stage form for the main path: stage 1 sums a slice of the sub-segments per (element block, slice) CTA into
// slices[slice][PS]; stage 2 adds the slices in a fixed order (deterministic, no atomics).
__global__ void __launch_bounds__(256)
reduce_partials_stage1(const float* __restrict__ partials, int nsub, size_t PS, int nslice, double* __restrict__ slices) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= PS) return;
  const int per = (nsub + nslice - 1) / nslice;
  const int s0 = blockIdx.y * per, s1 = min(nsub, s0 + per);
  double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
  int s = s0;
  for (; s + 3 < s1; s += 4) {
    acc0 += (double)partials[(size_t)s * PS + e];
    acc1 += (double)partials[(size_t)(s + 1) * PS + e];
    acc2 += (double)partials[(size_t)(s + 2) * PS + e];
    acc3 += (double)partials[(size_t)(s + 3) * PS + e];
  }
  for (; s < s1; ++s) acc0 += (double)partials[(size_t)s * PS + e];
  slices[(size_t)blockIdx.y * PS + e] = (acc0 + acc1) + (acc2 + acc3);
}

__global__ void __launch_bounds__(256)
reduce_partials_stage2(const double* __restrict__ slices, size_t PS, int nslice, int64_t rows, double* __restrict__ stats) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e == 0) stats[0] = (double)rows;
  if (e >= PS) return;
  double acc = 0.0;
  for (int k = 0; k < nslice; ++k) acc += slices[(size_t)k * PS + e];
  stats[1 + e] = acc;
}

// One thread per covariance element.  all_stats: [nranks][1 + d + d*d].
__global__ void __launch_bounds__(256)
finalize_cov(const double* __restrict__ all_stats, int nranks, int rank, int d, long long n_eff, long long n_new_total,
             long long start_clip, double n_stdev, const double* __restrict__ mean0, const double* __restrict__ C0,
             double* __restrict__ mean1, double* __restrict__ C1, double* __restrict__ lo1, double* __restrict__ hi1,
             double* __restrict__ rank_pre, long long* __restrict__ gate) {
  const size_t SL = 1 + (size_t)d + (size_t)d * d;
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;  // index into the d*d part
  if (e >= (size_t)d * d) return;
  const int a = (int)(e / d), b = (int)(e - (size_t)a * d);
  double nb = 0.0, sa = 0.0, sb_ = 0.0, gab = 0.0, pre_g = 0.0, pre_cnt = 0.0, pre_sb = 0.0;
  for (int r = 0; r < nranks; ++r) {
    const double* st = all_stats + (size_t)r * SL;
    if (r == rank) { pre_g = gab; pre_cnt = nb; pre_sb = sb_; }
    nb += st[0];
    sa += st[1 + a];
    sb_ += st[1 + b];
    gab += st[1 + d + e];
  }
  rank_pre[1 + d + e] = pre_g;
  if (a == 0) rank_pre[1 + b] = pre_sb;
  if (e == 0) { rank_pre[0] = pre_cnt; gate[4] = (long long)(nb + 0.5); }
  const double n = (double)n_eff;
  const double da = sa / nb, db = sb_ / nb;
  double sumB = gab - sa * sb_ / nb;
  if (n + nb > 0.0) sumB += n * nb / (n + nb) * da * db;
  const double den = n + fmax(1.0, nb - 1.0);
  const double c1 = (n - 1.0) / den * C0[e] + sumB / den;
  C1[e] = c1;
  if (a == b) {
    const double m1 = mean0[a] + nb / (n + nb) * da;
    mean1[a] = m1;
    if (n_new_total > start_clip) {
      const double sd = sqrt(c1);
      lo1[a] = m1 - n_stdev * sd;
      hi1[a] = m1 + n_stdev * sd;
    }
  }
}

}  // namespace

int launch_pass1(const sb_maha* h, const float* X, int64_t ld, const MahaWs& ws, cudaStream_t st) {
  const SubLayout& L = h->lay;
  if (L.nsub == 0) return SB_OK;
  const bool fast = h->d == 64 && !h->force_generic && (ld % 4 == 0) && (((uintptr_t)X & 15) == 0);
  const bool blocked = (h->d == 128 || h->d == 256) && !h->force_generic && (ld % 4 == 0) && (((uintptr_t)X & 15) == 0);
  if (h->profile) cudaEventRecord(h->ev[0], st);
  int umma_rc = 1;
  if (!h->force_generic && !h->legacy_mma) {
    umma_rc = launch_pass1_umma(h, X, ld, ws, st);  // tcgen05 path for d = 64 / 128 / 256
    if (umma_rc < 0) return umma_rc;
  }
  if (umma_rc == SB_OK) {
    // launched above
  } else if (blocked) {
    const int per_sub = h->d == 256 ? 2 : 1;
    int subs = L.nsub < kNumSMs * 2 ? L.nsub : kNumSMs * 2;
    const int grid = subs * per_sub;
#define SB_LAUNCH_BLOCKED(NB, THREADS)                                                                            \
    do {                                                                                                          \
      if (h->use_clip) {                                                                                          \
        if (h->gram_terms == 3) pass1_gram_blocked_tf32<NB, true, 3><<<grid, THREADS, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials); \
        else pass1_gram_blocked_tf32<NB, true, 1><<<grid, THREADS, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials);                    \
      } else {                                                                                                    \
        if (h->gram_terms == 3) pass1_gram_blocked_tf32<NB, false, 3><<<grid, THREADS, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials); \
        else pass1_gram_blocked_tf32<NB, false, 1><<<grid, THREADS, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials);                   \
      }                                                                                                           \
    } while (0)
    if (h->d == 256) SB_LAUNCH_BLOCKED(4, 256); else SB_LAUNCH_BLOCKED(2, 128);
#undef SB_LAUNCH_BLOCKED
  } else if (fast) {
    int grid = (int)ceil_div(L.nsub, 8);
    if (grid > kNumSMs) grid = kNumSMs;
    if (h->use_clip) {
      if (h->gram_terms == 3)
        pass1_gram64_tf32<true, 3><<<grid, 256, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials);
      else
        pass1_gram64_tf32<true, 1><<<grid, 256, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials);
    } else {
      if (h->gram_terms == 3)
        pass1_gram64_tf32<false, 3><<<grid, 256, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials);
      else
        pass1_gram64_tf32<false, 1><<<grid, 256, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials);
    }
  } else {
    int grid = L.nsub < kNumSMs * 4 ? L.nsub : kNumSMs * 4;
    size_t smem = (size_t)kGenRows * ((h->d + 3) & ~3) * sizeof(float);
    if (h->use_clip)
      pass1_generic<true><<<grid, 256, smem, st>>>(X, ld, h->d, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials);
    else
      pass1_generic<false><<<grid, 256, smem, st>>>(X, ld, h->d, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, ws.partials);
  }
  if (h->profile) cudaEventRecord(h->ev[1], st);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

int reduce_slices(int d) {
  const size_t PS = (size_t)d + (size_t)d * d;
  const int64_t blocks = ceil_div(PS, 256);
  int64_t n = ceil_div((int64_t)kNumSMs * 8, blocks);
  return (int)(n < 1 ? 1 : (n > 64 ? 64 : n));
}

int launch_reduce(const sb_maha* h, const MahaWs& ws, double* stats_out, cudaStream_t st) {
  const size_t PS = (size_t)h->d + (size_t)h->d * h->d;
  const int nslice = reduce_slices(h->d);
  dim3 grid((unsigned)ceil_div(PS, 256), (unsigned)nslice);
  reduce_partials_stage1<<<grid, 256, 0, st>>>(ws.partials, h->lay.nsub, PS, nslice, ws.slices);
  SB_LAUNCH_CHECK();
  reduce_partials_stage2<<<(int)ceil_div(PS, 256), 256, 0, st>>>(ws.slices, PS, nslice, h->rows_local, stats_out);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

int launch_finalize_cov(sb_maha* h, const double* all_stats, int nranks, int rank, const MahaWs& ws, cudaStream_t st) {
  const int d = h->d;
  finalize_cov<<<(int)ceil_div((int64_t)d * d, 256), 256, 0, st>>>(
      all_stats, nranks, rank, d, (long long)h->n_eff, (long long)(h->n + h->rows_total), (long long)h->start_clip,
      h->n_stdev, h->d_mean, h->d_C, h->d_mean1, h->d_C1, h->d_lo1, h->d_hi1, ws.rank_pre, h->d_gate);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

// ------------------------------------------------------------------------------------------------------------------
// sb_gram_stats_noalloc: the pass-1 statistics for callers that cannot hand over (or create) a device workspace.
//
// Who that is: (1) thin bindings that have no device allocator of their own - a JNI or cgo shim around the engine's
// combiner/transformer units only maps the request matrix and an output record, there is no torch caching allocator
// behind it to carve the sub-segment partials from; (2) request threads of a serving process that must keep
// cudaMalloc/cudaFree off the request path, because those calls synchronise the whole device and stall every other
// request stream.  For them the scratch that sb_maha_pass1_stats takes from the caller lives here, in a file-scope
// device array that every GPU loading this library carries until the process exits.
//
// Limits (both checked, SB_ERR_INVALID otherwise): d <= 64 features and rows <= 1 048 576 per call, i.e. at most
// 256 sub-segment partials of 64 + 64*64 floats = 4 259 840 bytes of static device memory per GPU.
//
// Sharing rules: there is one array per GPU and every call on that GPU uses it.  Nothing has to be cleared between
// calls - the statistics kernels store (never accumulate into) every slot that the reduction afterwards reads, so
// whatever the previous call left behind is overwritten before it is looked at - but two calls that overlap on the
// same GPU (different streams, or different host threads) race on the array; serialise them on one stream.
// The first call on a device looks up the array's device address; make that first call outside any stream capture.
namespace {

constexpr int kNoAllocMaxD = 64;
constexpr int kNoAllocMaxSub = 256;
constexpr int64_t kNoAllocMaxRows = (int64_t)kNoAllocMaxSub * kMaxSubLen;  // 1 048 576
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) float g_noalloc_partials[(size_t)kNoAllocMaxSub * (kNoAllocMaxD + kNoAllocMaxD * kNoAllocMaxD)];

// In host code `g_noalloc_partials` is only a stand-in for the symbol; the address a kernel can use has to be asked
// for, once per device, with that device current.
float* g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};

int noalloc_scratch(float** out) {
  int dev = -1;
  SB_CUDA(cudaGetDevice(&dev));
  SB_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, SB_ERR_INVALID, "sb_gram_stats_noalloc: device ordinal %d not supported", dev);
  float* p = __atomic_load_n(&g_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    void* sym = nullptr;
    SB_CUDA(cudaGetSymbolAddress(&sym, g_noalloc_partials));
    p = (float*)sym;
    __atomic_store_n(&g_noalloc_ptr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return SB_OK;
}

}  // namespace
}  // namespace sb

extern "C" int32_t sb_gram_stats_noalloc(const float* d_X, int64_t rows, int32_t d, int64_t ld, const float* d_shift,
                                         double* d_stats_out, uintptr_t stream) {
  SB_RANGE("sb_gram_stats_noalloc");
  using namespace sb;
  SB_REQUIRE(d_X && d_shift && d_stats_out, SB_ERR_INVALID, "sb_gram_stats_noalloc: null pointer");
  SB_REQUIRE(d >= 1 && d <= kNoAllocMaxD, SB_ERR_INVALID,
             "sb_gram_stats_noalloc: d=%d exceeds the static scratch limit of %d features", d, kNoAllocMaxD);
  SB_REQUIRE(rows >= 1 && rows <= kNoAllocMaxRows, SB_ERR_INVALID,
             "sb_gram_stats_noalloc: rows=%lld outside 1..%lld (static scratch holds %d sub-segment partials); use "
             "sb_maha_pass1_stats with a caller workspace for larger batches",
             (long long)rows, (long long)kNoAllocMaxRows, kNoAllocMaxSub);
  SB_REQUIRE(ld >= d, SB_ERR_INVALID, "sb_gram_stats_noalloc: ld < d");
  float* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // at most kNoAllocMaxSub sub-segments of equal length (a multiple of 16 rows, at least kMinSubLen)
  SubLayout L{};
  L.rows = rows;
  L.nshort = 0;
  L.shortlen = kShortLen;
  int64_t len = ceil_div(ceil_div(rows, kNoAllocMaxSub), 16) * 16;
  if (len < kMinSubLen) len = kMinSubLen;
  L.sublen = len;
  L.nsub = (int32_t)ceil_div(rows, len);
  const bool fast = d == 64 && (ld % 4 == 0) && (((uintptr_t)d_X & 15) == 0);
  if (fast) {
    int grid = (int)ceil_div(L.nsub, 8);
    pass1_gram64_tf32<false, 3><<<grid, 256, 0, st>>>(d_X, ld, L, d_shift, d_shift, d_shift, scratch);
  } else {
    size_t smem = (size_t)kGenRows * ((d + 3) & ~3) * sizeof(float);
    pass1_generic<false><<<L.nsub, 256, smem, st>>>(d_X, ld, d, L, d_shift, d_shift, d_shift, scratch);
  }
  SB_LAUNCH_CHECK();
  const size_t PS = (size_t)d + (size_t)d * d;
  reduce_partials<<<(int)ceil_div(PS, 256), 256, 0, st>>>(scratch, L.nsub, d, rows, d_stats_out);
  SB_LAUNCH_CHECK();
  return SB_OK;
}
