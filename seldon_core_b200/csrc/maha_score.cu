// Mahalanobis pass 2: projections, warm-up gate, inverse chain, score and flag per row.
//
// Replaces CoreMahalanobis.py:106-136 and 147-149 of the reference.  With V the top-k eigenvectors of the
// end-of-batch covariance, a = the start-of-call mean, z_i = V^T(clip(x_i) - a), y_i = V^T(x_i - a) and the plain
// prefix sums Z_i = sum_{j<i} z_j, Q_i = sum_{j<i} z_j z_j^T:
//     S_i = (n-1) V^T C0 V + Q_i - Z_i Z_i^T/(n+i),   u_i = y_i - Z_i/(n+i),
//     score_i = (n+i-1) u_i^T S_i^-1 u_i   once the gate |det(S_i/(n+i-1))| > 1e-8 has opened, else 0   (L122-136,148)
//     flag_i  = score_i > threshold                                                                     (L149)
// The reference walks the rows with a Sherman-Morrison update of the inverse; here the batch is cut into
// sub-segments, the prefix (Z, Q) at every sub-segment start comes from projecting the pass-1 partials
// (V^T s, V^T G V) and scanning them, the start inverse is computed in fp64, and inside a sub-segment a group of
// KP lanes (lane j holds row j of S^-1) applies S_{i+1} = S_i + (n+i)/(n+i+1) e_i e_i^T, e_i = z_i - Z_i/(n+i),
// to the inverse in fp32.  Sub-segments whose start matrix is not yet invertible or is ill conditioned (cold
// start) go through an exact fp64 kernel that factorises S_i for every row and evaluates the gate.
//
// Kernels: sub_project, sub_scan, sub_start, gate_init (prefix machinery, O(nsub k^2 d) work),
//          score_slow (fp64 warm-up), slow_fixup, score_chain<KP> (generic projection) and
//          score_chain64 (d = 64, k <= 16: projection on mma.sync TF32 with a 3xTF32 split).
// The scoring kernels read X once: 4*d + 8 B/row.
#include <limits.h>

#include "maha.cuh"

namespace sb {
namespace {

constexpr double kGateEps = 1e-8;     // CoreMahalanobis.py:120
// Rank guard of the gate.  The oracle's guard sits at fp64 rounding level (1e-12); here S is built from fp32 data
// (pass-1 Gram, fp32 inputs), so a singular matrix shows pivots near 1e-7 x the largest and anything below 1e-5 is
// indistinguishable from singular.  Matrices with a pivot ratio between the two tolerances are outside the parity
// contract ("warm-up gate").
constexpr double kRankTol = 1e-5;
constexpr double kWellTol = 1e-4;     // start matrices worse than this go to the fp64 kernel

__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xffffffffu, v, src); }

// -------------------------------------------------------------------------------------------- sub_project
// out[item][0..KP) = V^T s,  out[item][KP + a*KP + b] = (V^T G V)_ab  for the (s, G) record of each item.
// fp32 products with fp32 accumulation over <= d terms (the records themselves are fp32-accurate); the prefix sums
// over sub-segments that follow are fp64.  One CTA per item: G is staged through shared memory in slabs of FC rows
// (coalesced), T = G V per slab, then V^T T.
template <typename TIn>
__global__ void __launch_bounds__(256)
sub_project(const TIn* __restrict__ recs, size_t rec_stride, int nitems, int d, int KP, const float* __restrict__ Vf,
            double* __restrict__ out) {
  extern __shared__ __align__(16) float sp_smem[];
  const int FC = min(d, 8192 / d);           // rows of G per slab (slab <= 32 KB)
  float* Gs = sp_smem;                        // [FC][d]
  float* Ts = Gs + (((size_t)FC * d + 3) & ~(size_t)3);  // [FC][KP], 16-byte aligned
  const int KQ = KP + KP * KP;
  const int tid = threadIdx.x;
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const TIn* s = recs + (size_t)item * rec_stride;
    const TIn* G = s + d;
    double* o = out + (size_t)item * KQ;
    float qacc[4];  // KP*KP <= 1024 = 4 * 256
#pragma unroll
    for (int u = 0; u < 4; ++u) qacc[u] = 0.f;
    for (int f0 = 0; f0 < d; f0 += FC) {
      const int nf = min(FC, d - f0);
      __syncthreads();
      for (int idx = tid; idx < nf * d; idx += blockDim.x) Gs[idx] = (float)G[(size_t)f0 * d + idx];
      __syncthreads();
      // T = G V for this slab: each thread owns 4 consecutive components of one row (V rows read as float4)
      const int KP4 = KP >> 2;
      for (int idx = tid; idx < nf * KP4; idx += blockDim.x) {
        const int ff = idx / KP4, c4 = idx - ff * KP4;
        const float* grow = Gs + (size_t)ff * d;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int g = 0; g < d; ++g) {
          const float gv = grow[g];
          const float4 v = __ldg(reinterpret_cast<const float4*>(Vf + (size_t)g * KP) + c4);
          acc.x = fmaf(gv, v.x, acc.x);
          acc.y = fmaf(gv, v.y, acc.y);
          acc.z = fmaf(gv, v.z, acc.z);
          acc.w = fmaf(gv, v.w, acc.w);
        }
        reinterpret_cast<float4*>(Ts)[idx] = acc;
      }
      __syncthreads();
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int pq = tid + u * 256;
        if (pq < KP * KP) {
          const int a = pq / KP, b = pq - a * KP;
          float acc = qacc[u];
          for (int ff = 0; ff < nf; ++ff) acc = fmaf(__ldg(Vf + (size_t)(f0 + ff) * KP + a), Ts[ff * KP + b], acc);
          qacc[u] = acc;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int pq = tid + u * 256;
      if (pq < KP * KP) o[KP + pq] = (double)qacc[u];
    }
    if (tid < KP) {
      float acc = 0.f;
      for (int f = 0; f < d; ++f) acc = fmaf((float)s[f], __ldg(Vf + (size_t)f * KP + tid), acc);
      o[tid] = (double)acc;
    }
  }
}

// Exclusive scan over sub-segments of every projected component, seeded with the rank offset; row nsub = totals.
// Two kernels, both coalesced over the component index: (1) totals of chunks of 32 sub-segments, (2) every chunk
// adds up the totals of the chunks before it (a few hundred at most) and rewrites its 32 rows.
constexpr int kScanChunk = 32;

__global__ void __launch_bounds__(1024)
sub_scan_totals(const double* __restrict__ subproj, int nsub, int KQ, double* __restrict__ chunk_tot) {
  const int chunk = blockIdx.x;
  const int s0 = chunk * kScanChunk, s1 = min(nsub, s0 + kScanChunk);
  for (int comp = threadIdx.x; comp < KQ; comp += blockDim.x) {
    double acc = 0.0;
    for (int s = s0; s < s1; ++s) acc += subproj[(size_t)s * KQ + comp];
    chunk_tot[(size_t)chunk * KQ + comp] = acc;
  }
}

// Every chunk forms its own exclusive prefix - the rank offset plus the totals of the chunks before it, added in chunk
// order (the same sequence of additions in every CTA, so the prefixes are those of one running sum) - and rewrites its
// 32 rows.  The totals are a few hundred rows of KQ doubles that stay in L2; reading them again per chunk is cheaper
// than the single-CTA scan kernel that used to sit between the two (77 us cold on the critical path of every batch).
__global__ void __launch_bounds__(1024)
sub_scan_apply(double* __restrict__ subproj, int nsub, int KQ, const double* __restrict__ chunk_tot,
               const double* __restrict__ rank_off) {
  const int chunk = blockIdx.x, nchunk = gridDim.x;
  const int s0 = chunk * kScanChunk, s1 = min(nsub, s0 + kScanChunk);
  for (int comp = threadIdx.x; comp < KQ; comp += blockDim.x) {
    double run = rank_off[comp];
#pragma unroll 8
    for (int c = 0; c < chunk; ++c) run += chunk_tot[(size_t)c * KQ + comp];
    for (int s = s0; s < s1; ++s) {
      const double v = subproj[(size_t)s * KQ + comp];
      subproj[(size_t)s * KQ + comp] = run;
      run += v;
    }
    if (chunk == nchunk - 1) subproj[(size_t)nsub * KQ + comp] = run;
  }
}

// -------------------------------------------------------------------------------------------- sub_start
// One warp per sub-segment: S at the first row, Gauss-Jordan inverse in fp64 (S is SPD: no pivoting), flags.
template <int KP>
__global__ void __launch_bounds__(128)
sub_start(const double* __restrict__ subproj, const double* __restrict__ PC0, SubLayout lay, int k, long long n_eff,
          long long row0, float* __restrict__ sub_P, float* __restrict__ sub_Z, int32_t* __restrict__ sub_flags) {
  constexpr int KQ = KP + KP * KP;
  const int lane = threadIdx.x & 31;
  const int sub = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (sub >= lay.nsub) return;
  const int a = lane < KP ? lane : KP - 1;  // lanes >= KP mirror the last row (results discarded)
  const double* pz = subproj + (size_t)sub * KQ;
  const long long i0 = row0 + sub_begin(lay, sub);
  const double ni = (double)(n_eff + i0);
  const double za = pz[a];
  double row[2 * KP];
#pragma unroll
  for (int b = 0; b < KP; ++b) {
    double v = (double)(n_eff - 1) * PC0[a * KP + b];
    if (n_eff + i0 > 0) v += pz[KP + a * KP + b] - za * pz[b] / ni;
    if (a >= k || b >= k) v = (a == b) ? 1.0 : 0.0;
    row[b] = v;
    row[KP + b] = (a == b) ? 1.0 : 0.0;
  }
  double pmin = INFINITY, pmax = 0.0, det = 1.0;
  bool ok = true;
  double div = 1.0;
  bool testable = true;
  if (i0 == 0) { div = (double)(n_eff - 1); testable = n_eff - 1 > 0; }
  else if (n_eff + i0 - 1 == 0) testable = false;
  else div = ni - 1.0;
#pragma unroll
  for (int c = 0; c < KP; ++c) {
    const double piv = shfl_d(row[c], c);
    if (!(piv > 0.0) || !isfinite(piv)) { ok = false; break; }  // warp-uniform
    if (c < k) {
      pmin = fmin(pmin, piv);
      pmax = fmax(pmax, piv);
      det *= piv / div;
    }
    const double f = (a == c) ? 0.0 : row[c] / piv;
    const double rp = 1.0 / piv;
#pragma unroll
    for (int j = 0; j < 2 * KP; ++j) {
      const double pj = shfl_d(row[j], c);
      row[j] = (a == c) ? pj * rp : fma(-f, pj, row[j]);
    }
  }
  int flags = 0;
  if (ok && pmin >= kWellTol * pmax) flags |= SUBFLAG_WELL;
  if (ok && testable && pmin > kRankTol * pmax && fabs(det) > kGateEps) flags |= SUBFLAG_OPEN;
  if (lane < KP) {
#pragma unroll
    for (int b = 0; b < KP; ++b) sub_P[(size_t)sub * KP * KP + lane * KP + b] = ok ? (float)row[KP + b] : 0.f;
    sub_Z[(size_t)sub * KP + lane] = (float)za;
  }
  if (lane == 0) sub_flags[sub] = flags;
}

// gate[0] = b* (first sub-segment whose start passes the gate test), gate[1] = i* (global row, provisional),
// gate[2] = flagged rows, score_sum = 0
__global__ void gate_init(const int32_t* __restrict__ sub_flags, SubLayout lay, long long row0, long long* gate,
                          double* score_sum) {
  const int lane = threadIdx.x;
  int best = INT_MAX;
  for (int s = lane; s < lay.nsub; s += 32)
    if (sub_flags[s] & SUBFLAG_OPEN) { best = s; break; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (lane == 0) {
    const int bstar = best == INT_MAX ? lay.nsub : best;
    gate[0] = bstar;
    gate[1] = bstar < lay.nsub ? row0 + sub_begin(lay, bstar) : LLONG_MAX;
    gate[2] = 0;
    *score_sum = 0.0;
  }
}

// -------------------------------------------------------------------------------------------- score_slow
// Exact fp64 path, one warp per sub-segment that is before b* or not well conditioned.  Lane a holds row a of
// [S_i | u_i]; Gauss-Jordan with partial pivoting gives the determinant (gate) and the quadratic form.
template <int KP, bool CLIP>
__global__ void __launch_bounds__(32)
score_slow(const float* __restrict__ X, int64_t ld, int d, int k, SubLayout lay, const double* __restrict__ mean,
           const double* __restrict__ lo, const double* __restrict__ hi, const double* __restrict__ V,
           const double* __restrict__ PC0, const double* __restrict__ subproj, const int32_t* __restrict__ sub_flags,
           long long n_eff, long long row0, long long* __restrict__ gate, float* __restrict__ score,
           const float* __restrict__ in_mu, const float* __restrict__ in_is) {
  constexpr int KQ = KP + KP * KP;
  const int lane = threadIdx.x;
  const int a = lane < KP ? lane : KP - 1;
  const int bstar = (int)gate[0];
  for (int sub = blockIdx.x; sub < lay.nsub; sub += gridDim.x) {
    if (sub >= bstar && (sub_flags[sub] & SUBFLAG_WELL)) continue;
    bool open = sub >= bstar;
    const double* pz = subproj + (size_t)sub * KQ;
    double Za = pz[a];
    double Qrow[KP], S0row[KP];
#pragma unroll
    for (int b = 0; b < KP; ++b) {
      Qrow[b] = pz[KP + a * KP + b];
      S0row[b] = (double)(n_eff - 1) * PC0[a * KP + b];
    }
    const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
    for (int64_t r = r0; r < r1; ++r) {
      const long long i = row0 + r;
      const double ni = (double)(n_eff + i);
      // projections of row r: lane c owns component c
      double y = 0.0, z = 0.0;
      const float* x = X + r * ld;
      for (int f = 0; f < d; ++f) {
        const double xv = (double)(in_mu ? __fmul_rn(__fsub_rn(x[f], in_mu[f]), in_is[f]) : x[f]);  // input transform
        double xc = xv;
        if (CLIP) xc = fmin(fmax(xv, lo[f]), hi[f]);
        const double v = V[(size_t)f * KP + a];
        y = fma(xv - mean[f], v, y);
        z = fma(xc - mean[f], v, z);
      }
      const double pm = (n_eff + i > 0) ? Za / ni : 0.0;
      const double u = y - pm;
      double row[KP + 1];
#pragma unroll
      for (int b = 0; b < KP; ++b) {
        double v = S0row[b];
        if (n_eff + i > 0) v += Qrow[b] - Za * shfl_d(Za, b) / ni;
        if (a >= k || b >= k) v = (a == b) ? 1.0 : 0.0;
        row[b] = v;
      }
      row[KP] = (a < k) ? u : 0.0;
      const double uorig = row[KP];
      double div = 1.0;
      bool testable = true;
      if (i == 0) { div = (double)(n_eff - 1); testable = n_eff - 1 > 0; }
      else if (n_eff + i - 1 == 0) testable = false;
      else div = ni - 1.0;
      bool used = lane >= KP;  // mirror lanes never pivot
      int mycol = -1;
      double mypiv = 1.0, pmin = INFINITY, pmax = 0.0, det = 1.0;
      bool sing = false;
#pragma unroll
      for (int c = 0; c < KP; ++c) {
        double best = used ? -1.0 : fabs(row[c]);
        int bl = lane;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const double ov = __shfl_xor_sync(0xffffffffu, best, o);
          const int ol = __shfl_xor_sync(0xffffffffu, bl, o);
          if (ov > best || (ov == best && ol < bl)) { best = ov; bl = ol; }
        }
        if (!(best > 0.0) || !isfinite(best)) { sing = true; break; }  // warp-uniform
        const double piv = shfl_d(row[c], bl);
        if (c < k) {
          pmin = fmin(pmin, fabs(piv));
          pmax = fmax(pmax, fabs(piv));
          det *= piv / div;
        }
        const double f = (lane == bl) ? 0.0 : row[c] / piv;
#pragma unroll
        for (int j = c; j <= KP; ++j) {
          const double pj = shfl_d(row[j], bl);
          row[j] = fma(-f, pj, row[j]);
        }
        if (lane == bl) { used = true; mycol = c; mypiv = piv; }
      }
      const bool full_rank = !sing && pmin > kRankTol * pmax;
      double q = 0.0;
      if (full_rank) {
        const double uc = shfl_d(uorig, mycol < 0 ? 0 : mycol);
        double term = (lane < KP && mycol >= 0) ? uc * row[KP] / mypiv : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) term += __shfl_xor_sync(0xffffffffu, term, o);
        q = (ni - 1.0) * term;
      }
      if (!open && full_rank && testable && fabs(det) > kGateEps) {
        open = true;
        if (lane == 0) atomicMin(gate + 1, i);
      }
      if (lane == 0) score[r] = full_rank ? (float)q : 0.f;  // "score if open"; slow_fixup zeroes rows before i*
      // advance the prefix sums
#pragma unroll
      for (int b = 0; b < KP; ++b) Qrow[b] = fma(z, shfl_d(z, b), Qrow[b]);
      Za += z;
    }
  }
}

// rows of the slow sub-segments: zero what lies before the opening row, set flags, accumulate the batch sums
__global__ void __launch_bounds__(128)
slow_fixup(SubLayout lay, const int32_t* __restrict__ sub_flags, long long row0, double threshold,
           long long* __restrict__ gate, double* __restrict__ score_sum, float* __restrict__ score,
           int32_t* __restrict__ flag) {
  const int bstar = (int)gate[0];
  const long long istar = gate[1];
  const int lane = threadIdx.x & 31;
  const int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), nw = gridDim.x * (blockDim.x >> 5);
  long long nf = 0;
  double ss = 0.0;
  for (int sub = w; sub < lay.nsub; sub += nw) {
    if (sub >= bstar && (sub_flags[sub] & SUBFLAG_WELL)) continue;
    const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
    for (int64_t r = r0 + lane; r < r1; r += 32) {
      float s = score[r];
      if (row0 + r < istar) { s = 0.f; score[r] = 0.f; }
      const int f = (double)s > threshold;
      flag[r] = f;
      nf += f;
      ss += (double)s;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    nf += __shfl_xor_sync(0xffffffffu, nf, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if (lane == 0 && (nf != 0 || ss != 0.0)) {
    atomicAdd((unsigned long long*)(gate + 2), (unsigned long long)nf);
    atomicAdd(score_sum, ss);
  }
}

// -------------------------------------------------------------------------------------------- the fp32 chain
template <int KP>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = KP / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One row of the chain for the group this lane belongs to.  P = row j of S_i^-1 (updated to S_{i+1}^-1),
// buf = this group's 3*KP floats of shared memory.  All 32 lanes of the warp must call it together.
template <int KP>
__device__ __forceinline__ float chain_step(float (&P)[KP], float& Zrun, float z, float dl, bool need_u, bool active,
                                            float ni, float* buf, int j) {
  const float inv = ni > 0.f ? __frcp_rn(ni) : 0.f;
  const float e = fmaf(-Zrun, inv, z);
  const float u = e + dl;
  buf[j] = e;
  if (need_u) buf[KP + j] = u;
  __syncwarp();
  float c = 0.f;
#pragma unroll
  for (int i4 = 0; i4 < KP / 4; ++i4) {
    const float4 ev = *reinterpret_cast<const float4*>(buf + 4 * i4);
    c = fmaf(P[4 * i4 + 0], ev.x, c);
    c = fmaf(P[4 * i4 + 1], ev.y, c);
    c = fmaf(P[4 * i4 + 2], ev.z, c);
    c = fmaf(P[4 * i4 + 3], ev.w, c);
  }
  const float t = group_sum<KP>(e * c);
  float q = t;
  if (need_u) {  // warp-uniform
    float v = 0.f;
#pragma unroll
    for (int i4 = 0; i4 < KP / 4; ++i4) {
      const float4 uv = *reinterpret_cast<const float4*>(buf + KP + 4 * i4);
      v = fmaf(P[4 * i4 + 0], uv.x, v);
      v = fmaf(P[4 * i4 + 1], uv.y, v);
      v = fmaf(P[4 * i4 + 2], uv.z, v);
      v = fmaf(P[4 * i4 + 3], uv.w, v);
    }
    q = group_sum<KP>(u * v);
  }
  buf[2 * KP + j] = c;
  __syncwarp();
  const float w = ni / (ni + 1.f);
  const float gc = active ? w * c / fmaf(w, t, 1.f) : 0.f;
#pragma unroll
  for (int i4 = 0; i4 < KP / 4; ++i4) {
    const float4 cv = *reinterpret_cast<const float4*>(buf + 2 * KP + 4 * i4);
    P[4 * i4 + 0] = fmaf(-gc, cv.x, P[4 * i4 + 0]);
    P[4 * i4 + 1] = fmaf(-gc, cv.y, P[4 * i4 + 1]);
    P[4 * i4 + 2] = fmaf(-gc, cv.z, P[4 * i4 + 2]);
    P[4 * i4 + 3] = fmaf(-gc, cv.w, P[4 * i4 + 3]);
  }
  if (active) Zrun += z;
  return (ni - 1.f) * q;
}

// per-group cursor over the sub-segments assigned to it
struct GroupCursor {
  int sub;
  int64_t r, r_end;
};

template <int KP>
__device__ __forceinline__ void cursor_advance(GroupCursor& cur, int stride, const SubLayout& lay, int bstar,
                                               const int32_t* __restrict__ sub_flags, const float* __restrict__ sub_P,
                                               const float* __restrict__ sub_Z, float (&P)[KP], float& Zrun, int j) {
  while (cur.sub < lay.nsub && cur.r >= cur.r_end) {
    cur.sub += stride;
    if (cur.sub >= lay.nsub) break;
    if (cur.sub < bstar || !(sub_flags[cur.sub] & SUBFLAG_WELL)) continue;  // handled by score_slow
    cur.r = sub_begin(lay, cur.sub);
    cur.r_end = sub_begin(lay, cur.sub + 1);
    const float4* src = reinterpret_cast<const float4*>(sub_P + (size_t)cur.sub * KP * KP + (size_t)j * KP);
#pragma unroll
    for (int i4 = 0; i4 < KP / 4; ++i4) {
      const float4 v = src[i4];
      P[4 * i4 + 0] = v.x; P[4 * i4 + 1] = v.y; P[4 * i4 + 2] = v.z; P[4 * i4 + 3] = v.w;
    }
    Zrun = sub_Z[(size_t)cur.sub * KP + j];
  }
}

// generic projection (any d): lane j of a group accumulates component j of KP consecutive rows in registers
template <int KP, bool CLIP>
__global__ void __launch_bounds__(128)
score_chain(const float* __restrict__ X, int64_t ld, int d, SubLayout lay, const float* __restrict__ mean_f,
            const float* __restrict__ lo_f, const float* __restrict__ hi_f, const float* __restrict__ Vf,
            const float* __restrict__ sub_P, const float* __restrict__ sub_Z, const int32_t* __restrict__ sub_flags,
            long long n_eff, long long row0, double threshold, long long* __restrict__ gate,
            double* __restrict__ score_sum, float* __restrict__ score, int32_t* __restrict__ flag) {
  constexpr int GPW = 32 / KP;  // groups per warp
  extern __shared__ __align__(16) float smem[];
  float* Vs = smem;                      // [d][KP]
  float* ms = Vs + (size_t)d * KP;       // [d]
  float* los = ms + d;
  float* his = los + d;
  float* bufs = his + d;                 // [groups per CTA][3*KP]
  bufs = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(bufs) + 15) & ~(uintptr_t)15);
  for (int idx = threadIdx.x; idx < d * KP; idx += blockDim.x) Vs[idx] = Vf[idx];
  for (int f = threadIdx.x; f < d; f += blockDim.x) {
    ms[f] = mean_f[f];
    los[f] = CLIP ? lo_f[f] : 0.f;
    his[f] = CLIP ? hi_f[f] : 0.f;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, j = lane % KP;
  const int grp_in_cta = threadIdx.x / KP;
  float* buf = bufs + (size_t)grp_in_cta * 3 * KP;
  const int groups_total = gridDim.x * (blockDim.x / KP);
  const int gid = blockIdx.x * (blockDim.x / KP) + grp_in_cta;
  const int bstar = (int)gate[0];

  float P[KP];
#pragma unroll
  for (int i = 0; i < KP; ++i) P[i] = 0.f;
  float Zrun = 0.f;
  GroupCursor cur{gid - groups_total, 0, 0};
  cursor_advance<KP>(cur, groups_total, lay, bstar, sub_flags, sub_P, sub_Z, P, Zrun, j);
  long long nf = 0;
  double ss = 0.0;
  while (__any_sync(0xffffffffu, cur.sub < lay.nsub)) {
    const bool have = cur.sub < lay.nsub;
    const int64_t rb = cur.r;
    float zr[KP], dr[KP];
#pragma unroll
    for (int rr = 0; rr < KP; ++rr) { zr[rr] = 0.f; dr[rr] = 0.f; }
    if (have) {
      for (int f = 0; f < d; ++f) {
        const float v = Vs[f * KP + j], m = ms[f], l = los[f], h = his[f];
#pragma unroll
        for (int rr = 0; rr < KP; ++rr) {
          if (rb + rr < cur.r_end) {
            const float x = __ldg(X + (rb + rr) * ld + f);
            float xc = x;
            if (CLIP) xc = fminf(fmaxf(x, l), h);
            zr[rr] = fmaf(xc - m, v, zr[rr]);
            if (CLIP) dr[rr] = fmaf(x - xc, v, dr[rr]);
          }
        }
      }
    }
    float mysc = 0.f;
#pragma unroll
    for (int rr = 0; rr < KP; ++rr) {
      const bool active = have && (rb + rr < cur.r_end);
      const float dmag = CLIP ? group_sum<KP>(fabsf(dr[rr])) : 0.f;  // every lane takes part in the shuffle
      const bool clipped = active && dmag != 0.f;
      const bool need_u = CLIP && __any_sync(0xffffffffu, clipped);
      const float ni = (float)(double)(n_eff + row0 + rb + rr);
      const float sc = chain_step<KP>(P, Zrun, zr[rr], dr[rr], need_u, active, ni, buf, j);
      if (j == rr) mysc = sc;
    }
    if (have && rb + j < cur.r_end) {
      score[rb + j] = mysc;
      const int f = (double)mysc > threshold;
      flag[rb + j] = f;
      nf += f;
      ss += (double)mysc;
    }
    if (have) cur.r += KP;
    cursor_advance<KP>(cur, groups_total, lay, bstar, sub_flags, sub_P, sub_Z, P, Zrun, j);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    nf += __shfl_xor_sync(0xffffffffu, nf, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if (lane == 0 && (nf != 0 || ss != 0.0)) {
    atomicAdd((unsigned long long*)(gate + 2), (unsigned long long)nf);
    atomicAdd(score_sum, ss);
  }
  (void)GPW;
}

// ------------------------------------------------------------------------------- d = 64, KP = 16, tensor cores
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

constexpr int kZStride = 20;  // floats per projected row in shared memory (16 + pad, keeps float4 alignment)

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {  // packed fp32 FMA (FFMA2 on sm_100)
  unsigned long long ua = *reinterpret_cast<unsigned long long*>(&a), ub = *reinterpret_cast<unsigned long long*>(&b),
                     uc = *reinterpret_cast<unsigned long long*>(&c), ud;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(ud) : "l"(ua), "l"(ub), "l"(uc));
  return *reinterpret_cast<float2*>(&ud);
}

// One row of the chain for a 16-lane group; P = row j of S_i^-1 as 8 packed pairs.  Same algebra as chain_step<16>,
// with packed FMAs, four independent partial sums for c = S^-1 e, and reciprocals instead of divisions.
__device__ __forceinline__ float chain_step16(float2 (&P)[8], float& Zrun, float z, float dl, bool need_u, bool active,
                                              float inv, float w, float nim1, float* buf, int j) {
  const float e = fmaf(-Zrun, inv, z);
  const float u = e + dl;
  buf[j] = e;
  if (need_u) buf[16 + j] = u;
  __syncwarp();
  float2 a0 = make_float2(0.f, 0.f), a1 = a0;
#pragma unroll
  for (int i4 = 0; i4 < 4; ++i4) {
    const float4 ev = *reinterpret_cast<const float4*>(buf + 4 * i4);
    a0 = ffma2(P[2 * i4], make_float2(ev.x, ev.y), a0);
    a1 = ffma2(P[2 * i4 + 1], make_float2(ev.z, ev.w), a1);
  }
  const float c = (a0.x + a0.y) + (a1.x + a1.y);
  const float t = group_sum<16>(e * c);
  float q = t;
  if (need_u) {  // warp-uniform
    float2 b0 = make_float2(0.f, 0.f), b1 = b0;
#pragma unroll
    for (int i4 = 0; i4 < 4; ++i4) {
      const float4 uv = *reinterpret_cast<const float4*>(buf + 16 + 4 * i4);
      b0 = ffma2(P[2 * i4], make_float2(uv.x, uv.y), b0);
      b1 = ffma2(P[2 * i4 + 1], make_float2(uv.z, uv.w), b1);
    }
    q = group_sum<16>(u * ((b0.x + b0.y) + (b1.x + b1.y)));
  }
  buf[32 + j] = c;
  __syncwarp();
  const float gc = active ? __fdividef(w * c, fmaf(w, t, 1.f)) : 0.f;
  const float2 ngc = make_float2(-gc, -gc);
#pragma unroll
  for (int i4 = 0; i4 < 4; ++i4) {
    const float4 cv = *reinterpret_cast<const float4*>(buf + 32 + 4 * i4);
    P[2 * i4] = ffma2(ngc, make_float2(cv.x, cv.y), P[2 * i4]);
    P[2 * i4 + 1] = ffma2(ngc, make_float2(cv.z, cv.w), P[2 * i4 + 1]);
  }
  if (active) Zrun += z;
  return nim1 * q;
}

struct BlockPos {  // the 16-row block a group works on
  int sub;
  int64_t r, r_end;
};

// next block of the same sub-segment, or the first block of the next sub-segment assigned to this group
__device__ __forceinline__ BlockPos next_block(BlockPos p, int stride, const SubLayout& lay, int bstar,
                                               const int32_t* __restrict__ sub_flags, int step = 16) {
  p.r += step;
  while (p.sub < lay.nsub && p.r >= p.r_end) {
    p.sub += stride;
    if (p.sub >= lay.nsub) break;
    if (p.sub < bstar || !(sub_flags[p.sub] & SUBFLAG_WELL)) continue;  // handled by score_slow
    p.r = sub_begin(lay, p.sub);
    p.r_end = sub_begin(lay, p.sub + 1);
  }
  return p;
}

// A warp holds two groups of 16 lanes = two sub-segments.  Per step each group's next 16 rows are projected by the
// whole warp on mma.sync TF32 (M=16 rows, K=64 features, N=16 components, 3xTF32 split; lane (g,t) loads columns
// 16q+4t..+3 of rows g and g+8, so every warp load covers 64 contiguous bytes of 8 rows), the clipped projections z
// go to shared memory, the correction y - z = V^T(x - clip(x)) is added sparsely (only clipped elements contribute,
// exact fp32, shared-memory atomics), the next block's rows are prefetched into registers, and then both groups walk
// their 16 rows with chain_step16.
template <bool CLIP>
__global__ void __launch_bounds__(128, 3)
score_chain64(const float* __restrict__ X, int64_t ld, SubLayout lay, const float* __restrict__ mean_f,
              const float* __restrict__ lo_f, const float* __restrict__ hi_f, const float* __restrict__ Vf,
              const float* __restrict__ sub_P, const float* __restrict__ sub_Z, const int32_t* __restrict__ sub_flags,
              long long n_eff, long long row0, double threshold, long long* __restrict__ gate,
              double* __restrict__ score_sum, float* __restrict__ score, int32_t* __restrict__ flag) {
  constexpr int KP = 16, D = 64;
  __shared__ __align__(16) float ms[D], los[D], his[D];
  __shared__ __align__(16) float Vs[CLIP ? D * KP : 4];       // fp32 eigenvectors for the sparse clip correction
  __shared__ __align__(16) uint4 vfrag[16 * 32];               // B fragments {hi(k=t), hi(k=t+4), lo(k=t), lo(k=t+4)}
  __shared__ __align__(16) float zbuf[4][2][16 * kZStride];    // [warp][group][row][comp]  z = V^T(clip(x) - a)
  __shared__ __align__(16) float dbuf[4][2][16 * kZStride];    // y - z, non-zero only on clipped rows
  __shared__ __align__(16) float4 rowsc[4][2][16];             // per row: 1/(n+i), (n+i)/(n+i+1), n+i-1, clipped?
  __shared__ __align__(16) float cbufs[8][3 * KP];
  constexpr int kEvCap = 48;                                   // clipped elements listed per block and group
  __shared__ __align__(16) float2 evlist[4][2][CLIP ? kEvCap : 1];  // {row*64 + col as int bits, x - clip(x)}
  __shared__ int evcnt[4][2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int j = lane & 15, grp = lane >> 4;
  for (int f = threadIdx.x; f < D; f += blockDim.x) {
    ms[f] = mean_f[f];
    los[f] = CLIP ? lo_f[f] : 0.f;
    his[f] = CLIP ? hi_f[f] : 0.f;
  }
  if (CLIP)
    for (int idx = threadIdx.x; idx < D * KP; idx += blockDim.x) Vs[idx] = Vf[idx];
  for (int idx = threadIdx.x; idx < 16 * 32; idx += blockDim.x) {
    const int fr = idx >> 5, ln = idx & 31, ks = fr >> 1, nt = fr & 1, gg = ln >> 2, tt = ln & 3;
    const int f0 = 16 * (ks >> 1) + 4 * tt + 2 * (ks & 1);  // step ks uses physical rows f0 (k=t) and f0+1 (k=t+4)
    const float v0 = Vf[f0 * KP + 8 * nt + gg], v1 = Vf[(f0 + 1) * KP + 8 * nt + gg];
    uint4 fr4;
    fr4.x = to_tf32(v0);
    fr4.y = to_tf32(v1);
    fr4.z = to_tf32(v0 - __uint_as_float(fr4.x));
    fr4.w = to_tf32(v1 - __uint_as_float(fr4.y));
    vfrag[idx] = fr4;
  }
  for (int idx = threadIdx.x; idx < 4 * 2 * 16 * kZStride; idx += blockDim.x) (&dbuf[0][0][0])[idx] = 0.f;
  for (int idx = threadIdx.x; idx < 4 * 2 * 16; idx += blockDim.x) (&rowsc[0][0][0])[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (threadIdx.x < 8) (&evcnt[0][0])[threadIdx.x] = 0;
  __syncthreads();
  float* buf = cbufs[warp * 2 + grp];
  const int groups_total = gridDim.x * 8;
  const int gid = blockIdx.x * 8 + warp * 2 + grp;
  const int bstar = (int)gate[0];

  float2 P[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) P[i] = make_float2(0.f, 0.f);
  float Zrun = 0.f;
  BlockPos cur{gid - groups_total, 0, 0};
  cur.r = -16;  // next_block() adds 16 first
  cur = next_block(cur, groups_total, lay, bstar, sub_flags);
  int loaded_sub = -1;

  float4 xa[2][4], xb[2][4];  // rows g and g+8 of each group's current block
  auto load_rows = [&](const BlockPos& p) {
#pragma unroll
    for (int gs = 0; gs < 2; ++gs) {
      const int src = gs * 16;
      const int64_t rb = __shfl_sync(0xffffffffu, p.r, src);
      const int64_t re = __shfl_sync(0xffffffffu, p.r_end, src);
      const int psub = __shfl_sync(0xffffffffu, p.sub, src);
      const bool ghave = psub < lay.nsub;
      const bool va = ghave && rb + g < re, vb8 = ghave && rb + g + 8 < re;
      const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4* pa = reinterpret_cast<const float4*>(X + (rb + g) * ld);
      const float4* pb = reinterpret_cast<const float4*>(X + (rb + g + 8) * ld);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        xa[gs][q] = va ? ldg_stream_f4(pa + 4 * q + t) : zero4;
        xb[gs][q] = vb8 ? ldg_stream_f4(pb + 4 * q + t) : zero4;
      }
    }
  };
  load_rows(cur);

  long long nf = 0;
  double ss = 0.0;
  while (__any_sync(0xffffffffu, cur.sub < lay.nsub)) {
    const bool have = cur.sub < lay.nsub;
    // ---- projection of both groups' blocks by the whole warp
#pragma unroll
    for (int gs = 0; gs < 2; ++gs) {
      const int src = gs * 16;
      const int64_t rb = __shfl_sync(0xffffffffu, cur.r, src);
      const int64_t re = __shfl_sync(0xffffffffu, cur.r_end, src);
      const bool ghave = __shfl_sync(0xffffffffu, (int)have, src) != 0;
      if (!ghave) continue;  // warp-uniform
      const bool va = rb + g < re, vb8 = rb + g + 8 < re;
      float accz[2][4];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) accz[nt][e] = 0.f;
      float* zrow = zbuf[warp][gs];
      float* drow = dbuf[warp][gs];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 m4 = *reinterpret_cast<const float4*>(ms + 16 * q + 4 * t);
        const float mm[4] = {m4.x, m4.y, m4.z, m4.w};
        float ra[4] = {xa[gs][q].x, xa[gs][q].y, xa[gs][q].z, xa[gs][q].w};
        float rbv[4] = {xb[gs][q].x, xb[gs][q].y, xb[gs][q].z, xb[gs][q].w};
        if (CLIP) {
          const float4 l4 = *reinterpret_cast<const float4*>(los + 16 * q + 4 * t);
          const float4 h4 = *reinterpret_cast<const float4*>(his + 16 * q + 4 * t);
          const float ll[4] = {l4.x, l4.y, l4.z, l4.w}, hh[4] = {h4.x, h4.y, h4.z, h4.w};
          float ca[4], cb[4];
          bool anyc = false;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            ca[e] = fminf(fmaxf(ra[e], ll[e]), hh[e]);
            cb[e] = fminf(fmaxf(rbv[e], ll[e]), hh[e]);
            anyc = anyc || (ca[e] != ra[e]) || (cb[e] != rbv[e]);
          }
          if (anyc) {  // rare (0.3 % of the elements are clipped): list them, the group adds V^T(x - clip(x)) later
#pragma unroll
            for (int e = 0; e < 4; ++e) {
#pragma unroll
              for (int rr = 0; rr < 2; ++rr) {
                const float x = rr == 0 ? ra[e] : rbv[e];
                const float xc = rr == 0 ? ca[e] : cb[e];
                if (xc != x && (rr == 0 ? va : vb8)) {
                  const int row = rr == 0 ? g : g + 8;
                  const int col = 16 * q + 4 * t + e;
                  const int slot = atomicAdd(&evcnt[warp][gs], 1);
                  if (slot < kEvCap) {
                    evlist[warp][gs][slot] = make_float2(__int_as_float(row * 64 + col), x - xc);
                  } else {  // list full (a block dense in outliers): add this element's contribution directly
                    const float diff = x - xc;
                    const float* vrow = Vs + col * KP;
                    for (int c = 0; c < KP; ++c) atomicAdd(drow + row * kZStride + c, diff * vrow[c]);
                    reinterpret_cast<float*>(&rowsc[warp][gs][row])[3] = 1.f;
                  }
                }
              }
            }
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) { ra[e] = ca[e]; rbv[e] = cb[e]; }
        }
        uint32_t zb[2][4], zs[2][4];  // [row g / g+8][element]: hi and lo TF32 parts of clip(x) - a
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float da = ra[e] - mm[e], db = rbv[e] - mm[e];
          zb[0][e] = to_tf32(da);
          zs[0][e] = to_tf32(da - __uint_as_float(zb[0][e]));
          zb[1][e] = to_tf32(db);
          zs[1][e] = to_tf32(db - __uint_as_float(zb[1][e]));
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int ks = 2 * q + half, e0 = 2 * half;
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            const uint4 vf = vfrag[(ks * 2 + nt) * 32 + lane];
            mma_m16n8k8_tf32(accz[nt], zs[0][e0], zs[1][e0], zs[0][e0 + 1], zs[1][e0 + 1], vf.x, vf.y);
            mma_m16n8k8_tf32(accz[nt], zb[0][e0], zb[1][e0], zb[0][e0 + 1], zb[1][e0 + 1], vf.z, vf.w);
            mma_m16n8k8_tf32(accz[nt], zb[0][e0], zb[1][e0], zb[0][e0 + 1], zb[1][e0 + 1], vf.x, vf.y);
          }
        }
      }
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        const int c0 = 8 * nt + 2 * t;
        *reinterpret_cast<float2*>(zrow + g * kZStride + c0) = make_float2(accz[nt][0], accz[nt][1]);
        *reinterpret_cast<float2*>(zrow + (g + 8) * kZStride + c0) = make_float2(accz[nt][2], accz[nt][3]);
      }
    }
    // ---- per-row scalars of this group's block (lane j <-> row j), chain state on entering a sub-segment
    const int64_t rb = cur.r;
    {
      const float ni = (float)(double)(n_eff + row0 + rb + j);
      float* rs = reinterpret_cast<float*>(&rowsc[warp][grp][j]);
      rs[0] = ni > 0.f ? __fdividef(1.f, ni) : 0.f;
      rs[1] = __fdividef(ni, ni + 1.f);
      rs[2] = ni - 1.f;
    }
    if (have && cur.sub != loaded_sub) {
      const float4* src4 = reinterpret_cast<const float4*>(sub_P + (size_t)cur.sub * KP * KP + (size_t)j * KP);
#pragma unroll
      for (int i4 = 0; i4 < 4; ++i4) {
        const float4 v = src4[i4];
        P[2 * i4] = make_float2(v.x, v.y);
        P[2 * i4 + 1] = make_float2(v.z, v.w);
      }
      Zrun = sub_Z[(size_t)cur.sub * KP + j];
      loaded_sub = cur.sub;
    }
    // ---- prefetch the next block's rows while the chain runs
    const BlockPos nxt = have ? next_block(cur, groups_total, lay, bstar, sub_flags) : cur;
    load_rows(nxt);
    __syncwarp();
    const float* zrow = zbuf[warp][grp];
    float* drow = dbuf[warp][grp];
    if (CLIP) {
      // y - z = V^T(x - clip(x)): lane j adds component j of every listed element (exact fp32, one owner per word)
      const int nev = min(evcnt[warp][grp], kEvCap);  // group-uniform
      for (int k = 0; k < nev; ++k) {
        const float2 ev = evlist[warp][grp][k];
        const int rc = __float_as_int(ev.x), row = rc >> 6, col = rc & 63;
        drow[row * kZStride + j] = fmaf(ev.y, Vs[col * KP + j], drow[row * kZStride + j]);
        if (j == 0) reinterpret_cast<float*>(&rowsc[warp][grp][row])[3] = 1.f;
      }
      __syncwarp();
      if (j == 0) evcnt[warp][grp] = 0;
    }
    // ---- both groups walk their 16 rows
    float mysc = 0.f;
#pragma unroll 4
    for (int rr = 0; rr < 16; ++rr) {
      const bool active = have && (rb + rr < cur.r_end);
      const float4 sc4 = rowsc[warp][grp][rr];
      const float z = active ? zrow[rr * kZStride + j] : 0.f;
      bool clipped = false;
      float dl = 0.f;
      if (CLIP) {
        clipped = sc4.w != 0.f;  // group-uniform
        if (clipped) {
          dl = drow[rr * kZStride + j];
          drow[rr * kZStride + j] = 0.f;  // leave the buffer clean for the next block
        }
      }
      const bool need_u = CLIP && __any_sync(0xffffffffu, clipped);
      const float sc = chain_step16(P, Zrun, z, dl, need_u, active, sc4.x, sc4.y, sc4.z, buf, j);
      if (j == rr) mysc = sc;
    }
    if (CLIP) reinterpret_cast<float*>(&rowsc[warp][grp][j])[3] = 0.f;
    if (have && rb + j < cur.r_end) {
      score[rb + j] = mysc;
      const int f = (double)mysc > threshold;
      flag[rb + j] = f;
      nf += f;
      ss += (double)mysc;
    }
    __syncwarp();
    cur = nxt;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    nf += __shfl_xor_sync(0xffffffffu, nf, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if (lane == 0 && (nf != 0 || ss != 0.0)) {
    atomicAdd((unsigned long long*)(gate + 2), (unsigned long long)nf);
    atomicAdd(score_sum, ss);
  }
}

// ------------------------------------------------------------------ split pass 2 (default for d = 64, k <= 16)
// The fused kernel above carries the projection's registers (row prefetch, fragments) through the latency-bound
// chain and ends up with ~9 warps per SM.  Splitting lets each half run at its own occupancy:
//   project64 : row-independent, streams X once (4*d B/row), writes z (64 B/row), a clip flag (1 B/row) and y - z
//               for the rows that have a clipped feature (64 B for ~17 % of the rows);
//   chain16   : lean registers, 16+ warps per SM, reads z back (L2/HBM) and walks the inverse chains.
// Cost: ~150 B/row of extra DRAM traffic on top of the algorithmic 4*d + 8.
// clipped elements listed per 16-row block; denser blocks take the slow path.  The wide shapes keep the list short so
// that three CTAs fit in shared memory (78 KB -> 66 KB per CTA at d = 256; the expected count there is 11 per block)
__host__ __device__ constexpr int proj_ev_cap(int NB) { return NB == 1 ? 512 : 128; }

// shared-memory carve-up of project_rows<NB>: sizes in bytes
__host__ __device__ constexpr size_t proj_smem_bytes(int NB, bool clip) {
  return (size_t)5 * 64 * NB * 4                      // mean, lo, hi, input transform (mu, 1/sigma)
         + (clip ? (size_t)64 * NB * 16 * 4 : 16)     // fp32 eigenvectors (sparse clip correction)
         + (size_t)NB * 16 * 32 * 16                  // B fragments
         + (clip ? (size_t)4 * proj_ev_cap(NB) * 8 : 16)   // event lists, one per warp
         + (clip ? (size_t)4 * 16 * 16 * 4 : 16)      // y - z tiles, one per warp
         + 64                                         // event counters
         + (size_t)4 * 16 * 17 * 4;                   // z tiles for the Z^T Z fragments
}

// d = 64*NB, k <= 16.  Each warp projects blocks of 16 rows: 3xTF32 mma.sync over NB column chunks of 64 features.
// A warp owns whole sub-segments, so it can also sum z and z z^T over each of them (Z^T Z on the tensor cores, same
// 3xTF32 split): these are the projected sub-segment statistics that the prefix scan needs, i.e. V^T s and V^T G V of
// the sub-segment without ever projecting a d x d partial.
// STD: rows are standardised on load, (x - in_mu) * in_is with the roundings of the stand-alone kernel (see pass 1).
template <int NB, bool CLIP, bool STD>
__global__ void __launch_bounds__(128, NB == 1 ? 4 : 3)
project_rows(const float* __restrict__ X, int64_t ld, SubLayout lay, const float* __restrict__ mean_f,
             const float* __restrict__ lo_f, const float* __restrict__ hi_f, const float* __restrict__ in_mu,
             const float* __restrict__ in_is, const float* __restrict__ Vf,
             float* __restrict__ Zout, float* __restrict__ Dout, uint8_t* __restrict__ cflag,
             double* __restrict__ subproj) {
  constexpr int KP = 16, D = 64 * NB;
  extern __shared__ __align__(16) unsigned char proj_smem[];
  float* ms = reinterpret_cast<float*>(proj_smem);
  float* los = ms + D;
  float* his = los + D;
  float* smu = his + D;
  float* sis = smu + D;
  float* Vs = sis + D;
  uint4* vfrag = reinterpret_cast<uint4*>(Vs + (CLIP ? D * KP : 4));
  float2* evbase = reinterpret_cast<float2*>(vfrag + NB * 16 * 32);
  float* dtbase = reinterpret_cast<float*>(evbase + (CLIP ? 4 * proj_ev_cap(NB) : 2));
  int* evcnt = reinterpret_cast<int*>(dtbase + (CLIP ? 4 * 16 * KP : 4));
  float* ztbase = reinterpret_cast<float*>(evcnt + 16);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  float2* evlist = evbase + (CLIP ? warp * proj_ev_cap(NB) : 0);
  float* dt = dtbase + (CLIP ? warp * 16 * KP : 0);
  float* zt = ztbase + warp * 16 * 17;  // this block's z, [row][comp] padded to 17
  for (int f = threadIdx.x; f < D; f += blockDim.x) {
    ms[f] = mean_f[f];
    los[f] = CLIP ? lo_f[f] : 0.f;
    his[f] = CLIP ? hi_f[f] : 0.f;
    smu[f] = STD ? in_mu[f] : 0.f;
    sis[f] = STD ? in_is[f] : 0.f;
  }
  if (CLIP)
    for (int idx = threadIdx.x; idx < D * KP; idx += blockDim.x) Vs[idx] = Vf[idx];
  for (int idx = threadIdx.x; idx < NB * 16 * 32; idx += blockDim.x) {
    const int fr = idx >> 5, ln = idx & 31, cc = fr >> 4, ks = (fr >> 1) & 7, nt = fr & 1, gg = ln >> 2, tt = ln & 3;
    const int f0 = 64 * cc + 16 * (ks >> 1) + 4 * tt + 2 * (ks & 1);
    const float v0 = Vf[f0 * KP + 8 * nt + gg], v1 = Vf[(f0 + 1) * KP + 8 * nt + gg];
    uint4 fr4;
    fr4.x = to_tf32(v0);
    fr4.y = to_tf32(v1);
    fr4.z = to_tf32(v0 - __uint_as_float(fr4.x));
    fr4.w = to_tf32(v1 - __uint_as_float(fr4.y));
    vfrag[idx] = fr4;
  }
  if (threadIdx.x < 4) evcnt[threadIdx.x] = 0;
  __syncthreads();

  const int64_t rows = lay.rows;
  const int wtotal = gridDim.x * 4;
  float4 xa[4], xb[4];
  // rows [rb, rb+16) clipped to [.., rend): the loader of one 64-column chunk
  auto load_rows = [&](int64_t rb, int64_t rend, int cc) {
    const bool va = rb + g < rend, vb8 = rb + g + 8 < rend;
    const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* pa = reinterpret_cast<const float4*>(X + (rb + g) * ld + 64 * cc);
    const float4* pb = reinterpret_cast<const float4*>(X + (rb + g + 8) * ld + 64 * cc);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      xa[q] = va ? ldg_stream_f4(pa + 4 * q + t) : zero4;
      xb[q] = vb8 ? ldg_stream_f4(pb + 4 * q + t) : zero4;
    }
  };
  for (int sub = blockIdx.x * 4 + warp; sub < lay.nsub; sub += wtotal) {
  const int64_t r0 = sub_begin(lay, sub), r1 = sub_begin(lay, sub + 1);
  float qacc[2][4], zsum[2][2];
#pragma unroll
  for (int nt = 0; nt < 2; ++nt) {
    zsum[nt][0] = zsum[nt][1] = 0.f;
#pragma unroll
    for (int e = 0; e < 4; ++e) qacc[nt][e] = 0.f;
  }
  load_rows(r0, r1, 0);
  for (int64_t rb = r0; rb < r1; rb += 16) {
    const bool va = rb + g < r1, vb8 = rb + g + 8 < r1;
    float accz[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) accz[nt][e] = 0.f;
#pragma unroll 1
    for (int cc = 0; cc < NB; ++cc) {
      float ra[4][4], rbv[4][4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        ra[q][0] = xa[q].x; ra[q][1] = xa[q].y; ra[q][2] = xa[q].z; ra[q][3] = xa[q].w;
        rbv[q][0] = xb[q].x; rbv[q][1] = xb[q].y; rbv[q][2] = xb[q].z; rbv[q][3] = xb[q].w;
      }
      // next chunk (or the first chunk of the next block of this sub-segment) in flight while this one is multiplied
      if (cc + 1 < NB) load_rows(rb, r1, cc + 1); else load_rows(rb + 16, r1, 0);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int cbase = 64 * cc + 16 * q + 4 * t;
        const float4 m4 = *reinterpret_cast<const float4*>(ms + cbase);
        const float mm[4] = {m4.x, m4.y, m4.z, m4.w};
        if (STD) {
          const float4 u4 = *reinterpret_cast<const float4*>(smu + cbase);
          const float4 s4 = *reinterpret_cast<const float4*>(sis + cbase);
          const float uu[4] = {u4.x, u4.y, u4.z, u4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            ra[q][e] = __fmul_rn(__fsub_rn(ra[q][e], uu[e]), ss[e]);
            rbv[q][e] = __fmul_rn(__fsub_rn(rbv[q][e], uu[e]), ss[e]);
          }
        }
        if (CLIP) {
          const float4 l4 = *reinterpret_cast<const float4*>(los + cbase);
          const float4 h4 = *reinterpret_cast<const float4*>(his + cbase);
          const float ll[4] = {l4.x, l4.y, l4.z, l4.w}, hh[4] = {h4.x, h4.y, h4.z, h4.w};
          float ca[4], cb[4];
          uint32_t diff = 0;  // any bit that clipping changed (one LOP3 per element)
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            ca[e] = fminf(fmaxf(ra[q][e], ll[e]), hh[e]);
            cb[e] = fminf(fmaxf(rbv[q][e], ll[e]), hh[e]);
            diff |= __float_as_uint(ca[e]) ^ __float_as_uint(ra[q][e]);
            diff |= __float_as_uint(cb[e]) ^ __float_as_uint(rbv[q][e]);
          }
          if (diff != 0) {  // rare: list the clipped elements of this lane
#pragma unroll
            for (int e = 0; e < 4; ++e) {
#pragma unroll
              for (int rr = 0; rr < 2; ++rr) {
                const float x = rr == 0 ? ra[q][e] : rbv[q][e];
                const float xc = rr == 0 ? ca[e] : cb[e];
                if (xc != x && (rr == 0 ? va : vb8)) {
                  const int slot = atomicAdd(&evcnt[warp], 1);
                  if (slot < proj_ev_cap(NB))
                    evlist[slot] = make_float2(__int_as_float((rr == 0 ? g : g + 8) * D + cbase + e), x - xc);
                }
              }
            }
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) { ra[q][e] = ca[e]; rbv[q][e] = cb[e]; }
        }
        uint32_t zb[2][4], zs[2][4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float da = ra[q][e] - mm[e], db = rbv[q][e] - mm[e];
          // TF32 split by truncation: the tensor core reads the top 19 bits of an operand register and ignores the
          // rest, so the "hi" operand is the value itself (no instruction) and "lo" is what the truncation dropped
          // (exact in fp32; itself truncated by the hardware: 2^-20 relative per product instead of 2^-22 with
          // rounded parts, far inside the 1e-4 score tolerance)
          zb[0][e] = __float_as_uint(da);
          zs[0][e] = __float_as_uint(da - __uint_as_float(zb[0][e] & 0xffffe000u));
          zb[1][e] = __float_as_uint(db);
          zs[1][e] = __float_as_uint(db - __uint_as_float(zb[1][e] & 0xffffe000u));
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int ks = 2 * q + half, e0 = 2 * half;
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            const uint4 vf = vfrag[((cc * 8 + ks) * 2 + nt) * 32 + lane];
            mma_m16n8k8_tf32(accz[nt], zs[0][e0], zs[1][e0], zs[0][e0 + 1], zs[1][e0 + 1], vf.x, vf.y);
            mma_m16n8k8_tf32(accz[nt], zb[0][e0], zb[1][e0], zb[0][e0 + 1], zb[1][e0 + 1], vf.z, vf.w);
            mma_m16n8k8_tf32(accz[nt], zb[0][e0], zb[1][e0], zb[0][e0 + 1], zb[1][e0 + 1], vf.x, vf.y);
          }
        }
      }
    }
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
      const int c0 = 8 * nt + 2 * t;
      // rows past the end of the sub-segment contribute nothing
      const float z00 = va ? accz[nt][0] : 0.f, z01 = va ? accz[nt][1] : 0.f;
      const float z10 = vb8 ? accz[nt][2] : 0.f, z11 = vb8 ? accz[nt][3] : 0.f;
      if (va) *reinterpret_cast<float2*>(Zout + (rb + g) * KP + c0) = make_float2(z00, z01);
      if (vb8) *reinterpret_cast<float2*>(Zout + (rb + g + 8) * KP + c0) = make_float2(z10, z11);
      zsum[nt][0] += z00 + z10;
      zsum[nt][1] += z01 + z11;
      zt[g * 17 + c0] = z00; zt[g * 17 + c0 + 1] = z01;
      zt[(g + 8) * 17 + c0] = z10; zt[(g + 8) * 17 + c0 + 1] = z11;
    }
    __syncwarp();
    // Q += Z^T Z over the 16 rows: A[m = comp][k = row] and B[k = row][n = comp] are the same four values per step
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      const float f0 = zt[(8 * ks + t) * 17 + g], f1 = zt[(8 * ks + t) * 17 + g + 8];
      const float f2 = zt[(8 * ks + t + 4) * 17 + g], f3 = zt[(8 * ks + t + 4) * 17 + g + 8];
      const uint32_t b0 = __float_as_uint(f0), b1 = __float_as_uint(f1), b2 = __float_as_uint(f2), b3 = __float_as_uint(f3);
      const uint32_t s0 = __float_as_uint(f0 - __uint_as_float(b0 & 0xffffe000u));
      const uint32_t s1 = __float_as_uint(f1 - __uint_as_float(b1 & 0xffffe000u));
      const uint32_t s2 = __float_as_uint(f2 - __uint_as_float(b2 & 0xffffe000u));
      const uint32_t s3 = __float_as_uint(f3 - __uint_as_float(b3 & 0xffffe000u));
      // n-tile 0 uses (k=t, n=g) = f0 and (k=t+4, n=g) = f2; n-tile 1 uses f1, f3
      mma_m16n8k8_tf32(qacc[0], s0, s1, s2, s3, b0, b2);
      mma_m16n8k8_tf32(qacc[0], b0, b1, b2, b3, s0, s2);
      mma_m16n8k8_tf32(qacc[0], b0, b1, b2, b3, b0, b2);
      mma_m16n8k8_tf32(qacc[1], s0, s1, s2, s3, b1, b3);
      mma_m16n8k8_tf32(qacc[1], b0, b1, b2, b3, s1, s3);
      mma_m16n8k8_tf32(qacc[1], b0, b1, b2, b3, b1, b3);
    }
    unsigned rowmask = 0;
    if (CLIP) {
      __syncwarp();
      const int nev = evcnt[warp];  // warp-uniform
      if (nev > 0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) reinterpret_cast<float4*>(dt)[lane + 32 * i] = make_float4(0.f, 0.f, 0.f, 0.f);
        __syncwarp();
        // y - z = V^T(x - clip(x)); lane (half, j) owns component j of rows 8 half .. 8 half + 7: no atomics
        const int half = lane >> 4, jj = lane & 15;
        if (nev <= proj_ev_cap(NB)) {
          for (int k = 0; k < nev; ++k) {
            const float2 ev = evlist[k];
            const int rc = __float_as_int(ev.x), row = rc / D, col = rc - row * D;
            rowmask |= 1u << row;
            if ((row >> 3) == half) {
              float* acc = dt + row * KP + jj;
              *acc = fmaf(ev.y, Vs[col * KP + jj], *acc);
            }
          }
        } else {
          // more than proj_ev_cap(NB) clipped elements in 16 rows (degenerate data): recompute the block densely
          rowmask = 0xffffu;
          for (int row = half; row < 16; row += 2) {
            if (rb + row >= r1) break;
            const float* xr = X + (rb + row) * ld;
            float acc = 0.f;
            for (int f = 0; f < D; ++f) {
              float x = __ldg(xr + f);
              if (STD) x = __fmul_rn(__fsub_rn(x, smu[f]), sis[f]);
              acc = fmaf(x - fminf(fmaxf(x, los[f]), his[f]), Vs[f * KP + jj], acc);
            }
            dt[row * KP + jj] = acc;
          }
        }
        __syncwarp();
        for (unsigned m = rowmask; m != 0; m &= m - 1) {
          const int row = __ffs(m) - 1;
          if (lane < KP) Dout[(rb + row) * KP + lane] = dt[row * KP + lane];
        }
        __syncwarp();
        if (lane == 0) evcnt[warp] = 0;
      }
    }
    if (lane < 16 && rb + lane < r1) cflag[rb + lane] = (uint8_t)((rowmask >> lane) & 1u);
    __syncwarp();
  }
  // projected statistics of this sub-segment: sum z (16) and sum z z^T (16 x 16)
  {
    double* o = subproj + (size_t)sub * (KP + KP * KP);
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
      for (int h2 = 0; h2 < 2; ++h2) {
        float v = zsum[nt][h2];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        if (g == 0) o[8 * nt + 2 * t + h2] = (double)v;
      }
      const int n0 = 8 * nt + 2 * t;
      o[KP + g * KP + n0] = (double)qacc[nt][0];
      o[KP + g * KP + n0 + 1] = (double)qacc[nt][1];
      o[KP + (g + 8) * KP + n0] = (double)qacc[nt][2];
      o[KP + (g + 8) * KP + n0 + 1] = (double)qacc[nt][3];
    }
  }
  }
  (void)rows;
}

// chain over the projected rows.  Two rows per step (rank-2 form of the Sherman-Morrison update): with c1 = P e1,
// d2 = P e2 and the three dot products t11 = e1.c1, t12 = e1.d2, t22 = e2.d2 (one 3-wide shuffle butterfly),
//     g1 = w1/(1 + w1 t11),  c2 = d2 - g1 t12 c1,  t2 = t22 - g1 t12^2,  g2 = w2/(1 + w2 t2),
//     P <- P - g1 c1 c1^T - g2 c2 c2^T,   score1 = (n+i-1) t11,  score2 = (n+i) t2.
// All deviations e of a block are formed up front from the running prefix and parked in shared memory, so a step needs
// a single warp synchronisation (for the c broadcast).
// chain4: FOUR lanes per sub-segment (lane l owns rows 4l..4l+3 of S^-1 = 64 registers), eight sub-segments per warp,
// blocks of 8 rows.  Compared with one lane per row of S^-1 (16 lanes per chain, the first version of this kernel) the
// matrix-vector products and the rank-2 update are the same FFMA2 count, but the shuffle butterflies are two rounds
// instead of four and are shared by four times as many chains, and the operand loads (e, c) are shared by eight chains:
// about half the warp instructions per row.  Clipped rows use the expansion
//     u^T P u = e^T P e + 2 dl.(P e) + dl.(P dl),   dl = y - z = V^T(x - clip(x)),
// so only P dl is computed in addition (and only for the row of the pair that needs it).
constexpr int kChain4Rows = 8;
// 8 warps per SM = 2 per scheduler with up to 255 registers; 3 per scheduler (168 registers) spills the prefetched block
// and was measured slower (1.12 ms against 0.85 ms for 16Mi rows)
constexpr int kChain4Warps = 4;                    // warps per CTA
constexpr int kChain4Ctas = 2;                     // CTAs per SM
constexpr int kChain4Chains = 8 * kChain4Warps;    // chains per CTA
template <bool CLIP>
__global__ void __launch_bounds__(32 * kChain4Warps, kChain4Ctas)
chain4(const float* __restrict__ Zp, const float* __restrict__ Dp, const uint8_t* __restrict__ cflag, SubLayout lay,
       const float* __restrict__ sub_P, const float* __restrict__ sub_Z, const int32_t* __restrict__ sub_flags,
       long long n_eff, long long row0, double threshold, long long* __restrict__ gate,
       double* __restrict__ score_sum, float* __restrict__ score, int32_t* __restrict__ flag) {
  constexpr int KP = 16, RB = kChain4Rows;
  // per-chain strides are padded by 16 bytes: the eight chains of a warp read the same offsets of their own buffers
  // in one LDS.128, and unpadded strides (512 B, 256 B) would put all eight in the same banks
  constexpr int ES = RB * KP + 4, CS = 2 * 2 * KP + 4;
  extern __shared__ __align__(16) float chain4_smem[];
  constexpr int NCH = kChain4Chains;
  float* ebuf = chain4_smem;                    // [NCH chains][ES]      e of the block being walked
  float* cbufs = ebuf + NCH * ES;               // [NCH][CS]             c1 | c2, two parities
  float* dbuf = cbufs + NCH * CS;               // [2][NCH][ES] (CLIP)   y - z, zero for rows without a clipped feature
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int l = lane & 3, ci = warp * 8 + (lane >> 2);
  float* eb = ebuf + ci * ES;
  const int chains_total = gridDim.x * NCH;
  const int bstar = (int)gate[0];

  float2 P[4][8];
#pragma unroll
  for (int q = 0; q < 4; ++q)
#pragma unroll
    for (int i = 0; i < 8; ++i) P[q][i] = make_float2(0.f, 0.f);
  float4 Zrun = make_float4(0.f, 0.f, 0.f, 0.f);
  BlockPos cur{(int)(blockIdx.x * NCH + ci) - chains_total, 0, 0};
  cur.r = -RB;
  cur = next_block(cur, chains_total, lay, bstar, sub_flags, RB);
  int loaded_sub = -1;
  long long nf = 0;
  double ss = 0.0;
  // Software pipeline over the blocks of a chain: the clip flags are fetched two blocks ahead, the rows (z) one block
  // ahead into registers, the clip corrections (y - z, only rows whose flag is set; the others are zero-filled) one
  // block ahead with cp.async into the other half of dbuf.
  auto rows_of = [&](const BlockPos& p) { return p.sub < lay.nsub ? (int)min((int64_t)RB, p.r_end - p.r) : 0; };
  auto load_flags = [&](const BlockPos& p) {  // eight flag bytes of the block, raw
    uint2 fb = make_uint2(0u, 0u);
    const int na = rows_of(p);
    if (CLIP && na > 0) {
      if (na == RB) {
        fb = __ldg(reinterpret_cast<const uint2*>(cflag + p.r));  // block starts are multiples of 8
      } else {
        for (int rr = 0; rr < na; ++rr) {
          const unsigned b = cflag[p.r + rr] != 0 ? 1u : 0u;
          if (rr < 4) fb.x |= b << (8 * rr); else fb.y |= b << (8 * (rr - 4));
        }
      }
    }
    return fb;
  };
  auto flag_mask = [](uint2 fb) {  // bit r = row r has a clipped feature
    return ((fb.x | (fb.x >> 7) | (fb.x >> 14) | (fb.x >> 21)) & 0xfu) |
           (((fb.y | (fb.y >> 7) | (fb.y >> 14) | (fb.y >> 21)) & 0xfu) << 4);
  };
  auto load_rows = [&](const BlockPos& p, float4 (&z)[RB]) {
    const int na = rows_of(p);
#pragma unroll
    for (int rr = 0; rr < RB; ++rr)
      z[rr] = rr < na ? __ldg(reinterpret_cast<const float4*>(Zp + (p.r + rr) * KP) + l) : make_float4(0.f, 0.f, 0.f, 0.f);
    if (na > 0 && p.r + 2 * RB < p.r_end)  // two blocks further on towards L2 (512 B = four lines, one per lane)
      asm volatile("prefetch.global.L2 [%0];" ::"l"(Zp + (p.r + 2 * RB) * KP + 32 * l));
  };
  auto fetch_corr = [&](const BlockPos& p, unsigned cm, int half) {  // this lane's 16 bytes of every row of the block
    if (!CLIP) return;
    if (__any_sync(0xffffffffu, cm != 0)) {
      const uint32_t dst = (uint32_t)__cvta_generic_to_shared(dbuf + (half * NCH + ci) * ES + 4 * l);
#pragma unroll
      for (int rr = 0; rr < RB; ++rr) {
        const bool on = (cm >> rr) & 1u;
        const float* src = on ? Dp + (p.r + rr) * KP + 4 * l : Dp;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst + rr * KP * 4), "l"(src), "r"(on ? 16 : 0) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  BlockPos nxt = cur.sub < lay.nsub ? next_block(cur, chains_total, lay, bstar, sub_flags, RB) : cur;
  uint2 fb_cur = load_flags(cur), fb_nxt = load_flags(nxt);
  float4 zr[RB];
  load_rows(cur, zr);
  int par = 0;
  fetch_corr(cur, flag_mask(fb_cur), par);
  while (__any_sync(0xffffffffu, cur.sub < lay.nsub)) {
    const bool have = cur.sub < lay.nsub;
    const int64_t rb = cur.r;
    const int nact = rows_of(cur);  // active rows of this block
    const unsigned cmask_cur = flag_mask(fb_cur), cmask_nxt = flag_mask(fb_nxt);
    const BlockPos nxt2 = nxt.sub < lay.nsub ? next_block(nxt, chains_total, lay, bstar, sub_flags, RB) : nxt;
    const uint2 fb_nxt2 = load_flags(nxt2);
    if (have && cur.sub != loaded_sub) {
      const float4* src4 = reinterpret_cast<const float4*>(sub_P + (size_t)cur.sub * KP * KP + (size_t)(4 * l) * KP);
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int i4 = 0; i4 < 4; ++i4) {
          const float4 v = src4[q * 4 + i4];
          P[q][2 * i4] = make_float2(v.x, v.y);
          P[q][2 * i4 + 1] = make_float2(v.z, v.w);
        }
      Zrun = *reinterpret_cast<const float4*>(sub_Z + (size_t)cur.sub * KP + 4 * l);
      loaded_sub = cur.sub;
    }
    // deviations of the whole block from the running prefix: e_i = z_i - Z_i/(n+i)
    const bool blk_clip = CLIP && __any_sync(0xffffffffu, cmask_cur != 0);
    const float* db = dbuf + (par * NCH + ci) * ES;
    const double nb = (double)(n_eff + row0 + rb);
    float invr[RB];
#pragma unroll
    for (int rr = 0; rr < RB; ++rr) {
      const float ni = (float)(nb + (double)rr);
      const float inv = (rr < nact && ni > 0.f) ? __fdividef(1.f, ni) : 0.f;
      invr[rr] = inv;
      const float4 z = zr[rr];
      float4 e;
      e.x = fmaf(-Zrun.x, inv, z.x); e.y = fmaf(-Zrun.y, inv, z.y);
      e.z = fmaf(-Zrun.z, inv, z.z); e.w = fmaf(-Zrun.w, inv, z.w);
      if (rr >= nact) e = make_float4(0.f, 0.f, 0.f, 0.f);
      Zrun.x += z.x; Zrun.y += z.y; Zrun.z += z.z; Zrun.w += z.w;
      *reinterpret_cast<float4*>(eb + rr * KP + 4 * l) = e;
    }
    // the next block's rows and corrections travel while this block is walked
    fetch_corr(nxt, cmask_nxt, par ^ 1);
    load_rows(nxt, zr);
    asm volatile("cp.async.wait_group 1;" ::: "memory");  // this block's corrections have landed
    __syncwarp();
    float sc0 = 0.f, sc1 = 0.f;
#pragma unroll
    for (int pr = 0; pr < RB / 2; ++pr) {
      const int r1 = 2 * pr, r2 = 2 * pr + 1;
      const float4* e1p = reinterpret_cast<const float4*>(eb + r1 * KP);
      const float4* e2p = reinterpret_cast<const float4*>(eb + r2 * KP);
      float2 ca[4], da[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) ca[q] = da[q] = make_float2(0.f, 0.f);
#pragma unroll
      for (int i4 = 0; i4 < 4; ++i4) {
        const float4 x = e1p[i4], y = e2p[i4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          ca[q] = ffma2(P[q][2 * i4], make_float2(x.x, x.y), ca[q]);
          ca[q] = ffma2(P[q][2 * i4 + 1], make_float2(x.z, x.w), ca[q]);
          da[q] = ffma2(P[q][2 * i4], make_float2(y.x, y.y), da[q]);
          da[q] = ffma2(P[q][2 * i4 + 1], make_float2(y.z, y.w), da[q]);
        }
      }
      float c1[4], d2[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) { c1[q] = ca[q].x + ca[q].y; d2[q] = da[q].x + da[q].y; }
      const float4 e1m = e1p[l], e2m = e2p[l];  // this lane's four components of e1, e2
      float t11 = fmaf(e1m.x, c1[0], fmaf(e1m.y, c1[1], fmaf(e1m.z, c1[2], e1m.w * c1[3])));
      float t12 = fmaf(e1m.x, d2[0], fmaf(e1m.y, d2[1], fmaf(e1m.z, d2[2], e1m.w * d2[3])));
      float t22 = fmaf(e2m.x, d2[0], fmaf(e2m.y, d2[1], fmaf(e2m.z, d2[2], e2m.w * d2[3])));
#pragma unroll
      for (int o = 1; o <= 2; o <<= 1) {
        t11 += __shfl_xor_sync(0xffffffffu, t11, o);
        t12 += __shfl_xor_sync(0xffffffffu, t12, o);
        t22 += __shfl_xor_sync(0xffffffffu, t22, o);
      }
      // g = w/(1 + w t) with w = (n+i)/(n+i+1)  ==  1/(1 + 1/(n+i) + t)
      // With d1 = 1 + 1/(n+i) + t11 and s2 = 1 + 1/(n+i+1) + t22:  g1 = 1/d1,  t2 = t22 - t12^2/d1,
      // g2 = 1/(s2 - t12^2/d1) = d1/(d1 s2 - t12^2): the two reciprocals do not depend on each other.
      const bool act1 = invr[r1] > 0.f, act2 = invr[r2] > 0.f;
      const float d1 = 1.f + invr[r1] + t11, s2 = 1.f + invr[r2] + t22;
      const float rd1 = __fdividef(1.f, d1);
      const float rD = __fdividef(1.f, act1 ? fmaf(d1, s2, -t12 * t12) : s2);
      const float g1 = act1 ? rd1 : 0.f;
      const float g2 = act2 ? (act1 ? d1 * rD : rD) : 0.f;
      const float k12 = g1 * t12;
      float c2[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) c2[q] = fmaf(-k12, c1[q], d2[q]);
      const float t2 = fmaf(-k12, t12, t22);
      float q1 = t11, q2 = t2;
      if (CLIP && blk_clip) {
        const bool f1 = __any_sync(0xffffffffu, ((cmask_cur >> r1) & 1u) != 0);
        const bool f2 = __any_sync(0xffffffffu, ((cmask_cur >> r2) & 1u) != 0);
        if (f1 || f2) {  // warp-uniform
          float a11 = 0.f, b11 = 0.f, a22 = 0.f, b22 = 0.f, x12 = 0.f;
          if (f1) {
            const float4* dp = reinterpret_cast<const float4*>(db + r1 * KP);
            float2 pa[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) pa[q] = make_float2(0.f, 0.f);
#pragma unroll
            for (int i4 = 0; i4 < 4; ++i4) {
              const float4 x = dp[i4];
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                pa[q] = ffma2(P[q][2 * i4], make_float2(x.x, x.y), pa[q]);
                pa[q] = ffma2(P[q][2 * i4 + 1], make_float2(x.z, x.w), pa[q]);
              }
            }
            const float4 dm = dp[l];
            a11 = fmaf(dm.x, c1[0], fmaf(dm.y, c1[1], fmaf(dm.z, c1[2], dm.w * c1[3])));
            b11 = fmaf(dm.x, pa[0].x + pa[0].y, fmaf(dm.y, pa[1].x + pa[1].y,
                       fmaf(dm.z, pa[2].x + pa[2].y, dm.w * (pa[3].x + pa[3].y))));
          }
          if (f2) {
            const float4* dp = reinterpret_cast<const float4*>(db + r2 * KP);
            float2 pa[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) pa[q] = make_float2(0.f, 0.f);
#pragma unroll
            for (int i4 = 0; i4 < 4; ++i4) {
              const float4 x = dp[i4];
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                pa[q] = ffma2(P[q][2 * i4], make_float2(x.x, x.y), pa[q]);
                pa[q] = ffma2(P[q][2 * i4 + 1], make_float2(x.z, x.w), pa[q]);
              }
            }
            const float4 dm = dp[l];
            a22 = fmaf(dm.x, d2[0], fmaf(dm.y, d2[1], fmaf(dm.z, d2[2], dm.w * d2[3])));
            x12 = fmaf(dm.x, c1[0], fmaf(dm.y, c1[1], fmaf(dm.z, c1[2], dm.w * c1[3])));
            b22 = fmaf(dm.x, pa[0].x + pa[0].y, fmaf(dm.y, pa[1].x + pa[1].y,
                       fmaf(dm.z, pa[2].x + pa[2].y, dm.w * (pa[3].x + pa[3].y))));
          }
#pragma unroll
          for (int o = 1; o <= 2; o <<= 1) {
            a11 += __shfl_xor_sync(0xffffffffu, a11, o);
            b11 += __shfl_xor_sync(0xffffffffu, b11, o);
            a22 += __shfl_xor_sync(0xffffffffu, a22, o);
            b22 += __shfl_xor_sync(0xffffffffu, b22, o);
            x12 += __shfl_xor_sync(0xffffffffu, x12, o);
          }
          q1 = t11 + fmaf(2.f, a11, b11);
          const float cu = t12 + x12;  // c1 . u2
          q2 = fmaf(-g1 * cu, cu, t22 + fmaf(2.f, a22, b22));
        }
      }
      if (pr == l) { sc0 = q1; sc1 = q2; }
      // broadcast c1, c2 and apply both rank-1 terms to this lane's rows of P
      float* cb = cbufs + ci * CS + (pr & 1) * 2 * KP;
      *reinterpret_cast<float4*>(cb + 4 * l) = make_float4(c1[0], c1[1], c1[2], c1[3]);
      *reinterpret_cast<float4*>(cb + KP + 4 * l) = make_float4(c2[0], c2[1], c2[2], c2[3]);
      __syncwarp();
      float2 n1[4], n2[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        n1[q] = make_float2(-g1 * c1[q], -g1 * c1[q]);
        n2[q] = make_float2(-g2 * c2[q], -g2 * c2[q]);
      }
#pragma unroll
      for (int i4 = 0; i4 < 4; ++i4) {
        const float4 x1 = *reinterpret_cast<const float4*>(cb + 4 * i4);
        const float4 x2 = *reinterpret_cast<const float4*>(cb + KP + 4 * i4);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          P[q][2 * i4] = ffma2(n1[q], make_float2(x1.x, x1.y), P[q][2 * i4]);
          P[q][2 * i4 + 1] = ffma2(n1[q], make_float2(x1.z, x1.w), P[q][2 * i4 + 1]);
          P[q][2 * i4] = ffma2(n2[q], make_float2(x2.x, x2.y), P[q][2 * i4]);
          P[q][2 * i4 + 1] = ffma2(n2[q], make_float2(x2.z, x2.w), P[q][2 * i4 + 1]);
        }
      }
    }
    // lane l holds the quadratic forms of rows 2l and 2l+1 of the block
    {
      const int ra = 2 * l;
      const float s0 = ((float)(nb + (double)ra) - 1.f) * sc0;
      const float s1 = ((float)(nb + (double)(ra + 1)) - 1.f) * sc1;
      const int f0 = (double)s0 > threshold, f1 = (double)s1 > threshold;
      if (ra + 1 < nact) {
        *reinterpret_cast<float2*>(score + rb + ra) = make_float2(s0, s1);
        *reinterpret_cast<int2*>(flag + rb + ra) = make_int2(f0, f1);
        nf += f0 + f1;
        ss += (double)s0 + (double)s1;
      } else if (ra < nact) {
        score[rb + ra] = s0;
        flag[rb + ra] = f0;
        nf += f0;
        ss += (double)s0;
      }
    }
    cur = nxt;
    nxt = nxt2;
    fb_cur = fb_nxt;
    fb_nxt = fb_nxt2;
    par ^= 1;
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    nf += __shfl_xor_sync(0xffffffffu, nf, o);
    ss += __shfl_xor_sync(0xffffffffu, ss, o);
  }
  if (lane == 0 && (nf != 0 || ss != 0.0)) {
    atomicAdd((unsigned long long*)(gate + 2), (unsigned long long)nf);
    atomicAdd(score_sum, ss);
  }
}

// PC0 = V^T C0 V (KP x KP, zero padded), one block.  T = C0 V first: C0 is symmetric, so thread (feature f, column
// group) walks column f of C0 - consecutive threads read consecutive addresses - with the rows of V broadcast from
// shared memory; the loads are independent of the accumulation and pipeline freely (the first version read C0 by rows,
// two distinct addresses per warp load: 150 us at d = 256, a visible part of a 2 Mi-row shard's batch).
__global__ void __launch_bounds__(1024)
project_cov(const double* __restrict__ C0, const double* __restrict__ V, int d, int KP, double* __restrict__ PC0,
            int stage_v) {
  extern __shared__ double T[];  // [d][KP] = C0 V, then (where it fits) [d][KP] V staged behind it
  const double* Vs = V;
  if (stage_v) {
    double* vs = T + (size_t)d * KP;
    for (int idx = threadIdx.x; idx < d * KP; idx += blockDim.x) vs[idx] = V[idx];
    Vs = vs;
    __syncthreads();
  }
  const int ngrp = KP / 4;  // KP is 4, 8, 16 or 32: four columns per thread
  for (int idx = threadIdx.x; idx < d * ngrp; idx += blockDim.x) {
    const int f = idx % d, c0 = 4 * (idx / d);
    double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll 8
    for (int g = 0; g < d; ++g) {
      const double c = C0[(size_t)g * d + f];
      const double* v = Vs + g * KP + c0;
      a0 = fma(c, v[0], a0); a1 = fma(c, v[1], a1); a2 = fma(c, v[2], a2); a3 = fma(c, v[3], a3);
    }
    double* t = T + f * KP + c0;
    t[0] = a0; t[1] = a1; t[2] = a2; t[3] = a3;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < KP * KP; idx += blockDim.x) {
    const int a = idx / KP, b = idx - a * KP;
    double acc0 = 0.0, acc1 = 0.0;
    for (int f = 0; f + 1 < d; f += 2) {
      acc0 = fma(Vs[f * KP + a], T[f * KP + b], acc0);
      acc1 = fma(Vs[(f + 1) * KP + a], T[(f + 1) * KP + b], acc1);
    }
    if (d & 1) acc0 = fma(Vs[(d - 1) * KP + a], T[(d - 1) * KP + b], acc0);
    PC0[idx] = acc0 + acc1;
  }
}

// one thread per projected statistic (KQ = KP + KP^2), whole warps
inline int scan_threads(int KQ) { return KQ >= 1024 ? 1024 : ((KQ + 31) & ~31); }

template <int KP>
int launch_prefix_kp(sb_maha* h, const MahaWs& ws, double* rank_proj, cudaStream_t st) {
  const int d = h->d, nsub = h->lay.nsub;
  const int KQ = KP + KP * KP;
  const size_t PS = (size_t)d + (size_t)d * d;
  if (h->profile) cudaEventRecord(h->ev[4], st);
  const int stage_v = (size_t)2 * d * KP * sizeof(double) <= 200 * 1024;
  const size_t pc_smem = (size_t)(stage_v ? 2 : 1) * d * KP * sizeof(double);
  SB_CUDA(sb::func_attr_cached(project_cov, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  project_cov<<<1, 1024, pc_smem, st>>>(h->d_C, h->d_V, d, KP, h->d_PC0, stage_v);
  SB_LAUNCH_CHECK();
  int grid = nsub < kNumSMs * 8 ? nsub : kNumSMs * 8;
  const int FC = d < 8192 / d ? d : 8192 / d;
  const size_t sp_smem = ((((size_t)FC * d + 3) & ~(size_t)3) + (size_t)FC * KP) * sizeof(float);
  if (h->row0 == 0) {
    // no rows before this shard (single GPU, or rank 0): the offset of every prefix is zero
    SB_CUDA(cudaMemsetAsync(rank_proj, 0, (size_t)KQ * sizeof(double), st));
  } else {
    sub_project<double><<<1, 256, sp_smem, st>>>(ws.rank_pre + 1, PS, 1, d, KP, h->d_Vf, rank_proj);
    SB_LAUNCH_CHECK();
  }
  if (h->split_pass2) {
    // the projection kernel of pass 2 produces the per-sub-segment statistics; scan and start inverses follow it
    if (h->profile) cudaEventRecord(h->ev[5], st);
    return SB_OK;
  }
  sub_project<float><<<grid, 256, sp_smem, st>>>(ws.partials, PS, nsub, d, KP, h->d_Vf, ws.subproj);
  SB_LAUNCH_CHECK();
  const int nchunk = (int)ceil_div(nsub, kScanChunk);
  sub_scan_totals<<<nchunk, scan_threads(KQ), 0, st>>>(ws.subproj, nsub, KQ, ws.chunk_tot);
  SB_LAUNCH_CHECK();
  sub_scan_apply<<<nchunk, scan_threads(KQ), 0, st>>>(ws.subproj, nsub, KQ, ws.chunk_tot, rank_proj);
  SB_LAUNCH_CHECK();
  sub_start<KP><<<(int)ceil_div(nsub, 4), 128, 0, st>>>(ws.subproj, h->d_PC0, h->lay, h->k, (long long)h->n_eff,
                                                       (long long)h->row0, ws.sub_P, ws.sub_Z, ws.sub_flags);
  SB_LAUNCH_CHECK();
  gate_init<<<1, 32, 0, st>>>(ws.sub_flags, h->lay, (long long)h->row0, h->d_gate, h->d_score_sum);
  if (h->profile) cudaEventRecord(h->ev[5], st);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

template <int KP>
int launch_score_kp(sb_maha* h, const float* X, int64_t ld, float* score, int32_t* flag, const MahaWs& ws,
                    cudaStream_t st) {
  const int d = h->d;
  const SubLayout& L = h->split_pass2 ? h->lay2 : h->lay;
  const int nsub = L.nsub;
  const bool aligned = (ld % 4 == 0) && (((uintptr_t)X & 15) == 0);
  const bool fast = KP == 16 && d == 64 && !h->force_generic && aligned;
  const bool split = h->split_pass2 != 0;
  if (h->profile) { cudaEventRecord(h->ev[6], st); cudaEventRecord(h->ev[7], st); }
  if (split) {
    // projection (z, clip correction, per-sub-segment sums) -> scan -> start inverses -> gate
    const int NBv = d / 64;
    int pgrid = (int)ceil_div(nsub, 4);
    const int pmax = kNumSMs * (NBv == 1 ? 4 : 3);
    if (pgrid > pmax) pgrid = pmax;
#define SB_LAUNCH_PROJECT2(NB, CLIPV, STDV)                                                                         \
    do {                                                                                                           \
      const size_t psm = proj_smem_bytes(NB, CLIPV);                                                               \
      SB_CUDA(sb::func_attr_cached(project_rows<NB, CLIPV, STDV>, cudaFuncAttributeMaxDynamicSharedMemorySize,     \
                                   (int)psm));                                                                     \
      project_rows<NB, CLIPV, STDV><<<pgrid, 128, psm, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f,          \
                                                            h->d_in_mu, h->d_in_is, h->d_Vf, ws.proj_Z, ws.proj_D, \
                                                            ws.proj_flag, ws.subproj);                             \
    } while (0)
#define SB_LAUNCH_PROJECT(NB, CLIPV)                                                                               \
    do {                                                                                                           \
      if (h->d_in_mu) SB_LAUNCH_PROJECT2(NB, CLIPV, true); else SB_LAUNCH_PROJECT2(NB, CLIPV, false);              \
    } while (0)
    if (h->use_clip) {
      if (NBv == 1) SB_LAUNCH_PROJECT(1, true); else if (NBv == 2) SB_LAUNCH_PROJECT(2, true); else SB_LAUNCH_PROJECT(4, true);
    } else {
      if (NBv == 1) SB_LAUNCH_PROJECT(1, false); else if (NBv == 2) SB_LAUNCH_PROJECT(2, false); else SB_LAUNCH_PROJECT(4, false);
    }
#undef SB_LAUNCH_PROJECT2
#undef SB_LAUNCH_PROJECT
    SB_LAUNCH_CHECK();
    if (h->profile) cudaEventRecord(h->ev[7], st);
    constexpr int KQ = KP + KP * KP;
    const int nchunk = (int)ceil_div(nsub, kScanChunk);
    sub_scan_totals<<<nchunk, scan_threads(KQ), 0, st>>>(ws.subproj, nsub, KQ, ws.chunk_tot);
    SB_LAUNCH_CHECK();
    sub_scan_apply<<<nchunk, scan_threads(KQ), 0, st>>>(ws.subproj, nsub, KQ, ws.chunk_tot, ws.rank_proj);
    SB_LAUNCH_CHECK();
    sub_start<KP><<<(int)ceil_div(nsub, 4), 128, 0, st>>>(ws.subproj, h->d_PC0, L, h->k, (long long)h->n_eff,
                                                         (long long)h->row0, ws.sub_P, ws.sub_Z, ws.sub_flags);
    SB_LAUNCH_CHECK();
    gate_init<<<1, 32, 0, st>>>(ws.sub_flags, L, (long long)h->row0, h->d_gate, h->d_score_sum);
    SB_LAUNCH_CHECK();
  }
  // exact fp64 warm-up path (does nothing in the steady state: b* = 0 and every start is well conditioned)
  {
    int grid = nsub < kNumSMs * 16 ? nsub : kNumSMs * 16;
    if (h->use_clip)
      score_slow<KP, true><<<grid, 32, 0, st>>>(X, ld, d, h->k, L, h->d_mean, h->d_lo, h->d_hi, h->d_V, h->d_PC0,
                                               ws.subproj, ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                               h->d_gate, score, h->d_in_mu, h->d_in_is);
    else
      score_slow<KP, false><<<grid, 32, 0, st>>>(X, ld, d, h->k, L, h->d_mean, h->d_lo, h->d_hi, h->d_V, h->d_PC0,
                                                ws.subproj, ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                                h->d_gate, score, h->d_in_mu, h->d_in_is);
    SB_LAUNCH_CHECK();
    int g2 = (int)ceil_div(nsub, 4);
    if (g2 > kNumSMs * 4) g2 = kNumSMs * 4;
    slow_fixup<<<g2, 128, 0, st>>>(L, ws.sub_flags, (long long)h->row0, h->threshold, h->d_gate, h->d_score_sum, score,
                                   flag);
    SB_LAUNCH_CHECK();
  }
  if (h->profile) cudaEventRecord(h->ev[8], st);
  if (split) {
    if constexpr (KP == 16) {
      constexpr int NCH = kChain4Chains, ES = kChain4Rows * 16 + 4, CS = 2 * 2 * 16 + 4;
      int cgrid = (int)ceil_div(nsub, NCH);
      if (cgrid > kNumSMs * kChain4Ctas) cgrid = kNumSMs * kChain4Ctas;
      if (h->use_clip) {
        const size_t csm = (size_t)(NCH * ES + NCH * CS + 2 * NCH * ES) * sizeof(float);
        SB_CUDA(sb::func_attr_cached(chain4<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csm));
        chain4<true><<<cgrid, 32 * kChain4Warps, csm, st>>>(ws.proj_Z, ws.proj_D, ws.proj_flag, L, ws.sub_P, ws.sub_Z,
                                                           ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                                           h->threshold, h->d_gate, h->d_score_sum, score, flag);
      } else {
        const size_t csm = (size_t)(NCH * ES + NCH * CS) * sizeof(float);
        chain4<false><<<cgrid, 32 * kChain4Warps, csm, st>>>(ws.proj_Z, ws.proj_D, ws.proj_flag, L, ws.sub_P, ws.sub_Z,
                                                            ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                                            h->threshold, h->d_gate, h->d_score_sum, score, flag);
      }
    }
  } else if (fast) {
    if constexpr (KP == 16) {
      int grid = (int)ceil_div(nsub, 8);
      if (grid > kNumSMs * 3) grid = kNumSMs * 3;
      if (h->use_clip)
        score_chain64<true><<<grid, 128, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, h->d_Vf, ws.sub_P,
                                                 ws.sub_Z, ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                                 h->threshold, h->d_gate, h->d_score_sum, score, flag);
      else
        score_chain64<false><<<grid, 128, 0, st>>>(X, ld, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, h->d_Vf, ws.sub_P,
                                                  ws.sub_Z, ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                                  h->threshold, h->d_gate, h->d_score_sum, score, flag);
    }
  } else {
    const int groups_per_cta = 128 / KP;
    int grid = (int)ceil_div(nsub, groups_per_cta);
    if (grid > kNumSMs * 4) grid = kNumSMs * 4;
    size_t smem = ((size_t)d * KP + 3 * (size_t)d + 4 + (size_t)groups_per_cta * 3 * KP) * sizeof(float);
    if (h->use_clip) {
      SB_CUDA(sb::func_attr_cached(score_chain<KP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      score_chain<KP, true><<<grid, 128, smem, st>>>(X, ld, d, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, h->d_Vf, ws.sub_P,
                                                    ws.sub_Z, ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                                    h->threshold, h->d_gate, h->d_score_sum, score, flag);
    } else {
      SB_CUDA(sb::func_attr_cached(score_chain<KP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      score_chain<KP, false><<<grid, 128, smem, st>>>(X, ld, d, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, h->d_Vf,
                                                     ws.sub_P, ws.sub_Z, ws.sub_flags, (long long)h->n_eff,
                                                     (long long)h->row0, h->threshold, h->d_gate, h->d_score_sum,
                                                     score, flag);
    }
  }
  if (h->profile) cudaEventRecord(h->ev[9], st);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

}  // namespace

int launch_prefix(sb_maha* h, const MahaWs& ws, double* rank_proj, cudaStream_t st) {
  switch (h->KP) {
    case 4: return launch_prefix_kp<4>(h, ws, rank_proj, st);
    case 8: return launch_prefix_kp<8>(h, ws, rank_proj, st);
    case 16: return launch_prefix_kp<16>(h, ws, rank_proj, st);
    case 32: return launch_prefix_kp<32>(h, ws, rank_proj, st);
  }
  set_error("unsupported padded component count %d", h->KP);
  return SB_ERR_INVALID;
}

int launch_score(sb_maha* h, const float* X, int64_t ld, float* score, int32_t* flag, const MahaWs& ws,
                 cudaStream_t st) {
  switch (h->KP) {
    case 4: return launch_score_kp<4>(h, X, ld, score, flag, ws, st);
    case 8: return launch_score_kp<8>(h, X, ld, score, flag, ws, st);
    case 16: return launch_score_kp<16>(h, X, ld, score, flag, ws, st);
    case 32: return launch_score_kp<32>(h, X, ld, score, flag, ws, st);
  }
  set_error("unsupported padded component count %d", h->KP);
  return SB_ERR_INVALID;
}

}  // namespace sb
