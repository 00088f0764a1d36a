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
template <typename TIn>
__global__ void __launch_bounds__(128)
sub_project(const TIn* __restrict__ recs, size_t rec_stride, int nitems, int d, int KP, const double* __restrict__ V,
            double* __restrict__ out) {
  constexpr int FC = 32;
  __shared__ double Tc[FC * kMaxKP];
  const int KQ = KP + KP * KP;
  const int tid = threadIdx.x;
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const TIn* s = recs + (size_t)item * rec_stride;
    const TIn* G = s + d;
    double* o = out + (size_t)item * KQ;
    double qacc[8];  // KP*KP <= 1024 = 8 * 128
#pragma unroll
    for (int u = 0; u < 8; ++u) qacc[u] = 0.0;
    for (int f0 = 0; f0 < d; f0 += FC) {
      __syncthreads();
      for (int idx = tid; idx < FC * KP; idx += blockDim.x) {
        const int ff = idx / KP, c = idx - ff * KP, f = f0 + ff;
        double acc = 0.0;
        if (f < d) {
          for (int g = 0; g < d; ++g) {
            const size_t gi = f <= g ? (size_t)f * d + g : (size_t)g * d + f;  // upper triangle is authoritative
            acc = fma((double)G[gi], V[(size_t)g * KP + c], acc);
          }
        }
        Tc[idx] = acc;
      }
      __syncthreads();
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int pq = tid + u * 128;
        if (pq < KP * KP) {
          const int a = pq / KP, b = pq - a * KP;
          double acc = qacc[u];
          for (int ff = 0; ff < FC && f0 + ff < d; ++ff) acc = fma(V[(size_t)(f0 + ff) * KP + a], Tc[ff * KP + b], acc);
          qacc[u] = acc;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int pq = tid + u * 128;
      if (pq < KP * KP) o[KP + pq] = qacc[u];
    }
    if (tid < KP) {
      double acc = 0.0;
      for (int f = 0; f < d; ++f) acc = fma((double)s[f], V[(size_t)f * KP + tid], acc);
      o[tid] = acc;
    }
  }
}

// exclusive scan over sub-segments of every projected component, seeded with the rank offset; row nsub = totals
__global__ void __launch_bounds__(256)
sub_scan(double* __restrict__ subproj, int nsub, int KQ, const double* __restrict__ rank_off) {
  const int comp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (comp >= KQ) return;
  const int chunk = (nsub + 31) / 32;
  const int s0 = lane * chunk, s1 = min(nsub, s0 + chunk);
  double tot = 0.0;
  for (int s = s0; s < s1; ++s) tot += subproj[(size_t)s * KQ + comp];
  double incl = tot;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    double up = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += up;
  }
  double run = rank_off[comp] + (incl - tot);
  for (int s = s0; s < s1; ++s) {
    double v = subproj[(size_t)s * KQ + comp];
    subproj[(size_t)s * KQ + comp] = run;
    run += v;
  }
  const double total = __shfl_sync(0xffffffffu, incl, 31) + rank_off[comp];
  if (lane == 0) subproj[(size_t)nsub * KQ + comp] = total;
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
           long long n_eff, long long row0, long long* __restrict__ gate, float* __restrict__ score) {
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
        const double xv = (double)x[f];
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
__device__ __forceinline__ uint32_t to_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void mma_m16n8k8_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                                 uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

constexpr int kZStride = 20;  // floats per projected row in shared memory (16 + pad, keeps float4 alignment)

// A warp holds two groups of 16 lanes = two sub-segments.  Per step each group's next 16 rows are projected by the
// whole warp (M=16 rows, K=64 features, N=16 components; lane (g,t) loads columns 16q+4t..+3 of rows g and g+8, so
// every warp load covers 64 contiguous bytes of 8 rows), the projections go to shared memory, then both groups walk
// their 16 rows with chain_step.
template <bool CLIP>
__global__ void __launch_bounds__(128)
score_chain64(const float* __restrict__ X, int64_t ld, SubLayout lay, const float* __restrict__ mean_f,
              const float* __restrict__ lo_f, const float* __restrict__ hi_f, const float* __restrict__ Vf,
              const float* __restrict__ sub_P, const float* __restrict__ sub_Z, const int32_t* __restrict__ sub_flags,
              long long n_eff, long long row0, double threshold, long long* __restrict__ gate,
              double* __restrict__ score_sum, float* __restrict__ score, int32_t* __restrict__ flag) {
  constexpr int KP = 16, D = 64;
  __shared__ __align__(16) float ms[D], los[D], his[D];
  __shared__ __align__(16) float zbuf[4][2][16 * kZStride];  // [warp][group][row][comp]  clipped projection z
  __shared__ __align__(16) float dbuf[4][2][16 * kZStride];  // y - z
  __shared__ int clipf[4][2][16];
  __shared__ __align__(16) float cbufs[8][3 * KP];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int j = lane & 15, grp = lane >> 4;
  for (int f = threadIdx.x; f < D; f += blockDim.x) {
    ms[f] = mean_f[f];
    los[f] = CLIP ? lo_f[f] : 0.f;
    his[f] = CLIP ? hi_f[f] : 0.f;
  }
  // V fragments (constant for the whole batch): step ks uses physical rows 16q+4t+e0 (k=t) and +1 (k=t+4)
  uint32_t vb[8][2][2], vs[8][2][2];
#pragma unroll
  for (int ks = 0; ks < 8; ++ks) {
    const int f0 = 16 * (ks >> 1) + 4 * t + 2 * (ks & 1);
#pragma unroll
    for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float v = Vf[(f0 + h) * KP + 8 * nt + g];
        vb[ks][nt][h] = to_tf32(v);
        vs[ks][nt][h] = to_tf32(v - __uint_as_float(vb[ks][nt][h]));
      }
    }
  }
  __syncthreads();
  float* buf = cbufs[warp * 2 + grp];
  const int groups_total = gridDim.x * 8;
  const int gid = blockIdx.x * 8 + warp * 2 + grp;
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
    // ---- projection of both groups' blocks by the whole warp
#pragma unroll
    for (int gs = 0; gs < 2; ++gs) {
      const int src = gs * 16;  // any lane of group gs knows its cursor
      const int64_t rb = __shfl_sync(0xffffffffu, cur.r, src);
      const int64_t re = __shfl_sync(0xffffffffu, cur.r_end, src);
      const bool ghave = __shfl_sync(0xffffffffu, (int)have, src) != 0;
      if (!ghave) continue;  // warp-uniform
      const bool va = rb + g < re, vb8 = rb + g + 8 < re;
      float4 xa[4], xb[4];
      const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
      const float4* pa = reinterpret_cast<const float4*>(X + (rb + g) * ld);
      const float4* pb = reinterpret_cast<const float4*>(X + (rb + g + 8) * ld);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        xa[q] = va ? ldg_stream_f4(pa + 4 * q + t) : zero4;
        xb[q] = vb8 ? ldg_stream_f4(pb + 4 * q + t) : zero4;
      }
      float accz[2][4], accy[2][4];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) { accz[nt][e] = 0.f; accy[nt][e] = 0.f; }
      int clip_a = 0, clip_b = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 m4 = *reinterpret_cast<const float4*>(ms + 16 * q + 4 * t);
        const float4 l4 = *reinterpret_cast<const float4*>(los + 16 * q + 4 * t);
        const float4 h4 = *reinterpret_cast<const float4*>(his + 16 * q + 4 * t);
        const float mm[4] = {m4.x, m4.y, m4.z, m4.w}, ll[4] = {l4.x, l4.y, l4.z, l4.w}, hh[4] = {h4.x, h4.y, h4.z, h4.w};
        const float ra[4] = {xa[q].x, xa[q].y, xa[q].z, xa[q].w}, rbv[4] = {xb[q].x, xb[q].y, xb[q].z, xb[q].w};
        uint32_t zb[2][4], zs[2][4], yb[2][4], ys[2][4];  // [row g / g+8][element]
#pragma unroll
        for (int e = 0; e < 4; ++e) {
#pragma unroll
          for (int rr = 0; rr < 2; ++rr) {
            const float x = rr == 0 ? ra[e] : rbv[e];
            const bool valid = rr == 0 ? va : vb8;
            float xc = x;
            if (CLIP) xc = fminf(fmaxf(x, ll[e]), hh[e]);
            if (CLIP && valid && xc != x) { if (rr == 0) clip_a = 1; else clip_b = 1; }
            const float dz = valid ? xc - mm[e] : 0.f;
            zb[rr][e] = to_tf32(dz);
            zs[rr][e] = to_tf32(dz - __uint_as_float(zb[rr][e]));
            if (CLIP) {
              const float dy = valid ? x - mm[e] : 0.f;
              yb[rr][e] = to_tf32(dy);
              ys[rr][e] = to_tf32(dy - __uint_as_float(yb[rr][e]));
            }
          }
        }
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          const int ks = 2 * q + half, e0 = 2 * half;
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            mma_m16n8k8_tf32(accz[nt], zs[0][e0], zs[1][e0], zs[0][e0 + 1], zs[1][e0 + 1], vb[ks][nt][0], vb[ks][nt][1]);
            mma_m16n8k8_tf32(accz[nt], zb[0][e0], zb[1][e0], zb[0][e0 + 1], zb[1][e0 + 1], vs[ks][nt][0], vs[ks][nt][1]);
            mma_m16n8k8_tf32(accz[nt], zb[0][e0], zb[1][e0], zb[0][e0 + 1], zb[1][e0 + 1], vb[ks][nt][0], vb[ks][nt][1]);
            if (CLIP) {
              mma_m16n8k8_tf32(accy[nt], ys[0][e0], ys[1][e0], ys[0][e0 + 1], ys[1][e0 + 1], vb[ks][nt][0], vb[ks][nt][1]);
              mma_m16n8k8_tf32(accy[nt], yb[0][e0], yb[1][e0], yb[0][e0 + 1], yb[1][e0 + 1], vs[ks][nt][0], vs[ks][nt][1]);
              mma_m16n8k8_tf32(accy[nt], yb[0][e0], yb[1][e0], yb[0][e0 + 1], yb[1][e0 + 1], vb[ks][nt][0], vb[ks][nt][1]);
            }
          }
        }
      }
      float* zrow = zbuf[warp][gs];
      float* drow = dbuf[warp][gs];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        const int c0 = 8 * nt + 2 * t;
        *reinterpret_cast<float2*>(zrow + g * kZStride + c0) = make_float2(accz[nt][0], accz[nt][1]);
        *reinterpret_cast<float2*>(zrow + (g + 8) * kZStride + c0) = make_float2(accz[nt][2], accz[nt][3]);
        if (CLIP) {
          *reinterpret_cast<float2*>(drow + g * kZStride + c0) =
              make_float2(accy[nt][0] - accz[nt][0], accy[nt][1] - accz[nt][1]);
          *reinterpret_cast<float2*>(drow + (g + 8) * kZStride + c0) =
              make_float2(accy[nt][2] - accz[nt][2], accy[nt][3] - accz[nt][3]);
        }
      }
      if (CLIP) {
        clip_a |= __shfl_xor_sync(0xffffffffu, clip_a, 1);
        clip_a |= __shfl_xor_sync(0xffffffffu, clip_a, 2);
        clip_b |= __shfl_xor_sync(0xffffffffu, clip_b, 1);
        clip_b |= __shfl_xor_sync(0xffffffffu, clip_b, 2);
        if (t == 0) { clipf[warp][gs][g] = clip_a; clipf[warp][gs][g + 8] = clip_b; }
      }
    }
    __syncwarp();
    // ---- both groups walk their 16 rows
    const int64_t rb = cur.r;
    const float* zrow = zbuf[warp][grp];
    const float* drow = dbuf[warp][grp];
    float mysc = 0.f;
#pragma unroll 4
    for (int rr = 0; rr < 16; ++rr) {
      const bool active = have && (rb + rr < cur.r_end);
      const float z = active ? zrow[rr * kZStride + j] : 0.f;
      bool clipped = false;
      float dl = 0.f;
      if (CLIP) {
        clipped = active && clipf[warp][grp][rr] != 0;
        dl = clipped ? drow[rr * kZStride + j] : 0.f;
      }
      const bool need_u = CLIP && __any_sync(0xffffffffu, clipped);
      const float ni = (float)(double)(n_eff + row0 + rb + rr);
      const float sc = chain_step<KP>(P, Zrun, z, dl, need_u, active, ni, buf, j);
      if (j == rr) mysc = sc;
    }
    if (have && rb + j < cur.r_end) {
      score[rb + j] = mysc;
      const int f = (double)mysc > threshold;
      flag[rb + j] = f;
      nf += f;
      ss += (double)mysc;
    }
    if (have) cur.r += 16;
    __syncwarp();
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
}

// PC0 = V^T C0 V (KP x KP, zero padded), one block
__global__ void __launch_bounds__(256)
project_cov(const double* __restrict__ C0, const double* __restrict__ V, int d, int KP, double* __restrict__ PC0) {
  extern __shared__ double T[];  // [d][KP] = C0 V
  for (int idx = threadIdx.x; idx < d * KP; idx += blockDim.x) {
    const int f = idx / KP, c = idx - f * KP;
    double acc = 0.0;
    for (int g = 0; g < d; ++g) acc = fma(C0[(size_t)f * d + g], V[(size_t)g * KP + c], acc);
    T[idx] = acc;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < KP * KP; idx += blockDim.x) {
    const int a = idx / KP, b = idx - a * KP;
    double acc = 0.0;
    for (int f = 0; f < d; ++f) acc = fma(V[(size_t)f * KP + a], T[f * KP + b], acc);
    PC0[idx] = acc;
  }
}

template <int KP>
int launch_prefix_kp(sb_maha* h, const MahaWs& ws, double* rank_proj, cudaStream_t st) {
  const int d = h->d, nsub = h->lay.nsub;
  const int KQ = KP + KP * KP;
  const size_t PS = (size_t)d + (size_t)d * d;
  if (h->profile) cudaEventRecord(h->ev[4], st);
  SB_CUDA(cudaFuncSetAttribute(project_cov, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxD * kMaxKP * 8));
  project_cov<<<1, 256, (size_t)d * KP * sizeof(double), st>>>(h->d_C, h->d_V, d, KP, h->d_PC0);
  SB_LAUNCH_CHECK();
  int grid = nsub < kNumSMs * 8 ? nsub : kNumSMs * 8;
  sub_project<float><<<grid, 128, 0, st>>>(ws.partials, PS, nsub, d, KP, h->d_V, ws.subproj);
  SB_LAUNCH_CHECK();
  sub_project<double><<<1, 128, 0, st>>>(ws.rank_pre + 1, PS, 1, d, KP, h->d_V, rank_proj);
  SB_LAUNCH_CHECK();
  sub_scan<<<(int)ceil_div(KQ, 8), 256, 0, st>>>(ws.subproj, nsub, KQ, rank_proj);
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
  const int d = h->d, nsub = h->lay.nsub;
  const SubLayout& L = h->lay;
  // exact fp64 warm-up path (does nothing in the steady state: b* = 0 and every start is well conditioned)
  {
    int grid = nsub < kNumSMs * 16 ? nsub : kNumSMs * 16;
    if (h->use_clip)
      score_slow<KP, true><<<grid, 32, 0, st>>>(X, ld, d, h->k, L, h->d_mean, h->d_lo, h->d_hi, h->d_V, h->d_PC0,
                                               ws.subproj, ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                               h->d_gate, score);
    else
      score_slow<KP, false><<<grid, 32, 0, st>>>(X, ld, d, h->k, L, h->d_mean, h->d_lo, h->d_hi, h->d_V, h->d_PC0,
                                                ws.subproj, ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                                h->d_gate, score);
    SB_LAUNCH_CHECK();
    int g2 = (int)ceil_div(nsub, 4);
    if (g2 > kNumSMs * 4) g2 = kNumSMs * 4;
    slow_fixup<<<g2, 128, 0, st>>>(L, ws.sub_flags, (long long)h->row0, h->threshold, h->d_gate, h->d_score_sum, score,
                                   flag);
    SB_LAUNCH_CHECK();
  }
  const bool fast = KP == 16 && d == 64 && !h->force_generic && (ld % 4 == 0) && (((uintptr_t)X & 15) == 0);
  if (h->profile) cudaEventRecord(h->ev[6], st);
  if (fast) {
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
  } else {
    const int groups_per_cta = 128 / KP;
    int grid = (int)ceil_div(nsub, groups_per_cta);
    if (grid > kNumSMs * 4) grid = kNumSMs * 4;
    size_t smem = ((size_t)d * KP + 3 * (size_t)d + 4 + (size_t)groups_per_cta * 3 * KP) * sizeof(float);
    if (h->use_clip) {
      SB_CUDA(cudaFuncSetAttribute(score_chain<KP, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      score_chain<KP, true><<<grid, 128, smem, st>>>(X, ld, d, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, h->d_Vf, ws.sub_P,
                                                    ws.sub_Z, ws.sub_flags, (long long)h->n_eff, (long long)h->row0,
                                                    h->threshold, h->d_gate, h->d_score_sum, score, flag);
    } else {
      SB_CUDA(cudaFuncSetAttribute(score_chain<KP, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      score_chain<KP, false><<<grid, 128, smem, st>>>(X, ld, d, L, h->d_mean_f, h->d_lo_f, h->d_hi_f, h->d_Vf,
                                                     ws.sub_P, ws.sub_Z, ws.sub_flags, (long long)h->n_eff,
                                                     (long long)h->row0, h->threshold, h->d_gate, h->d_score_sum,
                                                     score, flag);
    }
  }
  if (h->profile) cudaEventRecord(h->ev[7], st);
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
