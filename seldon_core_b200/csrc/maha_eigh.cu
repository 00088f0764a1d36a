// Top-k eigenvectors of the d x d covariance by one-sided (Hestenes) Jacobi in fp64.
//
// Replaces the scipy.linalg.eigh(cov_batch, eigvals=(p-k, p-1)) call of CoreMahalanobis.py:103 (LAPACK, a
// third-party dependency of the reference).  Only the span of the k leading eigenvectors enters the scores
// (they are invariant to the basis of that subspace), so the solver is free to return any
// orthonormal basis; parity is tested on the projector V V^T.
//
// One CTA, 32 warps.  W starts as C (columns), V as I; a rotation of columns (p,q) that makes W_p and W_q
// orthogonal is applied to both.  At convergence W = C V has orthogonal columns, i.e. V diagonalises C^2 and
// hence C, with |lambda_j| = ||W_j||.  Per step the d/2 disjoint pairs of a round-robin tournament are rotated
// in parallel, one warp per pair, dot products by warp shuffles.  For 2*d*d*8 B <= 200 KB both matrices live
// in shared memory (d <= 111), otherwise in the (L2-resident) workspace.  d = 64 (jacobi64_regs) and d = 128 / 256
// (jacobi_cluster) keep the columns in registers instead and exchange them between neighbouring pairs.  This is a latency kernel
// (about 4*d^3 flops per sweep), not a roofline kernel; bench.py reports its time.
#include <cooperative_groups.h>
#include <stdlib.h>

#include "maha.cuh"

namespace sb {
namespace {

// pairs whose cos^2 is below this are in the quadratically convergent regime: after rotating them the remaining
// cosines are of the order of cos^2 <= 1e-14, below the 1e-11 working threshold
constexpr double kQuadCos2 = 1e-14;

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <bool SMEM>
__global__ void __launch_bounds__(1024, 1)
jacobi_topk(const double* __restrict__ C, int d, int k, int KP, double* gW, double* gV, double* __restrict__ Vd,
            float* __restrict__ Vf, double* __restrict__ evals, long long* __restrict__ gate,
            double* __restrict__ Vfull, int warm) {
  extern __shared__ double sm[];
  double* W = SMEM ? sm : gW;
  double* V = SMEM ? sm + (size_t)d * d : gV;
  __shared__ int rotated;
  __shared__ double norms[kMaxD];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = blockDim.x >> 5;

  if (warm) {
    // start from the basis of the previous batch: C changes little between batches, so W = C V is already close to
    // having orthogonal columns and 2-3 sweeps finish the job instead of ~11 from the identity
    for (int idx = tid; idx < d * d; idx += blockDim.x) V[idx] = Vfull[idx];
    __syncthreads();
    for (int idx = tid; idx < d * d; idx += blockDim.x) {
      const int c = idx / d, r = idx - c * d;
      const double* vcol = V + (size_t)c * d;
      double acc = 0.0;
      // C is symmetric: C[g][r] instead of C[r][g] makes the lanes of a warp read consecutive addresses
      for (int g = 0; g < d; ++g) acc = fma(C[(size_t)g * d + r], vcol[g], acc);
      W[idx] = acc;
    }
  } else {
    for (int idx = tid; idx < d * d; idx += blockDim.x) {
      int c = idx / d, r = idx - c * d;
      W[idx] = C[idx];  // column c of C = row c (C is symmetric)
      V[idx] = r == c ? 1.0 : 0.0;
    }
  }
  const int m = (d + 1) & ~1;  // players in the tournament (one bye if d is odd)
  int sweeps = 0;
  for (; sweeps < 40; ++sweeps) {
    __syncthreads();
    if (tid == 0) rotated = 0;
    __syncthreads();
    for (int step = 0; step < m - 1; ++step) {
      for (int pr = warp; pr < m / 2; pr += nwarp) {
        int p = pr == 0 ? m - 1 : (step + pr) % (m - 1);
        int q = (step + m - 1 - pr) % (m - 1);
        if (p >= d || q >= d) continue;
        if (p > q) { int tmp = p; p = q; q = tmp; }
        double* wp = W + (size_t)p * d;
        double* wq = W + (size_t)q * d;
        double al = 0.0, be = 0.0, ga = 0.0;
        for (int r = lane; r < d; r += 32) {
          double a = wp[r], b = wq[r];
          al = fma(a, a, al);
          be = fma(b, b, be);
          ga = fma(a, b, ga);
        }
        al = warp_sum(al); be = warp_sum(be); ga = warp_sum(ga);
        if (al == 0.0 || be == 0.0) continue;
        if (ga * ga <= 1e-22 * (al * be)) continue;  // |cos| <= 1e-11: columns orthogonal to working accuracy
        // The rotation only has to be orthogonal to fp64 accuracy (cs^2 + sn^2 = 1); its angle merely sets the
        // convergence rate, so tan(theta) is computed in fp32 (no fp64 divide / sqrt on the critical path).
        const float zeta = __fdividef((float)(be - al), (float)(2.0 * ga));
        const float tf = __fdividef(copysignf(1.f, zeta), fabsf(zeta) + sqrtf(fmaf(zeta, zeta, 1.f)));
        if (!(fabsf(tf) > 0.f)) continue;  // angle below fp32 resolution (or zeta overflowed): nothing to do
        const double tt = (double)tf;
        const double hh = fma(tt, tt, 1.0);
        double cs = (double)rsqrtf((float)hh);            // fp32 seed, two Newton steps -> full fp64 accuracy
        cs = cs * fma(-0.5 * hh, cs * cs, 1.5);
        cs = cs * fma(-0.5 * hh, cs * cs, 1.5);
        const double sn = cs * tt;
        double* vp = V + (size_t)p * d;
        double* vq = V + (size_t)q * d;
        for (int r = lane; r < d; r += 32) {
          double a = wp[r], b = wq[r];
          wp[r] = cs * a - sn * b;
          wq[r] = sn * a + cs * b;
          a = vp[r]; b = vq[r];
          vp[r] = cs * a - sn * b;
          vq[r] = sn * a + cs * b;
        }
        // a sweep whose largest |cos| was below kQuadCos leaves residuals of about cos^2 (quadratic convergence): no
        // confirming sweep is needed after it
        if (lane == 0 && ga * ga > kQuadCos2 * (al * be)) rotated = 1;
      }
      __syncthreads();
    }
    if (!rotated) break;
  }
  __syncthreads();
  // |lambda_j| = ||W_j||; rank the columns and emit the k largest (descending), zero-padded to KP columns
  for (int c = warp; c < d; c += nwarp) {
    double a = 0.0;
    for (int r = lane; r < d; r += 32) a = fma(W[(size_t)c * d + r], W[(size_t)c * d + r], a);
    a = warp_sum(a);
    if (lane == 0) norms[c] = sqrt(a);
  }
  __syncthreads();
  for (int idx = tid; idx < d * KP; idx += blockDim.x) {
    int f = idx / KP, c = idx - f * KP;
    if (c >= k) { Vd[idx] = 0.0; Vf[idx] = 0.f; }
  }
  for (int c = tid; c < d; c += blockDim.x) {
    int rank = 0;
    const double nc = norms[c];
    for (int o = 0; o < d; ++o) rank += (norms[o] > nc) || (norms[o] == nc && o < c);
    if (rank < k) {
      evals[rank] = nc;
      for (int f = 0; f < d; ++f) {
        double v = V[(size_t)c * d + f];
        Vd[(size_t)f * KP + rank] = v;
        Vf[(size_t)f * KP + rank] = (float)v;
      }
    }
  }
  for (int idx = tid; idx < d * d; idx += blockDim.x) Vfull[idx] = V[idx];
  if (tid == 0) gate[3] = sweeps;
}


// d = 64: the columns live in REGISTERS.  512 threads = 32 half-warps; half-warp h holds the two columns of its pair
// (lane l: rows 4l..4l+3 of W and V of both columns = 16 doubles).  Odd-even transposition ordering with an
// unconditional exchange after every rotation: step t pairs the positions (2h, 2h+1) for even t and (2h+1, 2h+2) for
// odd t; after 64 steps the order of the columns is reversed and every pair has met exactly once (one sweep).  Between
// two steps every half-warp keeps one column and passes the other to a neighbour through shared memory (1 KB written,
// 1 KB read per half-warp and step, one __syncthreads), against four columns read and four written per rotation when
// the matrices sit in shared memory; the three dot products take 4 shuffle rounds and serve two pairs per warp.
__global__ void __launch_bounds__(512, 1)
jacobi64_regs(const double* __restrict__ C, int k, int KP, double* __restrict__ Vd, float* __restrict__ Vf,
              double* __restrict__ evals, long long* __restrict__ gate, double* __restrict__ Vfull, int warm) {
  constexpr int D = 64, CS = 2 * D;  // a column slot: W rows 0..63, then V rows 0..63
  extern __shared__ double S[];      // [64 slots][128]; slots 0..31 = bufE, 32..63 = bufO while the sweeps run
  __shared__ int rotated[2];
  __shared__ double norms[D];
  const int tid = threadIdx.x, lane = tid & 31, hw = tid >> 4, l = tid & 15;
  double* bufE = S;
  double* bufO = S + 32 * CS;

  for (int idx = tid; idx < D * D; idx += blockDim.x) {
    const int c = idx >> 6, r = idx & 63;
    S[c * CS + D + r] = warm ? Vfull[idx] : (r == c ? 1.0 : 0.0);
  }
  if (tid < 2) rotated[tid] = 0;
  __syncthreads();
  // W = C V (columns): C is staged in shared memory behind S; thread (r, cg) forms rows r of the eight columns
  // 8 cg .. 8 cg + 7, so that C[g][r] (C is symmetric) is read once per g and conflict free
  {
    double* Cs = S + D * CS;
    for (int idx = tid; idx < D * D; idx += blockDim.x) Cs[idx] = C[idx];
    __syncthreads();
    const int r = tid & 63, cg = tid >> 6;
    double acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.0;
    if (warm) {
      for (int g = 0; g < D; ++g) {
        const double cv = Cs[g * D + r];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fma(cv, S[(8 * cg + j) * CS + D + g], acc[j]);
      }
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = Cs[(8 * cg + j) * D + r];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) S[(8 * cg + j) * CS + r] = acc[j];
  }
  __syncthreads();
  double Xw[4], Xv[4], Yw[4], Yv[4];
  auto load_col = [&](const double* slot, double (&w)[4], double (&v)[4]) {
    const double2 a = *reinterpret_cast<const double2*>(slot + 4 * l), b = *reinterpret_cast<const double2*>(slot + 4 * l + 2);
    const double2 c = *reinterpret_cast<const double2*>(slot + D + 4 * l), e = *reinterpret_cast<const double2*>(slot + D + 4 * l + 2);
    w[0] = a.x; w[1] = a.y; w[2] = b.x; w[3] = b.y;
    v[0] = c.x; v[1] = c.y; v[2] = e.x; v[3] = e.y;
  };
  auto store_col = [&](double* slot, const double (&w)[4], const double (&v)[4]) {
    *reinterpret_cast<double2*>(slot + 4 * l) = make_double2(w[0], w[1]);
    *reinterpret_cast<double2*>(slot + 4 * l + 2) = make_double2(w[2], w[3]);
    *reinterpret_cast<double2*>(slot + D + 4 * l) = make_double2(v[0], v[1]);
    *reinterpret_cast<double2*>(slot + D + 4 * l + 2) = make_double2(v[2], v[3]);
  };
  load_col(S + (2 * hw) * CS, Xw, Xv);
  load_col(S + (2 * hw + 1) * CS, Yw, Yv);
  __syncthreads();  // from here on S is the exchange buffer

  // rotate the pair (X, Y) so that the W columns become orthogonal; returns whether a rotation was applied
  auto rotate = [&](bool active) {
    double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      al = fma(Xw[i], Xw[i], al);
      be = fma(Yw[i], Yw[i], be);
      ga = fma(Xw[i], Yw[i], ga);
    }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) {
      al += __shfl_xor_sync(0xffffffffu, al, o);
      be += __shfl_xor_sync(0xffffffffu, be, o);
      ga += __shfl_xor_sync(0xffffffffu, ga, o);
    }
    if (!active || al == 0.0 || be == 0.0) return false;
    if (ga * ga <= 1e-22 * (al * be)) return false;  // |cos| <= 1e-11: orthogonal to working accuracy
    const bool big = ga * ga > kQuadCos2 * (al * be);
    // the rotation only has to be orthogonal to fp64 accuracy; its angle merely sets the convergence rate (fp32)
    const float zeta = __fdividef((float)(be - al), (float)(2.0 * ga));
    const float tf = __fdividef(copysignf(1.f, zeta), fabsf(zeta) + sqrtf(fmaf(zeta, zeta, 1.f)));
    if (!(fabsf(tf) > 0.f)) return false;
    const double tt = (double)tf;
    const double hh = fma(tt, tt, 1.0);
    double cs = (double)rsqrtf((float)hh);  // fp32 seed, two Newton steps -> full fp64 accuracy
    cs = cs * fma(-0.5 * hh, cs * cs, 1.5);
    cs = cs * fma(-0.5 * hh, cs * cs, 1.5);
    const double sn = cs * tt;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      double a = Xw[i], b = Yw[i];
      Xw[i] = cs * a - sn * b;
      Yw[i] = sn * a + cs * b;
      a = Xv[i]; b = Yv[i];
      Xv[i] = cs * a - sn * b;
      Yv[i] = sn * a + cs * b;
    }
    return big;  // small rotations do not call for another sweep
  };

  int sweeps = 0;
  for (; sweeps < 40; ++sweeps) {
    bool any = false;
    for (int t = 0; t < D; t += 2) {
      // even step: positions (2h, 2h+1) = (X, Y); afterwards Y moves down to the pair (2h-1, 2h) of half-warp h-1
      any |= rotate(true);
      store_col(bufE + hw * CS, Yw, Yv);
      __syncthreads();
      if (t == 0 && tid == 0) rotated[(sweeps + 1) & 1] = 0;
      if (hw < 31) load_col(bufE + (hw + 1) * CS, Yw, Yv);
      // odd step: positions (2h+1, 2h+2) = (X, Y) for h <= 30; half-warp 31 only keeps position 63 (its X)
      any |= rotate(hw < 31);
      if (hw < 31) store_col(bufO + (hw + 1) * CS, Xw, Xv);
      __syncthreads();
      // next even pair: first = what half-warp h-1 passed up (position 0 is the column parked in bufE[0]),
      // second = this half-warp's Y (half-warp 31: the X it kept)
      if (hw == 31) {
#pragma unroll
        for (int i = 0; i < 4; ++i) { Yw[i] = Xw[i]; Yv[i] = Xv[i]; }
      }
      load_col(hw == 0 ? bufE : bufO + hw * CS, Xw, Xv);
    }
    if (any) rotated[sweeps & 1] = 1;
    __syncthreads();
    if (!rotated[sweeps & 1]) break;
  }
  __syncthreads();
  store_col(S + (2 * hw) * CS, Xw, Xv);
  store_col(S + (2 * hw + 1) * CS, Yw, Yv);
  __syncthreads();
  // |lambda_j| = ||W_j||; rank the columns and emit the k largest (descending), zero-padded to KP columns
  {
    const int warp = tid >> 5, nwarp = blockDim.x >> 5;
    for (int c = warp; c < D; c += nwarp) {
      double a = 0.0;
      for (int r = lane; r < D; r += 32) a = fma(S[c * CS + r], S[c * CS + r], a);
      a = warp_sum(a);
      if (lane == 0) norms[c] = sqrt(a);
    }
  }
  __syncthreads();
  for (int idx = tid; idx < D * KP; idx += blockDim.x) {
    const int c = idx % KP;
    if (c >= k) { Vd[idx] = 0.0; Vf[idx] = 0.f; }
  }
  for (int c = tid; c < D; c += blockDim.x) {
    int rank = 0;
    const double nc = norms[c];
    for (int o = 0; o < D; ++o) rank += (norms[o] > nc) || (norms[o] == nc && o < c);
    if (rank < k) {
      evals[rank] = nc;
      for (int f = 0; f < D; ++f) {
        const double v = S[c * CS + D + f];
        Vd[(size_t)f * KP + rank] = v;
        Vf[(size_t)f * KP + rank] = (float)v;
      }
    }
  }
  for (int idx = tid; idx < D * D; idx += blockDim.x) Vfull[idx] = S[(idx >> 6) * CS + D + (idx & 63)];
  if (tid == 0) gate[3] = sweeps;
}


// d too large for one CTA's shared memory (d > 111): the same rotation scheme spread over several CTAs.  W and V live
// in the workspace (1 MB at d = 256, L2 resident); every step the d/2 column pairs are rotated by one warp each
// across the grid, and a cooperative grid barrier separates the steps.  Registers hold the two W and two V columns of
// a pair (d/32 doubles per lane and column).
template <int DMAX>
__global__ void __launch_bounds__(256, 1)
jacobi_topk_grid(const double* __restrict__ C, int d, int k, int KP, double* __restrict__ W, double* __restrict__ V,
                 double* __restrict__ Vd, float* __restrict__ Vf, double* __restrict__ evals,
                 long long* __restrict__ gate, double* __restrict__ Vfull, int warm, int* __restrict__ rot_flags) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  constexpr int PER = DMAX / 32;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthr = gridDim.x * blockDim.x;
  const int lane = threadIdx.x & 31;
  const int gwarp = tid >> 5, nwarp = nthr >> 5;
  __shared__ double norms[kMaxD];

  if (warm) {
    for (int idx = tid; idx < d * d; idx += nthr) V[idx] = Vfull[idx];
    grid.sync();
    for (int idx = tid; idx < d * d; idx += nthr) {
      const int c = idx / d, r = idx - c * d;
      const double* vcol = V + (size_t)c * d;
      double acc = 0.0;
      for (int g = 0; g < d; ++g) acc = fma(C[(size_t)g * d + r], __ldcg(vcol + g), acc);  // C symmetric: coalesced
      W[idx] = acc;
    }
  } else {
    for (int idx = tid; idx < d * d; idx += nthr) {
      const int c = idx / d, r = idx - c * d;
      W[idx] = C[idx];
      V[idx] = r == c ? 1.0 : 0.0;
    }
  }
  if (tid < 64) rot_flags[tid] = 0;
  grid.sync();
  const int m = (d + 1) & ~1;
  int sweeps = 0;
  for (; sweeps < 40; ++sweeps) {
    for (int step = 0; step < m - 1; ++step) {
      for (int pr = gwarp; pr < m / 2; pr += nwarp) {
        int p = pr == 0 ? m - 1 : (step + pr) % (m - 1);
        int q = (step + m - 1 - pr) % (m - 1);
        if (p >= d || q >= d) continue;
        if (p > q) { int tmp = p; p = q; q = tmp; }
        double* wp = W + (size_t)p * d;
        double* wq = W + (size_t)q * d;
        double a[PER], b[PER];
        double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
          const int r = lane + 32 * i;
          a[i] = r < d ? __ldcg(wp + r) : 0.0;
          b[i] = r < d ? __ldcg(wq + r) : 0.0;
          al = fma(a[i], a[i], al);
          be = fma(b[i], b[i], be);
          ga = fma(a[i], b[i], ga);
        }
        al = warp_sum(al); be = warp_sum(be); ga = warp_sum(ga);
        if (al == 0.0 || be == 0.0) continue;
        if (ga * ga <= 1e-22 * (al * be)) continue;
        const float zeta = __fdividef((float)(be - al), (float)(2.0 * ga));
        const float tf = __fdividef(copysignf(1.f, zeta), fabsf(zeta) + sqrtf(fmaf(zeta, zeta, 1.f)));
        if (!(fabsf(tf) > 0.f)) continue;
        const double tt = (double)tf;
        const double hh = fma(tt, tt, 1.0);
        double cs = (double)rsqrtf((float)hh);
        cs = cs * fma(-0.5 * hh, cs * cs, 1.5);
        cs = cs * fma(-0.5 * hh, cs * cs, 1.5);
        const double sn = cs * tt;
        double* vp = V + (size_t)p * d;
        double* vq = V + (size_t)q * d;
#pragma unroll
        for (int i = 0; i < PER; ++i) {
          const int r = lane + 32 * i;
          if (r < d) {
            __stcg(wp + r, cs * a[i] - sn * b[i]);
            __stcg(wq + r, sn * a[i] + cs * b[i]);
            const double x = __ldcg(vp + r), y = __ldcg(vq + r);
            __stcg(vp + r, cs * x - sn * y);
            __stcg(vq + r, sn * x + cs * y);
          }
        }
        if (lane == 0 && ga * ga > kQuadCos2 * (al * be)) rot_flags[sweeps] = 1;
      }
      grid.sync();
    }
    if (__ldcg(rot_flags + sweeps) == 0) break;  // every thread reads the same flag after the barrier
  }
  // column norms, ranks and output by block 0 (small)
  if (blockIdx.x == 0) {
    const int warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int c = warp; c < d; c += nw) {
      double acc = 0.0;
      for (int r = lane; r < d; r += 32) { const double w = __ldcg(W + (size_t)c * d + r); acc = fma(w, w, acc); }
      acc = warp_sum(acc);
      if (lane == 0) norms[c] = sqrt(acc);
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < d * KP; idx += blockDim.x) {
      const int c = idx % KP;
      if (c >= k) { Vd[idx] = 0.0; Vf[idx] = 0.f; }
    }
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
      int rank = 0;
      const double nc = norms[c];
      for (int o = 0; o < d; ++o) rank += (norms[o] > nc) || (norms[o] == nc && o < c);
      if (rank < k) {
        evals[rank] = nc;
        for (int f = 0; f < d; ++f) {
          const double v = __ldcg(V + (size_t)c * d + f);
          Vd[(size_t)f * KP + rank] = v;
          Vf[(size_t)f * KP + rank] = (float)v;
        }
      }
    }
    if (threadIdx.x == 0) gate[3] = sweeps;
  }
  for (int idx = tid; idx < d * d; idx += nthr) Vfull[idx] = __ldcg(V + idx);
}

// d = 128 and 256: the register-resident odd-even scheme of jacobi64_regs spread over a cluster of eight CTAs.
// One warp per column pair ("seat"; d/2 seats, d/16 per CTA), lane l holds rows (d/32) l .. of W and V of both columns.
// Between two steps every seat keeps one column and passes the other to the neighbouring seat: the writer stores it
// straight into the READER's shared memory - its own CTA's, or the neighbouring CTA's through distributed shared
// memory - and one cluster barrier per step (hardware, a few hundred ns) replaces the cooperative grid barrier and the
// L2 round trips of jacobi_topk_grid (about 4 us per step).  W = C V, the ranking and the outputs go through the
// workspace copies of W and V as before.
// NC CTAs per cluster (8 is the portable maximum, 16 needs the non-portable opt-in): the step is bound by the fp64 pipe
// and the shared-memory traffic of each SM, so more SMs per problem shorten it.
template <int D, int NC>
__global__ void __launch_bounds__(D * 16 / NC, 1)
jacobi_cluster(const double* __restrict__ C, int k, int KP, double* __restrict__ gW, double* __restrict__ gV,
               double* __restrict__ Vd, float* __restrict__ Vf, double* __restrict__ evals,
               long long* __restrict__ gate, double* __restrict__ Vfull, int warm, int* __restrict__ rot_flags,
               const int* __restrict__ skip) {
  if (skip != nullptr && *skip != 0) return;  // the subspace iteration already delivered this batch's vectors
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  constexpr int SEATS = D / 2, SPC = SEATS / NC, PER = D / 32, CS = 2 * D;  // slot: W rows, then V rows
  extern __shared__ double sm[];
  double* bufE = sm;                    // [SPC + 1][CS]: slot s is read by local seat s; slot SPC parks position 0
  double* bufO = sm + (SPC + 1) * CS;   // [SPC][CS]
  __shared__ double norms[D];
  const int cta = (int)cluster.block_rank();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int seat = cta * SPC + warp;
  const int gtid = cta * blockDim.x + tid, nthr = NC * blockDim.x;

  for (int idx = gtid; idx < D * D; idx += nthr) {
    const int c = idx / D, r = idx - c * D;
    gV[idx] = warm ? Vfull[idx] : (r == c ? 1.0 : 0.0);
  }
  if (gtid < 64) rot_flags[gtid] = 0;
  __threadfence();
  cluster.sync();
  for (int idx = gtid; idx < D * D; idx += nthr) {
    const int c = idx / D, r = idx - c * D;
    double acc;
    if (warm) {
      const double* vcol = gV + (size_t)c * D;
      acc = 0.0;
      for (int g = 0; g < D; ++g) acc = fma(C[(size_t)g * D + r], __ldcg(vcol + g), acc);  // C symmetric: coalesced
    } else {
      acc = C[idx];
    }
    gW[idx] = acc;
  }
  __threadfence();
  cluster.sync();

  double Xw[PER], Xv[PER], Yw[PER], Yv[PER];
  auto load_col = [&](const double* w, const double* v, double (&cw)[PER], double (&cv)[PER]) {
#pragma unroll
    for (int i = 0; i < PER; i += 2) {
      const double2 a = *reinterpret_cast<const double2*>(w + PER * lane + i);
      const double2 b = *reinterpret_cast<const double2*>(v + PER * lane + i);
      cw[i] = a.x; cw[i + 1] = a.y;
      cv[i] = b.x; cv[i + 1] = b.y;
    }
  };
  auto store_col = [&](double* w, double* v, const double (&cw)[PER], const double (&cv)[PER]) {
#pragma unroll
    for (int i = 0; i < PER; i += 2) {
      *reinterpret_cast<double2*>(w + PER * lane + i) = make_double2(cw[i], cw[i + 1]);
      *reinterpret_cast<double2*>(v + PER * lane + i) = make_double2(cv[i], cv[i + 1]);
    }
  };
  load_col(gW + (size_t)(2 * seat) * D, gV + (size_t)(2 * seat) * D, Xw, Xv);
  load_col(gW + (size_t)(2 * seat + 1) * D, gV + (size_t)(2 * seat + 1) * D, Yw, Yv);
  // where this seat's outgoing columns go: the slot of the seat that reads them, in that seat's CTA
  double* outE = seat == 0 ? bufE + SPC * CS
                           : cluster.map_shared_rank(bufE, (seat - 1) / SPC) + ((seat - 1) % SPC) * CS;
  double* outO = seat == SEATS - 1 ? nullptr
                                   : cluster.map_shared_rank(bufO, (seat + 1) / SPC) + ((seat + 1) % SPC) * CS;
  const double* inE = bufE + warp * CS;
  const double* inO = seat == 0 ? bufE + SPC * CS : bufO + warp * CS;
  cluster.sync();  // every CTA of the cluster is running before remote shared memory is touched

  auto rotate = [&](bool active) {
    double al = 0.0, be = 0.0, ga = 0.0;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      al = fma(Xw[i], Xw[i], al);
      be = fma(Yw[i], Yw[i], be);
      ga = fma(Xw[i], Yw[i], ga);
    }
    al = warp_sum(al); be = warp_sum(be); ga = warp_sum(ga);
    if (!active || al == 0.0 || be == 0.0) return false;
    if (ga * ga <= 1e-22 * (al * be)) return false;  // |cos| <= 1e-11: orthogonal to working accuracy
    const bool big = ga * ga > kQuadCos2 * (al * be);
    const float zeta = __fdividef((float)(be - al), (float)(2.0 * ga));
    const float tf = __fdividef(copysignf(1.f, zeta), fabsf(zeta) + sqrtf(fmaf(zeta, zeta, 1.f)));
    if (!(fabsf(tf) > 0.f)) return false;
    const double tt = (double)tf;
    const double hh = fma(tt, tt, 1.0);
    double cs = (double)rsqrtf((float)hh);  // fp32 seed, two Newton steps -> full fp64 accuracy
    cs = cs * fma(-0.5 * hh, cs * cs, 1.5);
    cs = cs * fma(-0.5 * hh, cs * cs, 1.5);
    const double sn = cs * tt;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      double a = Xw[i], b = Yw[i];
      Xw[i] = cs * a - sn * b;
      Yw[i] = sn * a + cs * b;
      a = Xv[i]; b = Yv[i];
      Xv[i] = cs * a - sn * b;
      Yv[i] = sn * a + cs * b;
    }
    return big;
  };

  int sweeps = 0;
  for (; sweeps < 40; ++sweeps) {
    bool any = false;
    for (int t = 0; t < D; t += 2) {
      // even step: positions (2h, 2h+1) = (X, Y); afterwards Y moves down to the pair (2h-1, 2h) of seat h-1
      any |= rotate(true);
      store_col(outE, outE + D, Yw, Yv);
      cluster.sync();
      if (seat < SEATS - 1) load_col(inE, inE + D, Yw, Yv);
      // odd step: positions (2h+1, 2h+2) = (X, Y) for h < SEATS-1; the last seat only keeps its X
      any |= rotate(seat < SEATS - 1);
      if (seat < SEATS - 1) store_col(outO, outO + D, Xw, Xv);
      cluster.sync();
      if (seat == SEATS - 1) {
#pragma unroll
        for (int i = 0; i < PER; ++i) { Yw[i] = Xw[i]; Yv[i] = Xv[i]; }
      }
      load_col(inO, inO + D, Xw, Xv);
    }
    if (any && lane == 0) rot_flags[sweeps] = 1;
    __threadfence();
    cluster.sync();
    if (__ldcg(rot_flags + sweeps) == 0) break;
  }
  store_col(gW + (size_t)(2 * seat) * D, gV + (size_t)(2 * seat) * D, Xw, Xv);
  store_col(gW + (size_t)(2 * seat + 1) * D, gV + (size_t)(2 * seat + 1) * D, Yw, Yv);
  __threadfence();
  cluster.sync();
  // column norms, ranks and output by CTA 0 (small); the full basis for the next warm start by everyone
  for (int idx = gtid; idx < D * D; idx += nthr) Vfull[idx] = __ldcg(gV + idx);
  if (cta == 0) {
    const int nw = blockDim.x >> 5;
    for (int c = warp; c < D; c += nw) {
      double acc = 0.0;
      for (int r = lane; r < D; r += 32) { const double w = __ldcg(gW + (size_t)c * D + r); acc = fma(w, w, acc); }
      acc = warp_sum(acc);
      if (lane == 0) norms[c] = sqrt(acc);
    }
    __syncthreads();
    for (int idx = tid; idx < D * KP; idx += blockDim.x) {
      const int c = idx % KP;
      if (c >= k) { Vd[idx] = 0.0; Vf[idx] = 0.f; }
    }
    for (int c = tid; c < D; c += blockDim.x) {
      int rank = 0;
      const double nc = norms[c];
      for (int o = 0; o < D; ++o) rank += (norms[o] > nc) || (norms[o] == nc && o < c);
      if (rank < k) {
        evals[rank] = nc;
        for (int f = 0; f < D; ++f) {
          const double v = __ldcg(gV + (size_t)c * D + f);
          Vd[(size_t)f * KP + rank] = v;
          Vf[(size_t)f * KP + rank] = (float)v;
        }
      }
    }
    if (tid == 0) gate[3] = sweeps;
  }
}

}  // namespace

int launch_eigh(sb_maha* h, const MahaWs& ws, cudaStream_t st) {
  const int d = h->d;
  const size_t need = 2 * (size_t)d * d * sizeof(double);
  const int warm = (h->warm_batches % 256) != 0;  // periodic cold start keeps the accumulated basis orthonormal
  h->warm_batches++;
  if (h->profile) cudaEventRecord(h->ev[2], st);
  if (d == 64 && !h->force_generic) {
    const size_t sm64 = (size_t)(64 * 128 + 64 * 64) * sizeof(double);  // columns (W | V) + the staged covariance
    SB_CUDA(sb::func_attr_cached(jacobi64_regs, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm64));
    jacobi64_regs<<<1, 512, sm64, st>>>(h->d_C1, h->k, h->KP, h->d_V, h->d_Vf, h->d_evals, h->d_gate, h->d_Vfull, warm);
  } else if ((d == 128 || d == 256) && !h->force_generic) {
    // batches after the first: warm-started block subspace iteration on one CTA; the Jacobi solver below runs only
    // when that reports no proven convergence (its `skip` word stays 0)
    int* conv = reinterpret_cast<int*>(h->d_gate + 47);
    const bool try_sub = h->d_Vsub != nullptr && !h->no_subspace && subspace_supported(d, h->KP);
    const int* skip = nullptr;
    if (try_sub) {
      if (h->sub_state == 0) {
        int rc = launch_subspace_init(h, st);
        if (rc) return rc;
        SB_CUDA(cudaMemsetAsync(h->d_gate + 46, 0, sizeof(long long), st));  // "block is orthonormal" word
        h->sub_state = 1;
      }
      SB_CUDA(cudaMemsetAsync(conv, 0, sizeof(int), st));
      SB_CUDA(cudaMemsetAsync(h->d_gate + 7, 0, sizeof(long long), st));
      if (h->sub_state >= 2) {
        int rc = launch_subspace(h, conv, st);
        if (rc) return rc;
        skip = conv;
      }
    }
    // cluster of 8 CTAs, columns in registers, neighbours exchange through (distributed) shared memory
    int* flags = reinterpret_cast<int*>(h->d_gate + 8);
    static const int nc = getenv("SB200_JACOBI_CLUSTER") ? atoi(getenv("SB200_JACOBI_CLUSTER")) : 16;
    auto launch = [&](auto kernel, int ncta) -> cudaError_t {
      const int spc = (d / 2) / ncta;
      const size_t smc = (size_t)(2 * spc + 1) * 2 * d * sizeof(double);
      cudaError_t e = sb::func_attr_cached(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smc);
      if (e == cudaSuccess && ncta > 8) e = sb::func_attr_cached(kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
      if (e != cudaSuccess) return e;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(ncta);
      cfg.blockDim = dim3(spc * 32);
      cfg.dynamicSmemBytes = smc;
      cfg.stream = st;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = ncta;
      attr[0].val.clusterDim.y = 1;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      return cudaLaunchKernelEx(&cfg, kernel, (const double*)h->d_C1, h->k, h->KP, ws.eig_W, ws.eig_V, h->d_V, h->d_Vf,
                                h->d_evals, h->d_gate, h->d_Vfull, warm, flags, skip);
    };
    // 16 CTAs per cluster where the device schedules it (a full GPU does), else the portable 8
    cudaError_t ce = cudaErrorInvalidValue;
    if (nc == 16 && !h->no_cluster16) {
      ce = d == 128 ? launch(jacobi_cluster<128, 16>, 16) : launch(jacobi_cluster<256, 16>, 16);
      if (ce != cudaSuccess) {
        (void)cudaGetLastError();
        h->no_cluster16 = 1;
      }
    }
    if (ce != cudaSuccess) ce = d == 128 ? launch(jacobi_cluster<128, 8>, 8) : launch(jacobi_cluster<256, 8>, 8);
    SB_CUDA(ce);
    if (try_sub) {
      int rc = launch_subspace_seed(h, skip, st);  // a Jacobi result becomes the head of the block
      if (rc) return rc;
      h->sub_state = 2;
    }
  } else if (need <= 200 * 1024) {
    SB_CUDA(sb::func_attr_cached(jacobi_topk<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    jacobi_topk<true><<<1, 1024, need, st>>>(h->d_C1, d, h->k, h->KP, nullptr, nullptr, h->d_V, h->d_Vf, h->d_evals,
                                            h->d_gate, h->d_Vfull, warm);
  } else {
    // cooperative multi-CTA solver: one warp per column pair
    const int pairs = (d + 1) / 2;
    int ctas = (int)ceil_div(pairs, 8);
    if (ctas > kNumSMs) ctas = kNumSMs;
    const double* Cp = h->d_C1;
    int dd = d, kk = h->k, kp = h->KP, wm = warm;
    double *Wp = ws.eig_W, *Vp = ws.eig_V, *Vdp = h->d_V, *evp = h->d_evals, *Vfl = h->d_Vfull;
    float* Vfp = h->d_Vf;
    long long* gp = h->d_gate;
    int* flags = reinterpret_cast<int*>(h->d_gate + 8);
    void* args[] = {&Cp, &dd, &kk, &kp, &Wp, &Vp, &Vdp, &Vfp, &evp, &gp, &Vfl, &wm, &flags};
    if (d <= 256) SB_CUDA(cudaLaunchCooperativeKernel((void*)jacobi_topk_grid<256>, dim3(ctas), dim3(256), args, 0, st));
    else SB_CUDA(cudaLaunchCooperativeKernel((void*)jacobi_topk_grid<512>, dim3(ctas), dim3(256), args, 0, st));
  }
  if (h->profile) cudaEventRecord(h->ev[3], st);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

}  // namespace sb
