// Top-k eigenvectors of the updated covariance by warm-started block subspace iteration (d = 128 / 256).
//
// Replaces the call `eigh(cov_batch, eigvals=(p-k, p-1))` of CoreMahalanobis.py:103 on the batches after the first:
// only the span of the k leading eigenvectors is used downstream (the score is invariant to the basis),
// and from one request batch to the next the covariance moves little, so the block of the previous batch is an
// excellent start.  A full Jacobi solve of a 256 x 256 matrix costs 1.6 ms on a 16-CTA cluster and is replicated on
// every rank of a row-sharded batch; this kernel needs one to six products C V (256 x 256 x 32 in fp64: ~50 us each on
// the fp64 pipe of ONE SM) and one to three 32 x 32 eigenproblems (~80 us each).
//
//   V (d x 32: the k wanted vectors + guard vectors, orthonormal)            kept in the handle between batches
//   repeat:  T = C V;  H = V^T T;  H = Y Theta Y^T (parallel two-sided Jacobi, 32 x 32);  V <- V Y, T <- T Y
//            residuals r_j = |T_j - theta_j V_j|;  accept or  V <- orth(T) (Cholesky QR) and go on
//
// Acceptance is a guarantee, not a heuristic.  With the Ritz values sorted, "top" = the k largest:
//   * every eigenvalue of C outside the block's span is at most sqrt(|C|_F^2 - sum_j theta_j^2) =: rho (the block is
//     invariant up to the residuals), the other Ritz values of the block are at most theta_j + r_j, so
//         gap = min_top theta - max(rho, max_rest(theta_j + r_j))
//     is a lower bound of the distance between the wanted eigenvalues and all others; gap <= 0 means the block cannot
//     be shown to hold the dominant subspace (no spectral gap at k, or a new dominant direction outside the block);
//   * Davis-Kahan: sin of the largest angle between span(top) and the true invariant subspace <= |R_top|_F / gap.
// The kernel accepts when |R_top|_F <= tol * gap (tol = 1e-5: a projector error an order below the 1e-4 parity
// bound) within max_prod products C V (the Rayleigh-Ritz step and the test run after products 1, 4, 6, 8), and otherwise reports failure; the caller then runs the Jacobi solver, which needs
// no gap.  Either way the block (with whatever progress was made) is kept for the next batch.
#include "maha.cuh"

namespace sb {
namespace {

constexpr int kB = 32;    // block width: k <= 16 wanted vectors + guard vectors
constexpr int kLD = 34;   // row stride of the d x 32 blocks in shared memory (doubles; even: 16-byte vector accesses)
constexpr int kLS = 33;   // row stride of the 32 x 32 matrices (odd: column walks are conflict free)
constexpr int kCR = 8;    // rows of C per staged chunk
constexpr int kNT = 512;  // threads: 16 warps, up to 128 registers each (the row updates keep 32 doubles per thread)

#ifdef SB_SUBSPACE_TIMING
#define SB_TICK(i) do { const long long now__ = clock64(); tph[i] += now__ - tk; tk = now__; } while (0)
#else
#define SB_TICK(i) do { } while (0)
#endif

template <int D>
__global__ void __launch_bounds__(kNT, 1)
subspace_topk(const double* __restrict__ C, int k, int KP, double* __restrict__ Vsub, double* __restrict__ Vd,
              float* __restrict__ Vf, double* __restrict__ evals, long long* __restrict__ gate,
              int* __restrict__ conv_flag, int* __restrict__ ortho_flag, int max_prod, double tol) {
  constexpr int TPR = kNT / D;    // threads per row (row-wise updates)
  constexpr int CPT = kB / TPR;   // columns per thread there
  extern __shared__ double sm[];
  double* V = sm;                 // [D][kLD]
  double* T = V + D * kLD;        // [D][kLD]
  double* H = T + D * kLD;        // [kB][kLS]
  double* Y = H + kB * kLS;       // [kB][kLS]
  double* L = Y + kB * kLS;       // [kB][kLS] Cholesky factor
  double* th = L + kB * kLS;      // [kB]
  double* r2 = th + kB;           // [kB]
  double* red = r2 + kB;          // [16] block reduction scratch
  double* Cs = red + 32;          // [2][kCR][D] rows of C on their way through (cp.async double buffer)
  __shared__ double rot_c[16], rot_s[16];
  __shared__ int s_rank[kB], s_fail, s_conv, s_any, s_next;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int row = tid / TPR, c0 = (tid % TPR) * CPT;
#ifdef SB_SUBSPACE_TIMING
  long long tph[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long tk = clock64();
#endif

  // |C|_F^2
  double f2 = 0.0;
  for (int idx = tid; idx < D * D; idx += kNT) { const double v = C[idx]; f2 = fma(v, v, f2); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) f2 += __shfl_xor_sync(0xffffffffu, f2, o);
  if (lane == 0) red[warp] = f2;
  for (int idx = tid; idx < D * kB; idx += kNT) V[(idx / kB) * kLD + (idx % kB)] = Vsub[idx];
  if (tid == 0) { s_fail = 0; s_conv = 0; s_next = 1; }
  __syncthreads();
  double normC2 = 0.0;
  for (int w = 0; w < kNT / 32; ++w) normC2 += red[w];
  __syncthreads();

  // V <- orth(V) by Cholesky QR.  Sets s_fail when the block has lost rank.
  auto orth = [&]() {
    for (int e = tid; e < kB * kB; e += kNT) {  // G = V^T V into H
      const int a = e >> 5, b = e & 31;
      double acc = 0.0;
      if (a <= b) {
        double a0 = 0.0, a1 = 0.0;
        for (int i = 0; i < D; i += 2) {
          a0 = fma(V[i * kLD + a], V[i * kLD + b], a0);
          a1 = fma(V[(i + 1) * kLD + a], V[(i + 1) * kLD + b], a1);
        }
        acc = a0 + a1;
        H[a * kLS + b] = acc;
        H[b * kLS + a] = acc;
      }
    }
    __syncthreads();
    if (warp == 0) {
      // Cholesky G = L L^T with lane = row and the row of L in registers; the pivot row comes by shuffle
      double Lr[kB];
      bool bad = false;
#pragma unroll
      for (int j = 0; j < kB; ++j) {
        double s = H[lane * kLS + j];
#pragma unroll
        for (int m = 0; m < j; ++m) s = fma(-Lr[m], __shfl_sync(0xffffffffu, Lr[m], j), s);
        const double djj = __shfl_sync(0xffffffffu, s, j);
        const double gjj = H[j * kLS + j];
        bad = bad || !(djj > 1e-13 * gjj) || !(djj > 1e-280);
        const double ljj = sqrt(bad ? 1.0 : djj);
        Lr[j] = lane == j ? ljj : (lane > j ? s / ljj : 0.0);
      }
#pragma unroll
      for (int j = 0; j < kB; ++j) L[lane * kLS + j] = Lr[j];
      if (lane == 0 && bad) s_fail = 1;
    }
    __syncthreads();
    if (s_fail) return;
    if (tid < D) {  // row tid:  x L^T = v  (forward substitution)
      double x[kB];
#pragma unroll
      for (int j = 0; j < kB; ++j) x[j] = V[tid * kLD + j];
#pragma unroll
      for (int j = 0; j < kB; ++j) {
        double s0 = x[j], s1 = 0.0;
#pragma unroll
        for (int m = 0; m + 1 < j; m += 2) {
          s0 = fma(-x[m], L[j * kLS + m], s0);
          s1 = fma(-x[m + 1], L[j * kLS + m + 1], s1);
        }
        if (j & 1) s0 = fma(-x[j - 1], L[j * kLS + j - 1], s0);
        x[j] = (s0 + s1) / L[j * kLS + j];
      }
#pragma unroll
      for (int j = 0; j < kB; ++j) V[tid * kLD + j] = x[j];
    }
    __syncthreads();
  };

  // the block as stored by the last run is its orthonormal Ritz basis; one seeded from a Jacobi result is not
  const int block_is_orthonormal = *ortho_flag;
  if (!block_is_orthonormal) {
    orth();
    orth();  // twice: a block seeded from two sources may be far from orthonormal
  }
  SB_TICK(0);
  int prod = 0;
  while (!s_fail) {
    ++prod;
    // T = C V: a thread owns a 4 x 4 tile (C symmetric: C[j][4 rows] is one aligned 32-byte read, coalesced over the rows)
    {
      // the rows of C stream through shared memory kCR at a time (16-byte cp.async, the next chunk in flight while
      // this one is multiplied): read straight from L2 the product was bound by load latency, not by the fp64 pipe
      const bool worker = tid < (D / 4) * 8;
      const int rg = tid >> 3, cgp = tid & 7;
      double acc[4][4];
#pragma unroll
      for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
      constexpr int kChunkVec = kCR * D / 2;  // 16-byte pieces per chunk
      auto stage = [&](int chunk, int buf) {
        const char* src = reinterpret_cast<const char*>(C + (size_t)chunk * kCR * D);
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(Cs + buf * kCR * D);
        for (int v = tid; v < kChunkVec; v += kNT)
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16 * v), "l"(src + 16 * (size_t)v) : "memory");
        asm volatile("cp.async.commit_group;" ::: "memory");
      };
      stage(0, 0);
      for (int ch = 0; ch < D / kCR; ++ch) {
        if (ch + 1 < D / kCR) stage(ch + 1, (ch + 1) & 1);
        else asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();
        if (worker) {
          const double* cb = Cs + (ch & 1) * kCR * D + 4 * rg;
#pragma unroll
          for (int jj = 0; jj < kCR; ++jj) {
            const int j = ch * kCR + jj;
            const double2 ca = *reinterpret_cast<const double2*>(cb + jj * D);
            const double2 cc = *reinterpret_cast<const double2*>(cb + jj * D + 2);
            const double2 va = *reinterpret_cast<const double2*>(V + j * kLD + 4 * cgp);
            const double2 vb = *reinterpret_cast<const double2*>(V + j * kLD + 4 * cgp + 2);
            const double cr[4] = {ca.x, ca.y, cc.x, cc.y}, vc[4] = {va.x, va.y, vb.x, vb.y};
#pragma unroll
            for (int a = 0; a < 4; ++a)
#pragma unroll
              for (int b = 0; b < 4; ++b) acc[a][b] = fma(cr[a], vc[b], acc[a][b]);
          }
        }
        __syncthreads();  // the buffer is free for the chunk after next
      }
      if (worker) {
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          double* tp = T + (4 * rg + a) * kLD + 4 * cgp;
          *reinterpret_cast<double2*>(tp) = make_double2(acc[a][0], acc[a][1]);
          *reinterpret_cast<double2*>(tp + 2) = make_double2(acc[a][2], acc[a][3]);
        }
      }
    }
    __syncthreads();
    SB_TICK(1);
    // Rayleigh-Ritz and the acceptance test after product 1 (a covariance that barely moved is accepted at once) and
    // then where the previous test predicts convergence: the span converges with or without it, and the 32 x 32
    // eigenproblem costs more than a product
    const bool rr = prod >= s_next || prod >= max_prod;
    if (rr) {
      for (int e = tid; e < kB * kB; e += kNT) {  // H = sym(V^T T), Y = I
        const int a = e >> 5, b = e & 31;
        Y[a * kLS + b] = a == b ? 1.0 : 0.0;
        if (a <= b) {  // V^T (C V) is symmetric up to rounding: the upper triangle is mirrored
          double x0 = 0.0, x1 = 0.0;
          for (int i = 0; i < D; i += 2) {
            x0 = fma(V[i * kLD + a], T[i * kLD + b], x0);
            x1 = fma(V[(i + 1) * kLD + a], T[(i + 1) * kLD + b], x1);
          }
          const double v = x0 + x1;
          H[a * kLS + b] = v;
          H[b * kLS + a] = v;
        }
      }
      __syncthreads();
      SB_TICK(2);
      // parallel two-sided Jacobi: 16 disjoint pairs per step (round robin), warp w owns pair w.  A sweep in which no
      // off-diagonal entry exceeded 1e-8 of its diagonal pair leaves them at ~1e-16 (quadratic convergence).
      for (int sweep = 0; sweep < 8; ++sweep) {
        if (tid == 0) s_any = 0;
        __syncthreads();
        for (int r = 0; r < kB - 1; ++r) {
          int p, q;
          if (warp == 0) { p = kB - 1; q = r; }
          else { p = (r + warp) % (kB - 1); q = (r - warp + (kB - 1)) % (kB - 1); }
          if (p > q) { const int t = p; p = q; q = t; }
          if (lane == 0) {
            const double app = H[p * kLS + p], aqq = H[q * kLS + q], apq = H[p * kLS + q];
            double c = 1.0, s = 0.0;
            const double scale2 = fabs(app * aqq);
            if (apq * apq > 1e-34 * scale2 && apq != 0.0) {
              // t = sign(tau) / (|tau| + sqrt(1 + tau^2)), tau = (aqq - app) / (2 apq), written with one square root
              // and one division: t = sign(al) apq / (|al| + hypot(al, apq)), al = (aqq - app) / 2
              const double al = 0.5 * (aqq - app);
              const double hyp = sqrt(fma(al, al, apq * apq));
              const double t = (al >= 0.0 ? apq : -apq) / (fabs(al) + hyp);
              c = rsqrt(fma(t, t, 1.0));
              s = t * c;
              if (apq * apq > 1e-16 * scale2) s_any = 1;
            }
            rot_c[warp] = c;
            rot_s[warp] = s;
          }
          __syncthreads();
          {  // columns p, q of H and Y (lane = row)
            const double c = rot_c[warp], s = rot_s[warp];
            const double hp = H[lane * kLS + p], hq = H[lane * kLS + q];
            H[lane * kLS + p] = c * hp - s * hq;
            H[lane * kLS + q] = s * hp + c * hq;
            const double yp = Y[lane * kLS + p], yq = Y[lane * kLS + q];
            Y[lane * kLS + p] = c * yp - s * yq;
            Y[lane * kLS + q] = s * yp + c * yq;
          }
          __syncthreads();
          {  // rows p, q of H (lane = column)
            const double c = rot_c[warp], s = rot_s[warp];
            const double hp = H[p * kLS + lane], hq = H[q * kLS + lane];
            H[p * kLS + lane] = c * hp - s * hq;
            H[q * kLS + lane] = s * hp + c * hq;
          }
          __syncthreads();
        }
        if (!s_any) break;
      }
      SB_TICK(3);
      if (tid < kB) th[tid] = H[tid * kLS + tid];
      // V <- V Y, T <- T Y (one matrix at a time: 32 doubles of the row in registers)
#pragma unroll 1
      for (int which = 0; which < 2; ++which) {
        double* M = which ? T : V;
        double v[kB];
#pragma unroll
        for (int m = 0; m < kB; ++m) v[m] = M[row * kLD + m];
        __syncthreads();
#pragma unroll 4
        for (int c = 0; c < CPT; ++c) {
          double a0 = 0.0, a1 = 0.0;
#pragma unroll
          for (int m = 0; m < kB; m += 2) {
            a0 = fma(v[m], Y[m * kLS + c0 + c], a0);
            a1 = fma(v[m + 1], Y[(m + 1) * kLS + c0 + c], a1);
          }
          M[row * kLD + c0 + c] = a0 + a1;
        }
      }
      __syncthreads();
      SB_TICK(4);
      for (int j = warp; j < kB; j += kNT / 32) {  // residual of Ritz pair j
        const double theta = th[j];
        double acc = 0.0;
        for (int i = lane; i < D; i += 32) {
          const double e = fma(-theta, V[i * kLD + j], T[i * kLD + j]);
          acc = fma(e, e, acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) r2[j] = acc;
      }
      __syncthreads();
      if (warp == 0) {  // rank the Ritz values, bound the gap, decide
        const double tj = th[lane];
        int rk = 0;
        for (int m = 0; m < kB; ++m) rk += (th[m] > tj) || (th[m] == tj && m < lane);
        s_rank[lane] = rk;
        const bool top = rk < k;
        double sumsq = tj * tj, lo_top = top ? tj : 1e300, hi_rest = top ? -1e300 : tj + sqrt(r2[lane]);
        double res2 = top ? r2[lane] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          sumsq += __shfl_xor_sync(0xffffffffu, sumsq, o);
          res2 += __shfl_xor_sync(0xffffffffu, res2, o);
          lo_top = fmin(lo_top, __shfl_xor_sync(0xffffffffu, lo_top, o));
          hi_rest = fmax(hi_rest, __shfl_xor_sync(0xffffffffu, hi_rest, o));
        }
        const double rho = sqrt(fmax(normC2 - sumsq, 0.0));
        const double gap = lo_top - fmax(rho, hi_rest);
        double th_min = tj;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) th_min = fmin(th_min, __shfl_xor_sync(0xffffffffu, th_min, o));
        if (lane == 0) {
          s_conv = gap > 0.0 && sqrt(res2) <= tol * gap;
          // the error of the wanted span shrinks per product by about (smallest Ritz value of the block) / (smallest
          // wanted one): test again where that rate brings the residual under the bound
          int more = 2;
          if (gap > 0.0 && res2 > 0.0 && th_min > 0.0 && th_min < lo_top) {
            const double need = log(tol * gap / sqrt(res2)) / log(fmax(th_min / lo_top, 1e-3));
            more = need < 1.0 ? 1 : (need > 4.0 ? 4 : (int)ceil(need));
          }
          s_next = prod + more;
        }
      }
      __syncthreads();
      SB_TICK(5);
      if (s_conv) break;
    }
    if (prod >= max_prod) break;
    for (int idx = tid; idx < D * kB; idx += kNT) V[(idx / kB) * kLD + (idx % kB)] = T[(idx / kB) * kLD + (idx % kB)];
    __syncthreads();
    orth();
    SB_TICK(6);
  }
#ifdef SB_SUBSPACE_TIMING
  if (tid == 0) for (int i = 0; i < 7; ++i) gate[16 + i] = tph[i];
#endif
  // keep the block for the next batch: Ritz vectors, wanted ones first (the loop only ends right after a Rayleigh-Ritz
  // step, so V is orthonormal and ranked); a block that lost rank is dropped (the old one stays)
  const int it = prod;
  for (int idx = tid; idx < D * kB; idx += kNT) {
    const int i = idx / kB, j = idx % kB;
    if (!s_fail) Vsub[i * kB + s_rank[j]] = V[i * kLD + j];
  }
  if (s_conv) {
    for (int idx = tid; idx < D * kB; idx += kNT) {
      const int i = idx / kB, j = idx % kB, rk = s_rank[j];
      if (rk < k) {
        Vd[(size_t)i * KP + rk] = V[i * kLD + j];
        Vf[(size_t)i * KP + rk] = (float)V[i * kLD + j];
      }
    }
    for (int idx = tid; idx < D * (KP - k); idx += kNT) {
      const int i = idx / (KP - k), j = k + idx % (KP - k);
      Vd[(size_t)i * KP + j] = 0.0;
      Vf[(size_t)i * KP + j] = 0.f;
    }
    if (tid < kB && s_rank[tid] < k) evals[s_rank[tid]] = th[tid];
  }
  if (tid == 0) {
    if (!s_fail) *ortho_flag = 1;
    *conv_flag = s_conv;
    gate[7] = s_conv ? it : -it;  // diagnostics: products C V taken, negative when the Jacobi solver has to run
    if (s_conv) gate[3] = it;
  }
}

// after a Jacobi solve: its k vectors become the head of the block (the guard vectors stay as they are)
__global__ void subspace_seed(const double* __restrict__ Vd, int d, int k, int KP, const int* __restrict__ conv_flag,
                              int* __restrict__ ortho_flag, double* __restrict__ Vsub) {
  if (conv_flag && *conv_flag) return;  // the block is already current
  if (blockIdx.x == 0 && threadIdx.x == 0) *ortho_flag = 0;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < d * k; idx += gridDim.x * blockDim.x) {
    const int i = idx / k, j = idx - i * k;
    Vsub[i * kB + j] = Vd[(size_t)i * KP + j];
  }
}

// first use: guard vectors = fixed pseudo-random directions (the first orthonormalisation makes them a basis)
__global__ void subspace_init(double* __restrict__ Vsub, int d) {
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < d * kB; idx += gridDim.x * blockDim.x) {
    unsigned x = (unsigned)idx * 2654435761u + 12345u;
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    Vsub[idx] = ((double)x / 4294967296.0 - 0.5);
  }
}

}  // namespace

int subspace_block_doubles(int d) { return d * kB; }
bool subspace_supported(int d, int KP) { return (d == 128 || d == 256) && KP <= 16; }

int launch_subspace_init(sb_maha* h, cudaStream_t st) {
  subspace_init<<<32, 256, 0, st>>>(h->d_Vsub, h->d);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

int launch_subspace(sb_maha* h, int* conv_flag, cudaStream_t st) {
  const int d = h->d;
  const size_t smem = ((size_t)2 * d * kLD + 3 * kB * kLS + 2 * kB + 32 + 2 * kCR * d) * sizeof(double);
  int* ortho = reinterpret_cast<int*>(h->d_gate + 46);
  if (d == 256) {
    SB_CUDA(sb::func_attr_cached(subspace_topk<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    subspace_topk<256><<<1, kNT, smem, st>>>(h->d_C1, h->k, h->KP, h->d_Vsub, h->d_V, h->d_Vf, h->d_evals, h->d_gate,
                                             conv_flag, ortho, 8, 1e-5);
  } else {
    SB_CUDA(sb::func_attr_cached(subspace_topk<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    subspace_topk<128><<<1, kNT, smem, st>>>(h->d_C1, h->k, h->KP, h->d_Vsub, h->d_V, h->d_Vf, h->d_evals, h->d_gate,
                                             conv_flag, ortho, 8, 1e-5);
  }
  SB_LAUNCH_CHECK();
  return SB_OK;
}

int launch_subspace_seed(sb_maha* h, const int* conv_flag, cudaStream_t st) {
  subspace_seed<<<16, 256, 0, st>>>(h->d_V, h->d, h->k, h->KP, conv_flag, reinterpret_cast<int*>(h->d_gate + 46), h->d_Vsub);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

}  // namespace sb
