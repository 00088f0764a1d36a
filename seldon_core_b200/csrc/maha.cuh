// Internal declarations shared by the Mahalanobis translation units.
#pragma once
#include "common.cuh"

struct sb_maha;

namespace sb {

// A request batch (the rows this rank holds) is cut into contiguous sub-segments.  Pass 1 emits one
// (row-sum, Gram) partial per sub-segment; after the eigenvectors are known their projections are
// prefix-summed so that pass 2 can start an independent inverse chain at every sub-segment
// (SURVEY App. A.4).  The first `nshort` sub-segments are short: a cold start spends them in the exact
// fp64 warm-up kernel.
struct SubLayout {
  int64_t rows;
  int64_t sublen;
  int32_t nshort;
  int32_t shortlen;
  int32_t nsub;
};

__host__ __device__ inline int64_t sub_begin(const SubLayout& L, int s) {
  int64_t b = s < L.nshort ? (int64_t)s * L.shortlen : (int64_t)L.nshort * L.shortlen + (int64_t)(s - L.nshort) * L.sublen;
  return b < L.rows ? b : L.rows;
}

constexpr int kTargetChains = kNumSMs * 16;  // independent inverse chains wanted in flight
constexpr int kShortLen = 128;
constexpr int kMinSubLen = 64;
constexpr int kMaxSubLen = 4096;  // bounds the fp32 drift of a Sherman-Morrison chain
constexpr int kMaxD = 512;
constexpr int kMaxKP = 32;

SubLayout make_layout(int64_t rows);
int64_t max_subsegments(int64_t max_rows);
// Split pass 2 takes the per-sub-segment statistics from its own projection kernel, not from the pass-1 partials, so it
// cuts the rows independently and finer: chain4 runs 8 chains per warp and wants kTargetChains2 of them in flight.
constexpr int kTargetChains2 = kNumSMs * 64;  // 2 CTAs x 4 warps x 8 chains per SM
constexpr int kMinSubLen2 = 256;
SubLayout make_layout2(int64_t rows);
SubLayout make_layout1(int64_t rows);
int64_t max_subsegments2(int64_t max_rows);

constexpr int SUBFLAG_OPEN = 1;  // warm-up gate test passes at the first row of the sub-segment
constexpr int SUBFLAG_WELL = 2;  // start matrix well enough conditioned for the fp32 chain

struct MahaWs {  // device pointers carved out of the caller's workspace
  float* partials;     // [nsub][d + d*d]
  double* subproj;     // [nsub + 1][KP + KP*KP]   (+1: totals after the scan)
  float* sub_P;        // [nsub][KP*KP]  inverse of S at the sub-segment start
  float* sub_Z;        // [nsub][KP]     prefix of projected deviations at the start
  int32_t* sub_flags;  // [nsub]
  double* stats;       // [1 + d + d*d]  this rank's totals
  double* rank_pre;    // [1 + d + d*d]  totals of the ranks before this one
  double* rank_proj;   // [KP + KP*KP]   their projection (offset of every prefix on this rank)
  double* slices;      // [reduce_slices(d)][d + d*d] stage-1 sums of the partials
  double* chunk_tot;   // [ceil(nsub/32)][KP + KP*KP] per-chunk totals of the projected partials
  float* proj_Z;       // [rows][16] clipped projections z (split pass 2, d = 64)
  float* proj_D;       // [rows][16] y - z, written only for rows with a clipped feature
  uint8_t* proj_flag;  // [rows] 1 = row has a clipped feature
  double* eig_W;       // [d*d] Jacobi work
  double* eig_V;       // [d*d]
  uint64_t bytes;
};

int reduce_slices(int d);
int launch_pass1_umma(const sb_maha* h, const float* X, int64_t ld, const MahaWs& ws, cudaStream_t st);
int launch_pass1(const sb_maha* h, const float* X, int64_t ld, const MahaWs& ws, cudaStream_t st);
int launch_reduce(const sb_maha* h, const MahaWs& ws, double* stats_out, cudaStream_t st);
int launch_finalize_cov(sb_maha* h, const double* all_stats, int nranks, int rank, const MahaWs& ws, cudaStream_t st);
int launch_eigh(sb_maha* h, const MahaWs& ws, cudaStream_t st);
// warm-started block subspace iteration for the top-k eigenvectors (maha_subspace.cu)
int subspace_block_doubles(int d);
bool subspace_supported(int d, int KP);
int launch_subspace_init(sb_maha* h, cudaStream_t st);
int launch_subspace(sb_maha* h, int* conv_flag, cudaStream_t st);
int launch_subspace_seed(sb_maha* h, const int* conv_flag, cudaStream_t st);
int launch_prefix(sb_maha* h, const MahaWs& ws, double* rank_proj, cudaStream_t st);
int launch_score(sb_maha* h, const float* X, int64_t ld, float* score, int32_t* flag, const MahaWs& ws,
                 cudaStream_t st);

}  // namespace sb

struct sb_maha {
  int d, k, KP;
  double threshold, n_stdev;
  int64_t start_clip, max_n;
  // host-side mirror of the scalar state
  int64_t n;
  int has_state, has_clip;
  // batch in flight
  int stage;  // 0 idle, 1 pass1 done, 2 finalized, 3 scored
  int64_t rows_local, rows_total, row0, n_eff;
  int use_clip;
  int split_pass2;  // decided in pass 1: projection kernel + chain kernel (d = 64/128/256, k <= 16, aligned rows)
  sb::SubLayout lay;   // pass 1 (and the non-split pass 2)
  sb::SubLayout lay2;  // split pass 2
  // options
  int gram_terms, force_generic, profile, fused_pass2, legacy_mma;
  cudaEvent_t ev[10];  // profile: [0,1] pass-1 kernel, [2,3] eigh, [4,5] prefix kernels, [6,7] projection, [8,9] chain
  // device state
  double *d_mean, *d_C, *d_lo, *d_hi;      // committed
  double *d_mean1, *d_C1, *d_lo1, *d_hi1;  // pending (this batch)
  float *d_mean_f, *d_lo_f, *d_hi_f;       // fp32 copies of the committed state
  double *d_V, *d_PC0;                     // [d][KP], [KP][KP]
  double* d_Vfull;                         // [d][d] full Jacobi basis of the previous batch (warm start)
  double* d_Vsub;                          // [d][32] block of the subspace iteration (d = 128 / 256), kept between batches
  int sub_state;                           // 0: block not initialised, 1: guard vectors set, 2: holds a previous basis
  int no_subspace;                         // SB_OPT_NO_SUBSPACE: always run the Jacobi solver
  int warm_batches;                        // batches since the last cold start of the eigensolver
  int no_cluster16;                        // a 16-CTA cluster launch failed once on this device: use 8
  const float *d_in_mu, *d_in_is;          // optional input transform (x - mu) * inv_sigma, caller-owned (sb_maha_set_input_transform)
  float* d_Vf;
  double* d_evals;                 // [KP]
  long long* d_gate;               // [0] b*, [1] i* (global row), [2] n_flagged, [3] eigh sweeps, [4] total rows
  double* d_score_sum;             // [1]
};
