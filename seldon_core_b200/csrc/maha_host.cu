// Mahalanobis detector: handle, workspace carving and the C-ABI entry points (see include/sb200.h).
//
// State and API follow CoreMahalanobis.py of the reference: constructor L26-40, _get_preds L68-151 split into
// pass1 / finalize / pass2 / commit so that a row-sharded batch needs exactly one exchange (the caller gathers
// the per-rank statistics records between pass1 and finalize).
#include <limits.h>
#include <stdlib.h>
#include <string.h>

#include "maha.cuh"

namespace sb {
namespace {

int padded_k(int k) { return k <= 4 ? 4 : k <= 8 ? 8 : k <= 16 ? 16 : 32; }

// carve the caller's workspace; returns the total number of bytes needed for `nsub_max` sub-segments
// (pass-1 layout) and `nsub2_max` (split pass-2 layout)
uint64_t carve(const sb_maha* h, int64_t nsub_max, int64_t nsub2_max, int64_t rows_max, void* base, MahaWs* ws) {
  const uint64_t d = h->d, KP = h->KP, KQ = KP + KP * KP, PS = d + d * d, SL = 1 + PS;
  uint64_t off = 0;
  auto take = [&](uint64_t bytes) {
    uint64_t o = off;
    off = align_up(off + bytes, 256);
    return o;
  };
  const bool split = (h->d == 64 || h->d == 128 || h->d == 256) && h->KP == 16;
  if (!split || nsub2_max < nsub_max) nsub2_max = nsub_max;
  uint64_t o_part = take((uint64_t)nsub_max * PS * sizeof(float));
  uint64_t o_proj = take((uint64_t)(nsub2_max + 1) * KQ * sizeof(double));
  uint64_t o_P = take((uint64_t)nsub2_max * KP * KP * sizeof(float));
  uint64_t o_Z = take((uint64_t)nsub2_max * KP * sizeof(float));
  uint64_t o_fl = take((uint64_t)nsub2_max * sizeof(int32_t));
  uint64_t o_st = take(SL * sizeof(double));
  uint64_t o_pre = take(SL * sizeof(double));
  uint64_t o_rp = take(KQ * sizeof(double));
  uint64_t o_sl = take((uint64_t)reduce_slices((int)d) * PS * sizeof(double));
  uint64_t o_ct = take((uint64_t)(nsub2_max / 32 + 2) * KQ * sizeof(double));
  uint64_t o_pz = take(split ? (uint64_t)rows_max * 16 * sizeof(float) : 0);
  uint64_t o_pd = take(split ? (uint64_t)rows_max * 16 * sizeof(float) : 0);
  uint64_t o_pf = take(split ? (uint64_t)rows_max : 0);
  uint64_t o_W = take(d * d * sizeof(double));
  uint64_t o_V = take(d * d * sizeof(double));
  if (ws) {
    char* b = (char*)base;
    ws->partials = (float*)(b + o_part);
    ws->subproj = (double*)(b + o_proj);
    ws->sub_P = (float*)(b + o_P);
    ws->sub_Z = (float*)(b + o_Z);
    ws->sub_flags = (int32_t*)(b + o_fl);
    ws->stats = (double*)(b + o_st);
    ws->rank_pre = (double*)(b + o_pre);
    ws->rank_proj = (double*)(b + o_rp);
    ws->slices = (double*)(b + o_sl);
    ws->chunk_tot = (double*)(b + o_ct);
    ws->proj_Z = (float*)(b + o_pz);
    ws->proj_D = (float*)(b + o_pd);
    ws->proj_flag = (uint8_t*)(b + o_pf);
    ws->eig_W = (double*)(b + o_W);
    ws->eig_V = (double*)(b + o_V);
    ws->bytes = off;
  }
  return off;
}

__global__ void to_float_state(const double* __restrict__ mean, const double* __restrict__ lo,
                               const double* __restrict__ hi, int d, float* __restrict__ mean_f,
                               float* __restrict__ lo_f, float* __restrict__ hi_f) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= d) return;
  mean_f[f] = (float)mean[f];
  lo_f[f] = (float)lo[f];
  hi_f[f] = (float)hi[f];
}

int refresh_float_state(sb_maha* h, cudaStream_t st) {
  to_float_state<<<(h->d + 127) / 128, 128, 0, st>>>(h->d_mean, h->d_lo, h->d_hi, h->d, h->d_mean_f, h->d_lo_f,
                                                     h->d_hi_f);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

int get_ws(const sb_maha* h, int64_t rows, void* d_ws, uint64_t ws_bytes, MahaWs* ws) {
  SB_REQUIRE(d_ws != nullptr && ((uintptr_t)d_ws & 255) == 0, SB_ERR_INVALID, "workspace must be 256-byte aligned");
  uint64_t need = carve(h, make_layout(rows).nsub, make_layout2(rows).nsub, rows, d_ws, ws);
  SB_REQUIRE(need <= ws_bytes, SB_ERR_WORKSPACE, "workspace too small: need %llu bytes, got %llu",
             (unsigned long long)need, (unsigned long long)ws_bytes);
  return SB_OK;
}

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" int32_t sb_maha_create(sb_maha_t** out, int32_t d, int32_t n_components, double threshold, double n_stdev,
                                  int64_t start_clip, int64_t max_n) {
  SB_RANGE("sb_maha_create");
  SB_REQUIRE(out != nullptr, SB_ERR_INVALID, "sb_maha_create: null handle pointer");
  SB_REQUIRE(d >= 1 && d <= kMaxD, SB_ERR_INVALID, "sb_maha_create: d=%d outside 1..%d", d, kMaxD);
  SB_REQUIRE(n_components >= 1, SB_ERR_INVALID, "sb_maha_create: n_components must be >= 1");
  const int k = n_components < d ? n_components : d;  // CoreMahalanobis.py:78
  SB_REQUIRE(k <= kMaxKP, SB_ERR_INVALID, "sb_maha_create: min(n_components, d)=%d exceeds %d", k, kMaxKP);
  // max_n == 1 makes the reference divide by zero at CoreMahalanobis.py:134 (n+i-2 == 0)
  SB_REQUIRE(max_n != 1, SB_ERR_INVALID, "sb_maha_create: max_n=1 is not usable (reference raises ZeroDivisionError)");
  sb_maha* h = new sb_maha();
  memset(h, 0, sizeof *h);
  h->d = d; h->k = k; h->KP = padded_k(k);
  h->threshold = threshold; h->n_stdev = n_stdev; h->start_clip = start_clip; h->max_n = max_n;
  h->gram_terms = 3;
  const size_t dd = (size_t)d * d, KP = h->KP;
  double* dbl = nullptr;
  const size_t n_dbl = 2 * (d + dd + 2 * (size_t)d) + (size_t)d * KP + KP * KP + KP + 1 + dd;
  cudaError_t e = cudaMalloc(&dbl, n_dbl * sizeof(double));
  if (e != cudaSuccess) { set_error("sb_maha_create: cudaMalloc: %s", cudaGetErrorString(e)); delete h; return SB_ERR_CUDA; }
  cudaMemset(dbl, 0, n_dbl * sizeof(double));
  double* p = dbl;
  h->d_mean = p; p += d; h->d_C = p; p += dd; h->d_lo = p; p += d; h->d_hi = p; p += d;
  h->d_mean1 = p; p += d; h->d_C1 = p; p += dd; h->d_lo1 = p; p += d; h->d_hi1 = p; p += d;
  h->d_V = p; p += (size_t)d * KP; h->d_PC0 = p; p += KP * KP; h->d_evals = p; p += KP; h->d_score_sum = p; p += 1;
  h->d_Vfull = p; p += dd;
  if (subspace_supported(d, (int)KP)) {
    e = cudaMalloc(&h->d_Vsub, (size_t)subspace_block_doubles(d) * sizeof(double));
    if (e != cudaSuccess) { set_error("sb_maha_create: cudaMalloc: %s", cudaGetErrorString(e)); cudaFree(dbl); delete h; return SB_ERR_CUDA; }
  }
  float* flt = nullptr;
  const size_t d4 = ((size_t)d + 3) & ~(size_t)3;  // keep every fp32 array 16-byte aligned
  e = cudaMalloc(&flt, (3 * d4 + (size_t)d * KP) * sizeof(float) + 48 * sizeof(long long));
  if (e != cudaSuccess) { set_error("sb_maha_create: cudaMalloc: %s", cudaGetErrorString(e)); cudaFree(dbl); delete h; return SB_ERR_CUDA; }
  cudaMemset(flt, 0, (3 * d4 + (size_t)d * KP) * sizeof(float) + 48 * sizeof(long long));
  h->d_gate = (long long*)flt;  // 8-byte aligned slots first
  float* q = flt + 96;  // 48 eight-byte slots: counters [0,8), Jacobi sweep flags behind them
  h->d_mean_f = q; q += d4; h->d_lo_f = q; q += d4; h->d_hi_f = q; q += d4; h->d_Vf = q;
  *out = h;
  return SB_OK;
}

extern "C" int32_t sb_maha_destroy(sb_maha_t* h) {
  SB_RANGE("sb_maha_destroy");
  if (!h) return SB_OK;
  if (h->profile)
    for (int i = 0; i < 10; ++i) cudaEventDestroy(h->ev[i]);
  cudaFree(h->d_mean);
  cudaFree(h->d_gate);
  if (h->d_Vsub) cudaFree(h->d_Vsub);
  delete h;
  return SB_OK;
}

extern "C" int32_t sb_maha_workspace_bytes(const sb_maha_t* h, int64_t max_rows, uint64_t* bytes) {
  SB_RANGE("sb_maha_workspace_bytes");
  SB_REQUIRE(h && bytes && max_rows >= 1, SB_ERR_INVALID, "sb_maha_workspace_bytes: bad arguments");
  *bytes = carve(h, max_subsegments(max_rows), max_subsegments2(max_rows), max_rows, nullptr, nullptr);
  return SB_OK;
}

extern "C" int64_t sb_maha_stats_len(const sb_maha_t* h) { return h ? 1 + (int64_t)h->d + (int64_t)h->d * h->d : 0; }

extern "C" int32_t sb_maha_set_input_transform(sb_maha_t* h, const float* d_mu, const float* d_inv_sigma) {
  SB_RANGE("sb_maha_set_input_transform");
  SB_REQUIRE(h, SB_ERR_INVALID, "sb_maha_set_input_transform: null handle");
  SB_REQUIRE((d_mu == nullptr) == (d_inv_sigma == nullptr), SB_ERR_INVALID,
             "sb_maha_set_input_transform: mu and inv_sigma go together");
  SB_REQUIRE(h->stage == 0, SB_ERR_STATE, "sb_maha_set_input_transform: a batch is in flight");
  SB_REQUIRE(d_mu == nullptr || ((((uintptr_t)d_mu | (uintptr_t)d_inv_sigma) & 15) == 0), SB_ERR_INVALID,
             "sb_maha_set_input_transform: the vectors must be 16-byte aligned");
  h->d_in_mu = d_mu;
  h->d_in_is = d_inv_sigma;
  return SB_OK;
}

extern "C" int32_t sb_maha_set_option(sb_maha_t* h, int32_t option, int64_t value) {
  SB_RANGE("sb_maha_set_option");
  SB_REQUIRE(h, SB_ERR_INVALID, "sb_maha_set_option: null handle");
  if (option == SB_OPT_GRAM_TERMS) {
    SB_REQUIRE(value == 1 || value == 3, SB_ERR_INVALID, "SB_OPT_GRAM_TERMS must be 1 or 3");
    h->gram_terms = (int)value;
  } else if (option == SB_OPT_FORCE_GENERIC) {
    h->force_generic = value != 0;
  } else if (option == SB_OPT_LEGACY_MMA) {
    h->legacy_mma = value != 0;
  } else if (option == SB_OPT_FUSED_PASS2) {
    h->fused_pass2 = value != 0;
  } else if (option == SB_OPT_NO_SUBSPACE) {
    h->no_subspace = value != 0;
  } else if (option == SB_OPT_PROFILE) {
    if (value && !h->profile)
      for (int i = 0; i < 10; ++i) SB_CUDA(cudaEventCreate(&h->ev[i]));
    h->profile = value != 0;
  } else {
    set_error("sb_maha_set_option: unknown option %d", option);
    return SB_ERR_INVALID;
  }
  return SB_OK;
}

extern "C" int32_t sb_maha_pass1_stats(sb_maha_t* h, const float* d_X, int64_t rows, int64_t ld, void* d_ws,
                                       uint64_t ws_bytes, double* d_stats_out, uintptr_t stream) {
  SB_RANGE("sb_maha_pass1_stats");
  SB_REQUIRE(h && rows >= 0 && (rows == 0 || (d_X && ld >= h->d)), SB_ERR_INVALID, "sb_maha_pass1_stats: bad arguments");
  SB_REQUIRE(h->stage == 0, SB_ERR_STATE, "sb_maha_pass1_stats: previous batch not committed");
  MahaWs ws;
  int rc = get_ws(h, rows > 0 ? rows : 1, d_ws, ws_bytes, &ws);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  h->rows_local = rows;
  if (rows == 0) {
    // a rank that holds no row of this request batch (fewer rows than ranks): its record is all zeros; it still takes
    // part in the exchange, forms the same covariance, eigenvectors and state, and scores nothing
    SB_CUDA(cudaMemsetAsync(d_stats_out ? d_stats_out : ws.stats, 0, (size_t)sb_maha_stats_len(h) * sizeof(double), st));
    h->n_eff = (h->max_n > 0 && h->n > h->max_n) ? h->max_n : h->n;
    h->use_clip = (h->n > h->start_clip) && h->has_clip;
    h->lay = h->lay2 = make_layout(1);
    h->stage = 1;
    return SB_OK;
  }
  h->lay = make_layout(rows);
  h->n_eff = (h->max_n > 0 && h->n > h->max_n) ? h->max_n : h->n;  // CoreMahalanobis.py:79-82
  h->use_clip = (h->n > h->start_clip) && h->has_clip;               // L85 (uncapped n)
  h->split_pass2 = h->KP == 16 && (h->d == 64 || h->d == 128 || h->d == 256) && !h->force_generic &&
                   !h->fused_pass2 && (ld % 4 == 0) && (((uintptr_t)d_X & 15) == 0);
  h->lay2 = h->split_pass2 ? make_layout2(rows) : h->lay;
  if (h->split_pass2) h->lay = make_layout1(rows);  // fewer, longer pass-1 sub-segments: the partials are only summed
  SB_REQUIRE(h->d_in_mu == nullptr || (h->split_pass2 && !h->legacy_mma), SB_ERR_INVALID,
             "sb_maha_pass1_stats: the fused input transform needs the tensor-core path (d = 64, 128 or 256, "
             "n_components <= 16, 16-byte aligned rows, no SB_OPT_FORCE_GENERIC / FUSED_PASS2 / LEGACY_MMA)");
  if ((rc = launch_pass1(h, d_X, ld, ws, st))) return rc;
  if ((rc = launch_reduce(h, ws, d_stats_out ? d_stats_out : ws.stats, st))) return rc;
  h->stage = 1;
  return SB_OK;
}

extern "C" int32_t sb_maha_finalize(sb_maha_t* h, const double* d_all_stats, int32_t nranks, int32_t rank,
                                    int64_t total_rows, int64_t row0, void* d_ws, uint64_t ws_bytes,
                                    uintptr_t stream) {
  SB_RANGE("sb_maha_finalize");
  SB_REQUIRE(h && nranks >= 1 && rank >= 0 && rank < nranks, SB_ERR_INVALID, "sb_maha_finalize: bad arguments");
  SB_REQUIRE(h->stage == 1, SB_ERR_STATE, "sb_maha_finalize: call sb_maha_pass1_stats first");
  SB_REQUIRE(total_rows >= 1 && (h->rows_local > 0 || (d_all_stats && nranks > 1)), SB_ERR_INVALID,
             "sb_maha_finalize: an empty shard needs the records of the other ranks");
  SB_REQUIRE(total_rows >= h->rows_local && row0 >= 0 && row0 + h->rows_local <= total_rows, SB_ERR_INVALID,
             "sb_maha_finalize: shard [%lld, +%lld) outside a batch of %lld rows", (long long)row0,
             (long long)h->rows_local, (long long)total_rows);
  MahaWs ws;
  int rc = get_ws(h, h->rows_local > 0 ? h->rows_local : 1, d_ws, ws_bytes, &ws);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  h->rows_total = total_rows;
  h->row0 = row0;
  const double* all = d_all_stats ? d_all_stats : ws.stats;
  if ((rc = launch_finalize_cov(h, all, nranks, rank, ws, st))) return rc;
  if ((rc = launch_eigh(h, ws, st))) return rc;
  if (h->rows_local > 0 && (rc = launch_prefix(h, ws, ws.rank_proj, st))) return rc;
  h->stage = 2;
  return SB_OK;
}

extern "C" int32_t sb_maha_pass2_score(sb_maha_t* h, const float* d_X, int64_t rows, int64_t ld, float* d_score,
                                       int32_t* d_flag, void* d_ws, uint64_t ws_bytes, uintptr_t stream) {
  SB_RANGE("sb_maha_pass2_score");
  SB_REQUIRE(h && (rows == 0 || (d_X && d_score && d_flag && ld >= h->d)), SB_ERR_INVALID, "sb_maha_pass2_score: bad arguments");
  SB_REQUIRE(h->stage == 2, SB_ERR_STATE, "sb_maha_pass2_score: call sb_maha_finalize first");
  SB_REQUIRE(rows == h->rows_local, SB_ERR_INVALID, "sb_maha_pass2_score: rows differ from pass 1");
  if (rows == 0) {  // empty shard: nothing to score; the sums of the last batch read as zero
    SB_CUDA(cudaMemsetAsync(h->d_gate + 2, 0, sizeof(long long), (cudaStream_t)stream));
    SB_CUDA(cudaMemsetAsync(h->d_score_sum, 0, sizeof(double), (cudaStream_t)stream));
    h->stage = 3;
    return SB_OK;
  }
  SB_REQUIRE(!h->split_pass2 || ((ld % 4 == 0) && (((uintptr_t)d_X & 15) == 0)), SB_ERR_INVALID,
             "sb_maha_pass2_score: X must have the alignment it had in pass 1");
  MahaWs ws;
  int rc = get_ws(h, rows, d_ws, ws_bytes, &ws);
  if (rc) return rc;
  if ((rc = launch_score(h, d_X, ld, d_score, d_flag, ws, (cudaStream_t)stream))) return rc;
  h->stage = 3;
  return SB_OK;
}

extern "C" int32_t sb_maha_commit(sb_maha_t* h, uintptr_t stream) {
  SB_RANGE("sb_maha_commit");
  SB_REQUIRE(h, SB_ERR_INVALID, "sb_maha_commit: null handle");
  SB_REQUIRE(h->stage >= 2, SB_ERR_STATE, "sb_maha_commit: nothing to commit");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t d = h->d;
  // CoreMahalanobis.py:139-144
  SB_CUDA(cudaMemcpyAsync(h->d_mean, h->d_mean1, d * sizeof(double), cudaMemcpyDeviceToDevice, st));
  SB_CUDA(cudaMemcpyAsync(h->d_C, h->d_C1, d * d * sizeof(double), cudaMemcpyDeviceToDevice, st));
  h->n += h->rows_total;
  h->has_state = 1;
  if (h->n > h->start_clip) {
    SB_CUDA(cudaMemcpyAsync(h->d_lo, h->d_lo1, d * sizeof(double), cudaMemcpyDeviceToDevice, st));
    SB_CUDA(cudaMemcpyAsync(h->d_hi, h->d_hi1, d * sizeof(double), cudaMemcpyDeviceToDevice, st));
    h->has_clip = 1;
  }
  int rc = refresh_float_state(h, st);
  if (rc) return rc;
  h->stage = 0;
  return SB_OK;
}

extern "C" int32_t sb_maha_batch(sb_maha_t* h, const float* d_X, int64_t rows, int64_t ld, float* d_score,
                                 int32_t* d_flag, void* d_ws, uint64_t ws_bytes, uintptr_t stream) {
  SB_RANGE("sb_maha_batch");
  int rc;
  if ((rc = sb_maha_pass1_stats(h, d_X, rows, ld, d_ws, ws_bytes, nullptr, stream))) return rc;
  if ((rc = sb_maha_finalize(h, nullptr, 1, 0, rows, 0, d_ws, ws_bytes, stream))) { h->stage = 0; return rc; }
  if ((rc = sb_maha_pass2_score(h, d_X, rows, ld, d_score, d_flag, d_ws, ws_bytes, stream))) { h->stage = 0; return rc; }
  return sb_maha_commit(h, stream);
}

extern "C" int32_t sb_maha_abort_batch(sb_maha_t* h) {
  SB_RANGE("sb_maha_abort_batch");
  if (h) h->stage = 0;
  return SB_OK;
}

extern "C" int32_t sb_maha_get_state(sb_maha_t* h, double* h_mean, double* h_C, double* h_clip_lo, double* h_clip_hi,
                                     int64_t* n, int32_t* has_state, int32_t* has_clip, uintptr_t stream) {
  SB_RANGE("sb_maha_get_state");
  SB_REQUIRE(h, SB_ERR_INVALID, "sb_maha_get_state: null handle");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t d = h->d;
  if (h_mean) SB_CUDA(cudaMemcpyAsync(h_mean, h->d_mean, d * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (h_C) SB_CUDA(cudaMemcpyAsync(h_C, h->d_C, d * d * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (h_clip_lo) SB_CUDA(cudaMemcpyAsync(h_clip_lo, h->d_lo, d * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (h_clip_hi) SB_CUDA(cudaMemcpyAsync(h_clip_hi, h->d_hi, d * sizeof(double), cudaMemcpyDeviceToHost, st));
  if (h_mean || h_C || h_clip_lo || h_clip_hi) SB_CUDA(cudaStreamSynchronize(st));  // the scalars are host mirrors
  if (n) *n = h->n;
  if (has_state) *has_state = h->has_state;
  if (has_clip) *has_clip = h->has_clip;
  return SB_OK;
}

extern "C" int32_t sb_maha_set_state(sb_maha_t* h, const double* h_mean, const double* h_C, const double* h_clip_lo,
                                     const double* h_clip_hi, int64_t n, int32_t has_state, int32_t has_clip,
                                     uintptr_t stream) {
  SB_RANGE("sb_maha_set_state");
  SB_REQUIRE(h && n >= 0, SB_ERR_INVALID, "sb_maha_set_state: bad arguments");
  SB_REQUIRE(!has_state || (h_mean && h_C), SB_ERR_INVALID, "sb_maha_set_state: mean and C required");
  SB_REQUIRE(!has_clip || (h_clip_lo && h_clip_hi), SB_ERR_INVALID, "sb_maha_set_state: clip bounds required");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t d = h->d;
  if (has_state) {
    SB_CUDA(cudaMemcpyAsync(h->d_mean, h_mean, d * sizeof(double), cudaMemcpyHostToDevice, st));
    SB_CUDA(cudaMemcpyAsync(h->d_C, h_C, d * d * sizeof(double), cudaMemcpyHostToDevice, st));
  } else {
    SB_CUDA(cudaMemsetAsync(h->d_mean, 0, d * sizeof(double), st));
    SB_CUDA(cudaMemsetAsync(h->d_C, 0, d * d * sizeof(double), st));
  }
  if (has_clip) {
    SB_CUDA(cudaMemcpyAsync(h->d_lo, h_clip_lo, d * sizeof(double), cudaMemcpyHostToDevice, st));
    SB_CUDA(cudaMemcpyAsync(h->d_hi, h_clip_hi, d * sizeof(double), cudaMemcpyHostToDevice, st));
  }
  h->n = n; h->has_state = has_state; h->has_clip = has_clip; h->stage = 0;
  h->warm_batches = 0;
  if (h->sub_state == 2) h->sub_state = 1;  // the kept block belongs to the state that was just replaced
  int rc = refresh_float_state(h, st);
  if (rc) return rc;
  SB_CUDA(cudaStreamSynchronize(st));
  return SB_OK;
}

extern "C" int32_t sb_maha_last_batch_sums(sb_maha_t* h, int64_t* n_flagged, double* score_sum, int64_t* gate_row,
                                           uintptr_t stream) {
  SB_RANGE("sb_maha_last_batch_sums");
  SB_REQUIRE(h, SB_ERR_INVALID, "sb_maha_last_batch_sums: null handle");
  cudaStream_t st = (cudaStream_t)stream;
  long long g[5];
  double s;
  SB_CUDA(cudaMemcpyAsync(g, h->d_gate, sizeof g, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaMemcpyAsync(&s, h->d_score_sum, sizeof s, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaStreamSynchronize(st));
  if (n_flagged) *n_flagged = g[2];
  if (score_sum) *score_sum = s;
  if (gate_row) *gate_row = g[1] == LLONG_MAX ? -1 : g[1];
  return SB_OK;
}

extern "C" int32_t sb_maha_kernel_times(sb_maha_t* h, float* out_ms5, uintptr_t stream) {
  SB_RANGE("sb_maha_kernel_times");
  SB_REQUIRE(h && out_ms5, SB_ERR_INVALID, "sb_maha_kernel_times: bad arguments");
  SB_REQUIRE(h->profile, SB_ERR_STATE, "sb_maha_kernel_times: enable SB_OPT_PROFILE first");
  SB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  for (int i = 0; i < 5; ++i) SB_CUDA(cudaEventElapsedTime(&out_ms5[i], h->ev[2 * i], h->ev[2 * i + 1]));
  return SB_OK;
}

extern "C" int32_t sb_maha_debug_counters(sb_maha_t* h, int64_t* out8, uintptr_t stream) {
  SB_RANGE("sb_maha_debug_counters");
  SB_REQUIRE(h && out8, SB_ERR_INVALID, "sb_maha_debug_counters: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  long long g[8];
  const char* off_env = getenv("SB200_GATE_OFFSET");  // kernel-timing builds park their clock counts in later slots
  const int off = off_env ? atoi(off_env) : 0;
  SB_REQUIRE(off >= 0 && off <= 40, SB_ERR_INVALID, "SB200_GATE_OFFSET out of range");
  SB_CUDA(cudaMemcpyAsync(g, h->d_gate + off, sizeof g, cudaMemcpyDeviceToHost, st));
  SB_CUDA(cudaStreamSynchronize(st));
  for (int i = 0; i < 8; ++i) out8[i] = g[i];
  if (off == 0) {
    out8[5] = h->lay.nsub;
    out8[6] = h->lay.sublen;
  }
  return SB_OK;
}

extern "C" int32_t sb_maha_get_eigvecs(sb_maha_t* h, double* h_V, uintptr_t stream) {
  SB_RANGE("sb_maha_get_eigvecs");
  SB_REQUIRE(h && h_V, SB_ERR_INVALID, "sb_maha_get_eigvecs: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = (size_t)h->d * h->KP;
  double* tmp = new double[n];
  cudaError_t e = cudaMemcpyAsync(tmp, h->d_V, n * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) { delete[] tmp; set_error("sb_maha_get_eigvecs: %s", cudaGetErrorString(e)); return SB_ERR_CUDA; }
  for (int f = 0; f < h->d; ++f)
    for (int c = 0; c < h->k; ++c) h_V[(size_t)f * h->k + c] = tmp[(size_t)f * h->KP + c];
  delete[] tmp;
  return SB_OK;
}
