// Feedback metrics of the Mahalanobis outlier detector: confusion matrix, accuracy / precision / recall / F1 / F2 over
// everything seen so far and over the last `roll_window` observations, and the outlier counts.
//
// Replaces components/outlier-detection/mahalanobis/utils.py:106-142 (performance, outlier_stats) as called from
// OutlierMahalanobis.send_feedback (OutlierMahalanobis.py:34-66).  The reference keeps every prediction and label in
// Python lists, re-flattens them on every call and runs six scikit-learn metric functions over the full history; here
// the history is four running counters plus a ring of the last `roll_window` (prediction, label) pairs, all in a
// caller-owned device buffer, and a call costs one pass over the new flags.
//
// State layout (int64 words): [0..3] tn, fp, fn, tp  [4] observations seen  [5] sum(pred)  [6] sum(truth)
//                             [7] ring capacity     [8 ..] ring, one byte per slot: bit 0 = prediction, bit 1 = label
#include "common.cuh"

namespace sb {
namespace {

constexpr int kFbHeader = 8;  // int64 words before the ring

__global__ void __launch_bounds__(256)
feedback_count(const int32_t* __restrict__ pred, const int32_t* __restrict__ truth, int64_t n, int64_t* __restrict__ state) {
  // labels outside {0, 1} are ignored by the confusion matrix (labels=[0,1]) but still enter the sums, as in the reference
  long long c[4] = {0, 0, 0, 0}, sp = 0, st = 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const int p = pred[i], t = truth[i];
    sp += p;
    st += t;
    if ((p == 0 || p == 1) && (t == 0 || t == 1)) ++c[2 * t + p];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int k = 0; k < 4; ++k) c[k] += __shfl_xor_sync(0xffffffffu, c[k], o);
    sp += __shfl_xor_sync(0xffffffffu, sp, o);
    st += __shfl_xor_sync(0xffffffffu, st, o);
  }
  if ((threadIdx.x & 31) == 0) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (c[k]) atomicAdd((unsigned long long*)(state + k), (unsigned long long)c[k]);
    if (sp) atomicAdd((unsigned long long*)(state + 5), (unsigned long long)sp);
    if (st) atomicAdd((unsigned long long*)(state + 6), (unsigned long long)st);
  }
}

// the last min(n, W) pairs of the batch go into the ring at the positions they have in the stream
__global__ void __launch_bounds__(256)
feedback_ring(const int32_t* __restrict__ pred, const int32_t* __restrict__ truth, int64_t n, int64_t* __restrict__ state) {
  const int64_t W = state[7], seen = state[4];
  uint8_t* ring = reinterpret_cast<uint8_t*>(state + kFbHeader);
  const int64_t first = n > W ? n - W : 0;
  for (int64_t i = first + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int p = pred[i], t = truth[i];
    // out-of-range labels keep a marker so that the rolling confusion matrix ignores them too
    const uint8_t code = ((p == 0 || p == 1) && (t == 0 || t == 1)) ? (uint8_t)(p | (t << 1)) : (uint8_t)0x80;
    ring[(seen + i) % W] = code;
  }
}

__device__ double prf_div(double a, double b) { return b == 0.0 ? 0.0 : a / b; }  // sklearn zero_division -> 0

// scikit-learn's binary scores from the counts (precision_recall_fscore_support with pos_label = 1)
__device__ void scores_from_counts(double tn, double fp, double fn, double tp, double total, double* out5) {
  out5[0] = total > 0.0 ? (tp + tn) / total : 0.0 / 0.0;  // accuracy_score of an empty list is nan
  const double prec = prf_div(tp, tp + fp), rec = prf_div(tp, tp + fn);
  out5[1] = prec;
  out5[2] = rec;
  for (int b = 1; b <= 2; ++b) {
    const double b2 = (double)(b * b);
    out5[2 + b] = prf_div((1.0 + b2) * tp, (1.0 + b2) * tp + b2 * fn + fp);
  }
}

// one thread: closes the batch (advance the counter) and writes the 18 metrics in the reference's order
__global__ void feedback_finish(int64_t n, int64_t* __restrict__ state, double* __restrict__ metrics) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const int64_t W = state[7];
  const int64_t seen = state[4] + n;
  state[4] = seen;
  const uint8_t* ring = reinterpret_cast<const uint8_t*>(state + kFbHeader);
  const int64_t live = seen < W ? seen : W;
  long long c[4] = {0, 0, 0, 0}, rp = 0, rt = 0, counted = 0;
  for (int64_t k = 0; k < live; ++k) {
    const uint8_t code = ring[(seen - 1 - k) % W];
    if (code & 0x80) continue;
    ++c[code & 3];
    rp += code & 1;
    rt += (code >> 1) & 1;
    ++counted;
  }
  const double tn = (double)state[0], fp = (double)state[1], fn = (double)state[2], tp = (double)state[3];
  metrics[0] = tn; metrics[1] = fp; metrics[2] = fn; metrics[3] = tp;
  // accuracy_score divides by the number of samples, the confusion matrix only counts labels in {0, 1}; they are the
  // same for binary feedback, which is what the reference supports
  scores_from_counts(tn, fp, fn, tp, (double)seen, metrics + 4);
  scores_from_counts((double)c[0], (double)c[1], (double)c[2], (double)c[3], (double)live, metrics + 9);
  (void)counted;
  metrics[14] = (double)rp;
  metrics[15] = (double)rt;
  metrics[16] = (double)state[5];
  metrics[17] = (double)state[6];
}

__global__ void feedback_reset(int64_t* __restrict__ state, int64_t words, int64_t W) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < words; i += (int64_t)gridDim.x * blockDim.x)
    state[i] = i == 7 ? W : 0;
}

}  // namespace
}  // namespace sb

using namespace sb;

extern "C" uint64_t sb_feedback_state_bytes(int32_t roll_window) {
  if (roll_window < 1) return 0;
  return (uint64_t)kFbHeader * 8 + (((uint64_t)roll_window + 7) & ~7ull);
}

extern "C" int32_t sb_feedback_reset(void* d_state, int32_t roll_window, uintptr_t stream) {
  SB_RANGE("sb_feedback_reset");
  SB_REQUIRE(d_state && roll_window >= 1, SB_ERR_INVALID, "sb_feedback_reset: bad arguments");
  SB_REQUIRE(((uintptr_t)d_state & 7) == 0, SB_ERR_INVALID, "sb_feedback_reset: state must be 8-byte aligned");
  const int64_t words = (int64_t)(sb_feedback_state_bytes(roll_window) / 8);
  feedback_reset<<<1, 256, 0, (cudaStream_t)stream>>>((int64_t*)d_state, words, roll_window);
  SB_LAUNCH_CHECK();
  return SB_OK;
}

extern "C" int32_t sb_feedback_update(const int32_t* d_pred, const int32_t* d_truth, int64_t n, void* d_state,
                                      double* d_metrics18, uintptr_t stream) {
  SB_RANGE("sb_feedback_update");
  SB_REQUIRE(n >= 0 && d_state && d_metrics18, SB_ERR_INVALID, "sb_feedback_update: bad arguments");
  SB_REQUIRE(n == 0 || (d_pred && d_truth), SB_ERR_INVALID, "sb_feedback_update: null flags");
  cudaStream_t st = (cudaStream_t)stream;
  int64_t* state = (int64_t*)d_state;
  if (n > 0) {
    int64_t want = ceil_div(n, (int64_t)256 * 8);
    const int grid = (int)(want < 1 ? 1 : (want > (int64_t)kNumSMs * 8 ? (int64_t)kNumSMs * 8 : want));
    feedback_count<<<grid, 256, 0, st>>>(d_pred, d_truth, n, state);
    SB_LAUNCH_CHECK();
    feedback_ring<<<4, 256, 0, st>>>(d_pred, d_truth, n, state);
    SB_LAUNCH_CHECK();
  }
  feedback_finish<<<1, 32, 0, st>>>(n, state, d_metrics18);
  SB_LAUNCH_CHECK();
  return SB_OK;
}
