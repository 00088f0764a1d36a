/* sb200.h - C ABI of libsb200.so: the B200-native numeric data plane behind the seldon_core component API.
 *
 * The reference (marcredhat/seldon-core 0.5.0-SNAPSHOT) is pure Python/Java on this path and has no FFI of
 * its own; each entry point below names the reference function whose per-batch arithmetic it replaces
 * (paths relative to the reference root).  INTEGRATION.md shows the ctypes stub a maintainer adds on the
 * reference side.
 *
 * Conventions
 *   - every function returns 0 on success or a negative SB_ERR_* code; sb_last_error() gives the message of
 *     the last failure on the calling thread;
 *   - all d_* pointers are DEVICE pointers owned by the caller (e.g. torch.Tensor.data_ptr()); h_* are host
 *     pointers; nothing is retained after the call returns except inside an opaque handle;
 *   - `stream` is a cudaStream_t passed as uintptr_t; calls are asynchronous and stream-ordered unless stated;
 *   - handles are not thread-safe: the Python layer serialises calls per instance (the reference wrapper
 *     serves 10 gRPC threads, python/seldon_core/wrapper.py:132-133);
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails with SB_ERR_CUDA.
 */
#ifndef SB200_H
#define SB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SB_OK 0
#define SB_ERR_INVALID (-1)   /* bad argument (shape, null pointer, unsupported size) */
#define SB_ERR_CUDA (-2)      /* CUDA runtime error, message in sb_last_error() */
#define SB_ERR_WORKSPACE (-3) /* caller-provided workspace too small */
#define SB_ERR_STATE (-4)     /* call sequence violated (e.g. pass2 before finalize) */

const char* sb_last_error(void);
int32_t sb_version(void);

/* ------------------------------------------------------------------------------------------------ codec
 * replaces python/seldon_core/utils.py:161-169 (grpc_datadef_to_array, tensor branch): the packed doubles of
 * a serialised Tensor start `values_off` bytes into `d_wire`; they are converted to fp32 rows of `d_out`
 * (leading dimension ld); the wire holds wire_nbytes bytes and a shape that claims more values than it carries is
 * rejected with SB_ERR_INVALID (the reference raises on the reshape, utils.py:169).  d_mu / d_inv_sigma (both or neither) fuse the standardise transformer
 * (components/outlier-detection/vae/CoreVAE.py:128): out = (x - mu) * inv_sigma, inv_sigma = 1/(sigma+1e-10). */
int32_t sb_unpack_tensor_f64(const uint8_t* d_wire, uint64_t wire_nbytes, uint64_t values_off, int64_t rows, int64_t cols,
                             const float* d_mu, const float* d_inv_sigma, float* d_out, int64_t ld,
                             uintptr_t stream);

/* replaces utils.py:174-176 (ndarray branch: MessageToDict + np.array) for a rows x cols numeric ListValue
 * with the fixed row stride of SURVEY App. C.  *d_status (int32, device) is set non-zero if any record tag
 * is not `0a 09 11` or a row header does not match (malformed / ragged payload). */
int32_t sb_unpack_listvalue(const uint8_t* d_wire, uint64_t nbytes, int64_t rows, int64_t cols, float* d_out,
                            int64_t ld, int32_t* d_status, uintptr_t stream);

/* replaces utils.py:235-259 (array_to_grpc_datadef, tensor: values=array.ravel().tolist()): n values of
 * in_dtype (SB_F32 / SB_I32) -> n little-endian doubles at d_wire_values. */
#define SB_F32 0
#define SB_I32 1
int32_t sb_pack_tensor_f64(const void* d_in, int32_t in_dtype, int64_t n, uint8_t* d_wire_values, uintptr_t stream);

/* replaces utils.py:281-304 (array_to_list_value).  cols == 0 encodes a 1-D list of `rows` values
 * (the flags of CoreMahalanobis.predict).  The number of bytes written is sb_listvalue_nbytes(rows, cols). */
int32_t sb_pack_listvalue(const void* d_in, int32_t in_dtype, int64_t rows, int64_t cols, uint8_t* d_wire,
                          uintptr_t stream);
uint64_t sb_listvalue_nbytes(int64_t rows, int64_t cols);
uint64_t sb_tensor_values_offset(int64_t rows, int64_t cols);

/* REST JSON: the text of a numeric "ndarray" value ([[..],[..]] or [..], whitespace allowed, NaN / Infinity /
 * -Infinity as CPython's json module writes them) -> fp32 matrix.  Replaces json.loads + np.array(datadef["ndarray"])
 * of python/seldon_core/utils.py:499-505 (extract_request_parts_json); numbers are converted with correct rounding to
 * binary64 (what float() does inside json.loads) and then narrowed to fp32 like the Tensor path.
 * Step 1: sb_json_ndarray_count fills d_counts4 = {numbers, '[' count, ']' count, characters that cannot occur in a
 * numeric array} and leaves per-chunk prefixes in the workspace (sb_json_workspace_bytes(nbytes) bytes, 8-byte
 * aligned).  The caller derives rows = '[' count - 1 (or 1 for a flat list) and cols = numbers / rows.
 * Step 2: sb_unpack_json_ndarray with the same text and workspace stores the values; *d_status is 0 when every token
 * was a number and every row has cols entries, 1 for malformed text, 2 for a number with more than 19 significant
 * digits whose rounding cannot be decided in 128-bit arithmetic. */
uint64_t sb_json_workspace_bytes(uint64_t nbytes);
int32_t sb_json_ndarray_count(const uint8_t* d_text, uint64_t nbytes, void* d_ws, uint64_t ws_bytes, int64_t* d_counts4,
                              uintptr_t stream);
int32_t sb_unpack_json_ndarray(const uint8_t* d_text, uint64_t nbytes, int64_t rows, int64_t cols, const void* d_ws,
                               float* d_out, int64_t ld, int32_t* d_status, uintptr_t stream);

/* REST JSON, encode direction: the text json.dumps writes for a rows x cols matrix (cols == 0: a flat list of `rows`
 * values) of in_dtype SB_F32 (every value widened to binary64 and written as float.__repr__ writes it: the shortest
 * decimal that reads back as the same double) or SB_I32 (ints).  Replaces the `.tolist()` + json.dumps of the response
 * path, python/seldon_core/utils.py:345-384 (construct_response_json: "ndarray": array.tolist() /
 * "tensor": {"values": array.ravel().tolist()}).  d_text must hold sb_json_text_capacity(rows, cols) bytes; the number of
 * bytes written goes to *d_nbytes (device); d_ws: sb_json_encode_workspace_bytes(rows, cols) bytes, 8-byte aligned. */
uint64_t sb_json_text_capacity(int64_t rows, int64_t cols);
uint64_t sb_json_encode_workspace_bytes(int64_t rows, int64_t cols);
int32_t sb_pack_json_ndarray(const void* d_in, int32_t in_dtype, int64_t rows, int64_t cols, void* d_ws, uint64_t ws_bytes,
                             uint8_t* d_text, uint64_t* d_nbytes, uintptr_t stream);

/* --------------------------------------------------------------------------------------------- combiner
 * replaces examples/combiners/mnist_combiner/MnistCombiner.py:8-11 and the N-child rule of
 * engine/src/main/java/io/seldon/engine/predictors/AverageCombinerUnit.java:35-79: out = (sum_j in_j) / N,
 * summed in child order.  h_inputs is a HOST array of n_children device pointers (n_children <= 16). */
int32_t sb_combiner_mean(const float* const* h_inputs, int32_t n_children, int64_t n_elems, float* d_out,
                         uintptr_t stream);

/* standalone standardise transformer (CoreVAE.py:128) on an fp32 device matrix. */
int32_t sb_standardise(const float* d_in, int64_t rows, int64_t cols, int64_t ld_in, const float* d_mu,
                       const float* d_inv_sigma, float* d_out, int64_t ld_out, uintptr_t stream);

/* ------------------------------------------------------------------------------------- feedback metrics
 * replaces components/outlier-detection/mahalanobis/utils.py:106-142 (performance, outlier_stats: confusion matrix,
 * accuracy / precision / recall / F1 / F2 over all feedback and over the last roll_window observations, outlier
 * counts) as called by OutlierMahalanobis.send_feedback (OutlierMahalanobis.py:34-66, roll_window = 100 at L30).
 * The history is a caller-owned device buffer of sb_feedback_state_bytes(roll_window) bytes (four counters, two
 * sums, a ring of the last roll_window pairs); reset once, then one update per feedback batch.
 * d_metrics18 receives, in the reference's order: tn fp fn tp | acc prec rec f1 f2 (total) | acc prec rec f1 f2
 * (rolling) | predicted outliers (rolling), labelled outliers (rolling), predicted (total), labelled (total). */
uint64_t sb_feedback_state_bytes(int32_t roll_window);
int32_t sb_feedback_reset(void* d_state, int32_t roll_window, uintptr_t stream);
int32_t sb_feedback_update(const int32_t* d_pred, const int32_t* d_truth, int64_t n, void* d_state,
                           double* d_metrics18, uintptr_t stream);

/* ------------------------------------------------------------------------------------------ mahalanobis
 * replaces components/outlier-detection/mahalanobis/CoreMahalanobis.py.
 * create: the constructor L26-40 (same parameters; d = number of features, fixed per handle).
 * One request batch = pass1 -> finalize -> pass2 -> commit, which together are _get_preds (L68-151):
 *   pass1   : L84-100  clip + additive statistics (count, sum, Gram about the start-of-call mean)
 *   finalize: L100-112 covariance update, top-k eigenvectors (Jacobi instead of scipy eigh), projected
 *             start-of-sub-segment inverses; with nranks > 1 the statistics of all row shards (gathered by the
 *             caller, e.g. ncclAllGather through torch.distributed) are combined in rank order
 *   pass2   : L106-136, L147-149 projections, warm-up gate, inverse chain, score and flag per row
 *   commit  : L139-144 state update (mean, C, n, clip)
 * sb_maha_batch runs the four on one GPU. */
typedef struct sb_maha sb_maha_t;

int32_t sb_maha_create(sb_maha_t** h, int32_t d, int32_t n_components, double threshold, double n_stdev,
                       int64_t start_clip, int64_t max_n);
int32_t sb_maha_destroy(sb_maha_t* h);

/* bytes of device workspace needed for a (local) batch of up to max_rows rows */
int32_t sb_maha_workspace_bytes(const sb_maha_t* h, int64_t max_rows, uint64_t* bytes);
/* number of doubles in one rank's statistics record: 1 + d + d*d */
int64_t sb_maha_stats_len(const sb_maha_t* h);

int32_t sb_maha_pass1_stats(sb_maha_t* h, const float* d_X, int64_t rows, int64_t ld, void* d_ws,
                            uint64_t ws_bytes, double* d_stats_out, uintptr_t stream);
/* d_all_stats: nranks records of sb_maha_stats_len() doubles in rank order (NULL with nranks == 1: use this
 * rank's own record); total_rows = rows of the whole request batch, row0 = index of this rank's first row */
int32_t sb_maha_finalize(sb_maha_t* h, const double* d_all_stats, int32_t nranks, int32_t rank, int64_t total_rows,
                         int64_t row0, void* d_ws, uint64_t ws_bytes, uintptr_t stream);
int32_t sb_maha_pass2_score(sb_maha_t* h, const float* d_X, int64_t rows, int64_t ld, float* d_score,
                            int32_t* d_flag, void* d_ws, uint64_t ws_bytes, uintptr_t stream);
int32_t sb_maha_commit(sb_maha_t* h, uintptr_t stream);
/* drop a batch that failed half way (state untouched) */
int32_t sb_maha_abort_batch(sb_maha_t* h);

int32_t sb_maha_batch(sb_maha_t* h, const float* d_X, int64_t rows, int64_t ld, float* d_score, int32_t* d_flag,
                      void* d_ws, uint64_t ws_bytes, uintptr_t stream);

/* Fuses the standardise transformer (CoreVAE.py:128, (X - mu) / (sigma + 1e-10); sb_standardise above) into the
 * detector: with d_mu / d_inv_sigma set (fp32[d] device vectors owned by the caller, 16-byte aligned, kept alive while
 * set; NULL, NULL switches it off) every batch is scored as if sb_standardise had been applied to it first - the same
 * two fp32 roundings per value, so the results agree with the two-step graph to the detector's own run-to-run spread
 * (a few 1e-7 relative in the scores, equal flags) - without the 8*d B/row of the separate pass.  Only on the tensor-core path (d = 64, 128, 256, n_components <= 16, aligned rows);
 * other shapes report SB_ERR_INVALID at sb_maha_pass1_stats. */
int32_t sb_maha_set_input_transform(sb_maha_t* h, const float* d_mu, const float* d_inv_sigma);

/* Pass-1 statistics without any workspace: count, sum and Gram of (x - shift) over `rows` rows of an fp32 device
 * matrix into d_stats_out (1 + d + d*d doubles, the record layout of sb_maha_pass1_stats).  Has no counterpart in
 * the reference (which never leaves the host).  It exists for callers that cannot allocate device memory: bindings
 * with no device allocator of their own (a JNI/cgo shim has no caching allocator to carve a workspace from) and
 * serving threads that must keep cudaMalloc - a device-wide synchronisation - off the request path.  The
 * sub-segment partials live in a static device array inside the library (4 259 840 bytes per GPU, held until exit),
 * hence the limits d <= 64 and rows <= 1 048 576; larger arguments are rejected with SB_ERR_INVALID.  The function
 * calls no allocating routine.  The array is shared by every call on a GPU: nothing needs clearing between calls,
 * but calls that overlap on the same GPU race - keep them on one stream.  The first call on each device resolves the
 * array's device address; make it before starting any stream capture.  d_shift: d floats on the device. */
int32_t sb_gram_stats_noalloc(const float* d_X, int64_t rows, int32_t d, int64_t ld, const float* d_shift,
                              double* d_stats_out, uintptr_t stream);

/* Synchronising accessors (public attributes of the reference class / pickling, persistence.py:84).
 * Any h_* pointer may be NULL.  h_C is d*d row-major; has_state = 0 while mean/C are still the scalar 0 of
 * L35-36; has_clip = 0 while clip is None. */
int32_t sb_maha_get_state(sb_maha_t* h, double* h_mean, double* h_C, double* h_clip_lo, double* h_clip_hi,
                          int64_t* n, int32_t* has_state, int32_t* has_clip, uintptr_t stream);
int32_t sb_maha_set_state(sb_maha_t* h, const double* h_mean, const double* h_C, const double* h_clip_lo,
                          const double* h_clip_hi, int64_t n, int32_t has_state, int32_t has_clip,
                          uintptr_t stream);
/* sums over the last batch (metrics(), L180-192): number of flags set and sum of scores; row where the
 * warm-up gate opened (-1: never) - diagnostics. */
int32_t sb_maha_last_batch_sums(sb_maha_t* h, int64_t* n_flagged, double* score_sum, int64_t* gate_row,
                                uintptr_t stream);
/* with SB_OPT_PROFILE: device time in ms of the last batch's { pass-1 statistics kernel, eigensolver,
 * prefix/start-inverse kernels, pass-2 projection kernel (0 when pass 2 runs fused), pass-2 chain kernel }.
 * Synchronises the stream. */
int32_t sb_maha_kernel_times(sb_maha_t* h, float* out_ms5, uintptr_t stream);
/* diagnostics: out8 = { first open sub-segment, gate row, flagged rows, Jacobi sweeps, batch rows,
 * sub-segments, sub-segment length, subspace-iteration count (> 0: accepted, < 0: not accepted and the Jacobi solver
 * ran, 0: not attempted) } of the last batch.  Synchronises. */
int32_t sb_maha_debug_counters(sb_maha_t* h, int64_t* out8, uintptr_t stream);
/* top-k eigenvectors of the last batch, d x k row-major (diagnostics / subspace parity tests) */
int32_t sb_maha_get_eigvecs(sb_maha_t* h, double* h_V, uintptr_t stream);
/* tuning knobs: SB_OPT_GRAM_TERMS 1|3 (TF32 products per Gram element, default 3),
 * SB_OPT_FORCE_GENERIC 0|1 (skip the d=64 tensor-core kernels), SB_OPT_PROFILE 0|1 (record CUDA events around
 * the main kernels on the caller's stream), SB_OPT_FUSED_PASS2 0|1 (d=64: project and chain in one kernel instead
 * of the default projection kernel + chain kernel), SB_OPT_LEGACY_MMA 0|1 (pass-1 Gram on mma.sync instead of the
 * default tcgen05.mma kernel) */
#define SB_OPT_GRAM_TERMS 1
#define SB_OPT_FORCE_GENERIC 2
#define SB_OPT_PROFILE 3
#define SB_OPT_FUSED_PASS2 4
#define SB_OPT_LEGACY_MMA 5
/* SB_OPT_NO_SUBSPACE 0|1: d = 128 / 256 take the top-k eigenvectors of every batch after the first from a warm-started
 * block subspace iteration that accepts only under a proven spectral gap (csrc/maha_subspace.cu) and otherwise falls
 * back to the Jacobi solver; 1 = always run the Jacobi solver */
#define SB_OPT_NO_SUBSPACE 6
int32_t sb_maha_set_option(sb_maha_t* h, int32_t option, int64_t value);
/* number of kernel launches issued through this handle so far (bench.py's gpu_launches) */
int64_t sb_launch_count(void);

/* ---------------------------------------------------------------------------------------------- routers
 * replaces components/routers/epsilon-greedy/EpsilonGreedy.py (kind 0) and
 * components/routers/thompson-sampling/ThompsonSampling.py (kind 1), batched: T interleaved
 * route()/send_feedback() calls in one launch, decisions bit-identical to the reference under a seed.
 * create = __init__ (L35-85 / L35-76): seed > 0 seeds both MT19937 streams like random.seed/np.random.seed;
 * best_branch <= 0 draws the start branch like random.choice(range(n)). */
typedef struct sb_router sb_router_t;
#define SB_ROUTER_EPSILON_GREEDY 0
#define SB_ROUTER_THOMPSON 1
#define SB_REWARD_PER_CALL 0  /* d_reward[t]                                  */
#define SB_REWARD_TABLE 1     /* d_reward[t*n_branches + arm]                  */
#define SB_REWARD_BERNOULLI 2 /* d_reward[t] = u_t; reward = u_t < d_arm_p[arm] */
int32_t sb_router_create(sb_router_t** h, int32_t kind, int32_t n_branches, double epsilon, int64_t seed,
                         int32_t best_branch);
int32_t sb_router_destroy(sb_router_t* h);
/* d_rows_per_call may be NULL (1 row per call).  route only: reward_mode = -1 (no feedback applied). */
int32_t sb_router_run(sb_router_t* h, int64_t T, const int64_t* d_rows_per_call, const float* d_reward,
                      const float* d_arm_p, int32_t reward_mode, int32_t* d_decisions, uintptr_t stream);
/* route() of the class API (EpsilonGreedy.py:87-106, ThompsonSampling.py:78-91): one decision, written by the kernel
 * into pinned host memory owned by the handle and returned in *h_out.  Synchronises the stream. */
int32_t sb_router_route(sb_router_t* h, int32_t* h_out, uintptr_t stream);
/* send_feedback(routing = arm, rows, reward) of the previous request followed by the next route(), in ONE launch: the
 * request loop of the wrapper (seldon_methods.py:59-103 then L196-243) costs one launch and one synchronisation per
 * request instead of two launches.  Same decisions as sb_router_feedback + sb_router_route. */
int32_t sb_router_feedback_route(sb_router_t* h, int32_t arm, int64_t rows, double reward, int32_t* h_out,
                                 uintptr_t stream);
/* Thompson: the n beta samples of the last route (self.branch_values, ThompsonSampling.py:81-82).  Synchronises. */
int32_t sb_router_last_values(sb_router_t* h, double* h_values, uintptr_t stream);
/* single feedback for an externally chosen arm (send_feedback of the class API) */
int32_t sb_router_feedback(sb_router_t* h, int32_t arm, int64_t rows, double reward, uintptr_t stream);
/* state blob for pickling: success[n], tries[n], best_branch, then both generator states.  Synchronises. */
int64_t sb_router_state_len(const sb_router_t* h);
int32_t sb_router_get_state(sb_router_t* h, int64_t* h_state, uintptr_t stream);
int32_t sb_router_set_state(sb_router_t* h, const int64_t* h_state, uintptr_t stream);
/* out2 = { accept/reject comparisons of the legacy gamma sampler (ThompsonSampling.py:81 -> np.random.beta) that fell
 * inside the libm guard band and were decided again with correctly rounded logarithms, beta draws taken draw by draw
 * because they outran the kernel's look-ahead window } since the handle was created.  Synchronises. */
int32_t sb_router_counters(sb_router_t* h, int64_t* out2, uintptr_t stream);
/* test knob: SB_ROUTER_OPT_DEBUG_FALLBACK_PERIOD p > 0 takes every p-th beta draw through the draw-by-draw fallback
 * (decisions must not change); 0 = off */
#define SB_ROUTER_OPT_DEBUG_FALLBACK_PERIOD 1
int32_t sb_router_set_option(sb_router_t* h, int32_t option, int64_t value);

/* RANDOM_ABTEST, the engine's built-in two-way router: replaces
 * engine/src/main/java/io/seldon/engine/predictors/RandomABTestUnit.java:29-59 - `new Random(1337)` (L31),
 * `rand.nextFloat()` per call (L45), branch 0 iff nextFloat() <= ratioA (L52) - batched: T calls per launch, the
 * java.util.Random stream consumed in parallel by jump-ahead (the generator is an affine map modulo 2^48).
 * d_state2: two device words; word 0 is the generator state (start: sb_abtest_initial_state(1337)), advanced by T
 * calls when the launch completes; word 1 is scratch. */
uint64_t sb_abtest_initial_state(int64_t seed);
int32_t sb_abtest_route(uint64_t* d_state2, float ratio_a, int64_t T, int32_t* d_decisions, uintptr_t stream);

#ifdef __cplusplus
}
#endif
#endif /* SB200_H */
