/* libsmi -- C ABI of the B200-native Stein Mixture Inference / SVGD update step.
 *
 * The reference (aleatory-science/smi_experiments) has no native boundary: its hot path
 * is the Python call `engine.update(state, *args, **kwargs)` (reference
 * src/engine_run.py:25,30; src/model_select/early_stopping.py:42) that runs NumPyro's
 * `SteinVI.update` as one XLA computation.  This header is the boundary a replacement
 * exports instead; every entry point cites the stage of that computation it replaces
 * (SURVEY.md section 8a rows a1-a10).  INTEGRATION.md shows the jax.ffi / ctypes
 * binding a maintainer of the reference would add.
 *
 * Conventions
 *   - return 0 on success, a negative SMI_E* code otherwise; never throws; the message
 *     of the last failure on the calling thread is smi_last_error().
 *   - the caller owns every buffer.  All pointers are DEVICE pointers (row-major fp32 /
 *     int32, 16-byte aligned) unless a parameter is documented as host.  Workspaces are
 *     sized by the *_workspace_bytes query and passed in; nothing is allocated inside.
 *   - no host synchronisation and no default-stream use: work is enqueued on `stream`
 *     (a cudaStream_t passed as void*), so every call is CUDA-graph capturable.
 *   - scalar results (loss, bandwidth, step counter, PRNG key) stay on the device.
 *   - sm_100a only.  There is no CPU fallback.
 */
#ifndef SMI_H_
#define SMI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SMI_VERSION 100

#define SMI_OK 0
#define SMI_EINVAL (-1)   /* bad argument                                  */
#define SMI_ENOSPC (-2)   /* workspace too small                           */
#define SMI_ECUDA (-3)    /* CUDA runtime / driver error                   */
#define SMI_EARCH (-4)    /* device is not sm_100                          */
#define SMI_ELIMIT (-5)   /* shape outside what the kernels support        */

/* jnp.median index arithmetic */
#define SMI_MEDIAN_EXACT_MIDPOINT 0
#define SMI_MEDIAN_JAX_FP32_INDEX 1

/* which Stein-force implementation ran / is requested */
#define SMI_PATH_AUTO 0
#define SMI_PATH_SMALL 1   /* m <= 64: difference-form kernel, shared-memory sort      */
#define SMI_PATH_TC 2      /* any m: tcgen05 3xTF32 Gram + radix select + tcgen05 force */

#define SMI_OPT_NONE 0
#define SMI_OPT_ADAM 1     /* numpyro.optim.Adam    (reference experiments/uci_benchmark.py:191)   */
#define SMI_OPT_ADAGRAD 2  /* numpyro.optim.Adagrad (reference experiments/img_classifier.py:173)  */

#define SMI_MODEL_GAUSS 0    /* reference experiments/variance_collapse.py:70-74 */
#define SMI_MODEL_BNN_REG 1  /* reference src/models/uci_bnn.py:21-68, src/models/reg_1d_bnn.py:22-67 */
#define SMI_MODEL_BNN_CLS 2  /* reference src/models/mlp.py:14-65 */

#define SMI_ACT_RELU 0
#define SMI_ACT_TANH 1

#define SMI_NOISE_GAMMA_PREC 0 /* prec ~ Gamma(1, 0.1), y ~ N(f, prec^-1/2) */
#define SMI_NOISE_FIXED 1      /* y ~ N(f, sigma)                            */

#define SMI_GUIDE_NORMAL 0 /* AutoNormal: psi = sorted(loc, rho) leaves, u = loc + softplus(rho) eps */
#define SMI_GUIDE_DELTA 1  /* AutoDelta : psi = loc leaves, S = 1                                    */

#define SMI_T_IDENTITY 0
#define SMI_T_EXP 1

#define SMI_MAX_SITES 16

/* One latent site of the model (numpyro.sample statement), in model order. */
typedef struct SmiSite {
  int32_t size;      /* number of scalars                                         */
  int32_t transform; /* SMI_T_*: unconstrained -> constrained                     */
  int32_t lat;       /* offset in the latent vector (model order)                 */
  int32_t loc;       /* offset of <site>_auto_loc   inside the raveled particle   */
  int32_t rho;       /* offset of <site>_auto_scale inside the raveled particle; -1 for AutoDelta */
} SmiSite;

/* Model + guide descriptor (host struct, passed by pointer, copied at call time). */
typedef struct SmiModelDesc {
  int32_t kind;    /* SMI_MODEL_*                                                 */
  int32_t guide;   /* SMI_GUIDE_*                                                 */
  int32_t in_dim;  /* features (Gaussian target: dimension)                       */
  int32_t hidden;  /* hidden width (two hidden layers)                            */
  int32_t out_dim; /* 1 for regression, classes for classification               */
  int32_t act;     /* SMI_ACT_*                                                   */
  int32_t noise;   /* SMI_NOISE_* (regression)                                    */
  float sigma;     /* fixed observation noise (SMI_NOISE_FIXED)                   */
  int32_t n_sites;
  int32_t latent_dim;   /* L = sum of site sizes                                  */
  int32_t particle_dim; /* d = 2L (AutoNormal) or L (AutoDelta)                   */
  SmiSite sites[SMI_MAX_SITES];
} SmiModelDesc;

int smi_version(void);
const char* smi_last_error(void);
/* 0 when the current device is compute capability 10.x, SMI_EARCH otherwise. */
int smi_device_check(void);
/* Number of kernels libsmi has launched from the calling thread since the last reset (launches
 * replayed by a CUDA graph are not seen here: count one eager step).  reset != 0 zeroes it.   */
long long smi_launch_count(int reset);
/* Per-kernel timing for the roofline report: while enabled, CUDA events are recorded on the
 * launching stream around the hot kernels.  smi_profile(1) clears and starts, smi_profile(0)
 * stops; smi_profile_report synchronises the device and writes one "name count total_ms" line
 * per kernel into buf (host).  Not for use during CUDA-graph capture.                        */
int smi_profile(int on);
int smi_profile_report(char* buf, size_t cap);

/* ---- Host-buffer staging: the reference hands engine.update host (NumPy) arrays every step
 *      (src/engine_run.py:52, experiments/uci_benchmark.py:228); jax.device_put stages them.
 *   smi_stage_create : two pinned host buffers of `nbytes` used alternately, each guarded by the
 *                      CUDA event of its last copy (the host may run ahead of the stream).
 *   smi_stage_h2d    : memcpy src -> pinned, cudaMemcpyAsync pinned -> dst on `stream`; returns
 *                      without waiting for the copy.  nbytes <= the handle's capacity.
 *   Not for use during CUDA-graph capture.                                                    */
int smi_stage_create(size_t nbytes, void** handle);
int smi_stage_h2d(void* handle, const void* src, size_t nbytes, void* dst, void* stream);
/* `count` host arrays in ONE copy: array i goes to dst + off_i, off_0 = 0 and
 * off_{i+1} = off_i + nbytes[i] rounded up to 256 (sum <= the handle's capacity).             */
int smi_stage_h2dv(void* handle, const void* const* srcs, const size_t* nbytes, int count, void* dst, void* stream);
int smi_stage_destroy(void* handle);
/* Widening of staged raw arguments on the device: the reference hands its classifier uint8 pixels and int64 labels
 * (datasets/ood_detection/load_ood_detect.py:39-46, experiments/img_classifier.py:66-73), which jnp converts inside
 * the jitted update; here the raw bytes travel H2D and these kernels widen them (no torch op on the step's path). */
int smi_convert_u8_f32(const unsigned char* in, size_t n, float* out, void* stream);
int smi_convert_i64_i32(const long long* in, size_t n, int* out, void* stream);

/* ---- Stein force: replaces RBFKernel.compute + _apply_kernel + _kernel_grad + force sum
 *      + ||F|| (NumPyro steinvi.py::_svgd_loss_and_grads).
 *   X, G : (m, d) particles and per-particle ELBO gradients.
 *   F    : (m, d) out (may be NULL when an optimizer is fused and F is not wanted).
 *   loss : out, ||F||_F.   h : out, the bandwidth used.
 *   h_fixed > 0 skips the median heuristic and uses that bandwidth.
 *   F = c_att * K G + (2 alpha / (h m)) (rowsum(K) * X - K X),  K_ij = exp(-|x_i-x_j|^2 / h),
 *   h = median(D)^2 / ln m + 1.1920929e-7 (m == 1: h = 1).                               */
size_t smi_stein_force_workspace_bytes(int m, int d, int path);
int smi_stein_force_f32(const float* X, const float* G, int m, int d, float c_att, float alpha,
                        int median_mode, float h_fixed, int path, float* F, float* loss, float* h,
                        void* ws, size_t ws_bytes, void* stream);

/* Stages of the above, separately testable.
 *   smi_sqdist_f32          : D2 (m, m) = max(|x_i|^2 + |x_j|^2 - 2 x_i.x_j, 0)  (Gram form, row a4)
 *   smi_median_bandwidth_f32: h from D2 (exact order statistics of sqrt(D2), row a4)
 *   rank_out (optional, 2 x uint32 device): the two selected D2 bit patterns (lo, hi).   */
size_t smi_sqdist_workspace_bytes(int m, int d, int path);
int smi_sqdist_f32(const float* X, int m, int d, int path, float* D2, void* ws, size_t ws_bytes,
                   void* stream);
size_t smi_median_bandwidth_workspace_bytes(int m);
int smi_median_bandwidth_f32(const float* D2, int m, int median_mode, float* h, uint32_t* rank_out,
                             void* ws, size_t ws_bytes, void* stream);

/* No-allocation variant of smi_median_bandwidth_f32.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops,
 * harnesses that forbid it).  Same inputs / outputs, no workspace argument: the 8.7 KB of select scratch live in a
 * file-scope device array of libsmi.so (16 KB per GPU that loads the library, held until exit), resolved once per
 * device with cudaGetSymbolAddress.  Limit: m <= SMI_NOALLOC_MAX_M, larger m returns SMI_ELIMIT with a message.
 * Calls no cudaMalloc / cudaFree / allocating routine.  All calls on a device share the one array (it keeps the
 * previous call's contents; what must start at zero is cleared on the stream by every call): calls that OVERLAP on
 * the same GPU race -- serialise them on one stream.                                                           */
#define SMI_NOALLOC_MAX_M 32768
int smi_median_bandwidth_noalloc_max_m(void);
int smi_median_bandwidth_noalloc_f32(const float* D2, int m, int median_mode, float* h, uint32_t* rank_out,
                                     void* stream);

/* ---- Row-sharded Stein force (no counterpart in the single-device reference).
 *   One process per GPU owns particle rows [row0, row0 + rows).  X and G hold ALL m particles
 *   (the caller all-gathers them over NCCL).  Stages, with the caller's collectives between:
 *     smi_stein_rows_dist
 *     for pass in 0..2: smi_stein_rows_hist -> all-reduce SUM hist[2048] -> smi_stein_rows_scan
 *     smi_stein_rows_tail -> all-reduce SUM cnt_min[0], MIN cnt_min[1]
 *     smi_stein_rows_bandwidth -> smi_stein_rows_force -> all-reduce SUM sumsq; loss = sqrt(sumsq)
 *   `hist` is a caller-owned 2048 x uint32 buffer, zero before pass 0.  Integer reductions are
 *   exact, so h is bit-identical for every world size.  ws is private to the rank and must be
 *   the same buffer for every stage of one step.                                           */
size_t smi_stein_rows_workspace_bytes(int rows, int m, int d);
int smi_stein_rows_dist(const float* X, int m, int d, int row0, int rows, void* ws, size_t ws_bytes,
                        void* stream);
int smi_stein_rows_hist(int m, int d, int row0, int rows, int pass, uint32_t* hist, void* ws,
                        size_t ws_bytes, void* stream);
int smi_stein_rows_scan(int m, int d, int row0, int rows, int pass, int median_mode, uint32_t* hist,
                        void* ws, size_t ws_bytes, void* stream);
int smi_stein_rows_tail(int m, int d, int row0, int rows, uint32_t* cnt_min, void* ws, size_t ws_bytes,
                        void* stream);
int smi_stein_rows_bandwidth(int m, int d, int row0, int rows, int median_mode, float alpha,
                             float h_fixed, const uint32_t* cnt_le, const uint32_t* min_gt, float* h,
                             void* ws, size_t ws_bytes, void* stream);
int smi_stein_rows_force(const float* X, const float* G, int m, int d, int row0, int rows, float c_att,
                         float* F_rows, double* sumsq, void* ws, size_t ws_bytes, void* stream);

/* ---- Block-symmetric variant of the row-sharded distance stage ("must overlap"; no counterpart in
 *   the single-device reference).  The world is a ring of row blocks: rank i computes its own diagonal block
 *   (upper tiles + mirror) and the `right` = world / 2 blocks after it, and receives the blocks to its left as the
 *   transposes of what the ranks before it computed, so every unordered block pair is computed once, as on one GPU.
 *   Only the particles of half the ring are needed before the Gram; the rest may arrive until smi_stein_rows_force.
 *     smi_stein_sym_colsum -> all-gather `part` (64 x d doubles, slab-ordered) -> smi_stein_sym_mean
 *     smi_stein_sym_split(own rows), smi_stein_sym_split(each right block as it arrives) -> smi_stein_sym_gram
 *     per right block o < world / 2 (and o = world / 2 for odd worlds): smi_stein_sym_pack_t -> send to rank i + o;
 *     per left block received from rank i - o: smi_stein_sym_unpack
 *     then smi_stein_rows_hist / scan / tail / bandwidth / force exactly as above (same workspace).
 *   h keeps the bits of the single-GPU path (slab-ordered mean, same accumulation sequence per particle pair).
 *   smi_stein_sym_supported: rows % 256 == 0, rows a multiple of m / 64, no split-K shape; else use the API above. */
int smi_stein_sym_supported(int m, int d, int rows);
int smi_stein_sym_colsum(const float* X, int m, int d, int row0, int rows, double* part, void* ws, size_t ws_bytes,
                         void* stream);
int smi_stein_sym_mean(const double* part, int m, int d, int row0, int rows, void* ws, size_t ws_bytes, void* stream);
int smi_stein_sym_split(const float* X, int m, int d, int row0, int rows, int r0, int nr, void* ws, size_t ws_bytes,
                        void* stream);
int smi_stein_sym_gram(int m, int d, int row0, int rows, int first, int count, void* ws, size_t ws_bytes,
                       void* stream);   /* column blocks [first, first + count) after the own one, ring order */
int smi_stein_sym_pack_t(int m, int d, int row0, int rows, int col0, float* out, void* ws, size_t ws_bytes,
                         void* stream);
int smi_stein_sym_unpack(const float* in, int m, int d, int row0, int rows, int col0, void* ws, size_t ws_bytes,
                         void* stream);

/* ---- Optimizers (row a8).  `grad_sign` multiplies F before use: the Stein step hands the
 *      optimizer -F, so pass -1.  `step` is a device int32 holding the number of updates
 *      applied so far; it is read, and incremented by one when `bump_step` is non-zero.
 *      Hyper-parameters are doubles: (1 - b1), (1 - b2), (1 - momentum) are formed in double and
 *      then rounded to fp32, as JAX does with weakly-typed Python scalars.                  */
int smi_adam_f32(float* x, float* mom, float* vel, const float* F, float grad_sign, int32_t* step,
                 int bump_step, double lr, double b1, double b2, double eps, size_t n, void* stream);
int smi_adagrad_f32(float* x, float* acc, float* mom, const float* F, float grad_sign,
                    int32_t* step, int bump_step, double lr, double momentum, size_t n, void* stream);

/* Gated forms used inside run loops: identical arithmetic, but the whole update is skipped when
 * *skip != 0 (skip may be NULL), and the step counter is NOT bumped here -- smi_step_commit_f32
 * closes the step.                                                                          */
int smi_adam_gated_f32(float* x, float* mom, float* vel, const float* F, float grad_sign,
                       const int32_t* step, double lr, double b1, double b2, double eps, size_t n,
                       const int32_t* skip, void* stream);
int smi_adagrad_gated_f32(float* x, float* acc, float* mom, const float* F, float grad_sign, double lr,
                          double momentum, size_t n, const int32_t* skip, void* stream);

/* ---- Per-particle mixture-ELBO gradients (row a3; NumPyro stein_loss.py).
 *   psi  : (m, d) all particles.   eps : (nrows, S, L) N(0,1) noise of the owned rows (NULL for
 *   AutoDelta).  idx : (nrows, S, B) int32 subsample rows into x / y, or NULL = every draw uses
 *   rows 0..B-1 of x / y.
 *   x    : (N, in_dim) fp32.  y : (N) fp32 targets, or int32 labels for SMI_MODEL_BNN_CLS.
 *   lik_scale : plate scaling N/B.   tau : loss_temperature.
 *   rows : particles [row0, row0 + nrows) are differentiated (particle-row sharding, row a3/e).
 *   G    : (nrows, d) out = d/d psi_i mean_s ELBO_is.   elbo : (nrows) out.                */
size_t smi_elbo_grad_workspace_bytes(const SmiModelDesc* md, int m, int S, int B, int nrows);
int smi_elbo_grad_f32(const SmiModelDesc* md, const float* psi, const float* eps,
                      const int32_t* idx, const float* x, const void* y, int m, int S, int B,
                      float lik_scale, float tau, int row0, int nrows, float* G, float* elbo,
                      void* ws, size_t ws_bytes, void* stream);

/* As smi_elbo_grad_f32 with the loss temperature read from device memory at execution time (so a
 * captured CUDA graph follows a schedule), and the ASVGD schedule itself (numpyro
 * ASVGD._cyclical_annealing; reference src/methods/asvgd.py:4-27, src/engine_run.py:16-24):
 *   *tau = cyc(t) / m,  t = *step - step0,
 *   cyc(t) = (1 - last) (((t - 1) mod cycle_len + 1) / norm)^trans_speed + last,
 *   norm = (num_steps + 1) / num_cycles, cycle_len = num_steps / num_cycles (integer),
 *   last = min(t / ((num_cycles - 1) cycle_len), 1) (integer division; 1 when num_cycles = 1).      */
int smi_elbo_grad_tau_f32(const SmiModelDesc* md, const float* psi, const float* eps,
                          const int32_t* idx, const float* x, const void* y, int m, int S, int B,
                          float lik_scale, const float* tau_dev, int row0, int nrows, float* G,
                          float* elbo, void* ws, size_t ws_bytes, void* stream);
int smi_asvgd_tau_f32(const int32_t* step, int step0, int num_steps, int num_cycles, int trans_speed,
                      int m, float* tau, void* stream);

/* ---- Counter PRNG (row a9): threefry2x32 key chain of one update on the device.
 *   key_in  : device uint32[2] state key.  key_out : device uint32[2] next state key (must not
 *   alias key_in; may be NULL).  n_total > 0 declares a data plate (it consumes one subkey even
 *   when idx is NULL).
 *   eps     : (nrows, S, L) out (NULL for AutoDelta).  idx : (nrows, S, B) out or NULL.
 *   Chain (partitionable threefry layout): split(key,3) ->
 *   step key -> attractive key -> per particle -> per draw -> guide key -> one subkey per
 *   RNG-consuming statement (data plate first, then latent sites in model order).          */
int smi_prng_step_f32(const SmiModelDesc* md, const uint32_t* key_in, uint32_t* key_out, int m,
                      int S, int n_total, int B, int row0, int nrows, float* eps, int32_t* idx,
                      void* stream);

/* Threefry stream layout of every PRNG entry below and of the predictive draws: JAX flipped the default of
 * jax_threefry_partitionable in 0.5.0 and the reference leaves JAX unpinned (requirements.txt), so both exist.
 * Process-wide; returns the previous layout; any other argument only queries.  Default: partitionable.  */
#define SMI_PRNG_PARTITIONABLE 1
#define SMI_PRNG_LEGACY 0
int smi_prng_layout(int layout);
/* Rule of the data-plate subsample (numpyro/primitives.py::_subsample_fn).
 *   CPU backend of the reference: partial Fisher-Yates, j_t = random.randint(split(key, B)[t], (), 0, n - t):
 *     MULSHIFT (default): one 32-bit word, (bits * span) >> 32.
 *     JAX: jax.random.randint's construction -- two words from split(key), ((hi % span) * (2^32 % span) +
 *          lo % span) % span in wrapping uint32 arithmetic.
 *   Any other backend of the reference: CHOICE = jax.random.choice(key, n, (B,), replace=False) =
 *     permutation(key, n)[:B], rounds of a stable sort by fresh 32-bit words (n <= 16384 here).
 * Process-wide; returns the previous rule; any other argument queries.                                       */
#define SMI_RANDINT_MULSHIFT 0
#define SMI_RANDINT_JAX 1
#define SMI_RANDINT_CHOICE 2
int smi_prng_randint(int mode);

/* jax.random primitives on the device (layout per smi_prng_layout), used by engine.init
 * (numpyro init_to_uniform / guide initialisation) and exposed for known-answer tests.
 *   smi_prng_split      : keys_out (num, 2) = jax.random.split(key, num)
 *   smi_prng_uniform_f32: out (n) = jax.random.uniform(key, (n,), minval=lo, maxval=hi)
 *   smi_prng_normal_f32 : out (n) = jax.random.normal(key, (n,))
 *   smi_prng_bits_u32   : out (n) = jax.random.bits(key, (n,), uint32)                      */
int smi_prng_split(const uint32_t* key, int num, uint32_t* keys_out, void* stream);
int smi_prng_uniform_f32(const uint32_t* key, size_t n, float lo, float hi, float* out, void* stream);
int smi_prng_normal_f32(const uint32_t* key, size_t n, float* out, void* stream);
int smi_prng_bits_u32(const uint32_t* key, size_t n, uint32_t* out, void* stream);

/* ---- Loss collection without a host sync (SteinVI.run's fori_collect of the per-step loss;
 *      replaces the per-step device_get at reference src/engine_run.py:56).
 *   ring[(step - 1) mod ring_len] = *loss, `step` being the device counter the optimizer
 *   bumped for this update.                                                               */
int smi_loss_record_f32(const float* loss, const int32_t* step, float* ring, int ring_len, void* stream);

/* ---- End of one update inside a device-side run loop: replaces the
 *      loop-carried bookkeeping of SteinVI.run / fori_collect and of OnlineEarlyStop.step
 *      (reference src/model_select/early_stopping.py:24-62), with no host round trip.
 *   If skip != NULL and *skip != 0: nothing happens (the run has stopped; the gated optimizer
 *   left the particles alone).  Otherwise, with t = *step:
 *     *step = t + 1;  ring[t mod ring_len] = *loss (ring may be NULL);
 *     key <- key_next (both uint32[2], may both be NULL);
 *     es (may be NULL) = {t_es, fast_avg, slow_avg} float32:
 *        tau = min(t_es, period_x), avg_x = loss / (tau + 1) + (tau / (tau + 1)) * decay_x * avg_x
 *        for x in {fast (period), slow (10 * period)};  t_es += 1;
 *        *skip = 1 when fast_avg > 2 * slow_avg.
 *   fast_decay = exp(log(1e-3) / period), slow_decay = exp(log(1e-3) / (10 period)), evaluated
 *   by the caller in float32 as the reference does.                                        */
int smi_step_commit_f32(const float* loss, int32_t* step, float* ring, int ring_len,
                        const uint32_t* key_next, uint32_t* key, float* es, int32_t* skip,
                        float fast_decay, float slow_decay, int period, void* stream);

/* ---- Posterior predictive + regression metrics: the step AFTER the path.
 *      Replaces MixtureGuidePredictive / Predictive as used by the reference's samplers
 *      (src/metrics.py:216-237, experiments/uci_benchmark.py:297-316) for the regression BNNs,
 *      and rmse_fn / nll_reg_fn / lppd_fn (src/metrics.py:222-268).
 *   psi (m, d) particles; key uint32[2]; x (n, in_dim); y (n) observed targets or NULL.
 *   AutoNormal: draw s picks particle a_s uniformly and samples its Normal; AutoDelta: draw s is
 *   particle s mod m (Predictive(num_samples=1, batch_ndims=1)).
 *   Outputs (row-major, any may be NULL except y_loc): assign (S) int32; latents (S, L) constrained,
 *   model order; y_loc (S, n); y_draw (S, n) ~ N(y_loc, sigma_s); loglik (S, n) (needs y).   */
int smi_predictive_reg_f32(const SmiModelDesc* md, const float* psi, int m, const uint32_t* key,
                           int num_samples, const float* x, int n, const float* y, int32_t* assign,
                           float* latents, float* y_loc, float* y_draw, float* loglik, void* stream);
size_t smi_metric_workspace_bytes(int num_samples, int n);
/* out[0] = sqrt(mean_{s,i} (y_i - y_loc[s][i])^2)                                            */
int smi_metric_rmse_f32(const float* y_loc, const float* y, int num_samples, int n, float* out, void* ws,
                        size_t ws_bytes, void* stream);
/* lpd_i = logsumexp_s loglik[s][i] - log S (written to lpd (n) when not NULL);
 * out2[0] = -mean_i lpd_i (nll_reg_fn), out2[1] = sum_i lpd_i (lppd_fn)                      */
int smi_metric_nll_f32(const float* loglik, int num_samples, int n, float* out2, float* lpd, void* ws,
                       size_t ws_bytes, void* stream);

/* Classification metrics of the reference (src/metrics.py:17-160, 270-321: nll_fn, acc_fn, brier_fn, ece_fn,
 * mce_fn, conf_fn) per posterior draw.  logits (S, n, C) fp32; y (n) true labels; ypred (S, n) the sampled
 * "y" site or NULL (then acc and conf are NaN); bins <= 128 (the reference uses 100).
 * out6 (S, 6) = {nll, acc, brier, ece, mce, conf}.  ece / mce restate binned_conf_acc including its
 * treatment of empty bins.                                                                      */
int smi_metric_classification_f32(const float* logits, const int32_t* y, const int32_t* ypred, int num_samples,
                                  int n, int num_classes, int bins, float* out6, void* stream);

/* Posterior-predictive draws for the classifier BNN (reference src/models/mlp.py:14-65 under
 * MixtureGuidePredictive / Predictive, src/metrics.py:216-237; experiments/img_classifier.py:126-150):
 *   smi_predictive_latents_f32 : latents (S, L) constrained draws of any model + assign (S), the guide half
 *                                (same key chain as smi_predictive_reg_f32);
 *   smi_predictive_cls_layer1  : h1 (S, n, H) = act(x W1_s + b1_s) as ONE stacked tcgen05 3xTF32 GEMM
 *                                (M = n data rows, N = S x H, K = in_dim);
 *   smi_predictive_cls_head    : logits (S, n, O) and the Categorical draw y (S, n) int32 (may be NULL),
 *                                y = argmax_o(logits + Gumbel noise) as jax.random.categorical.           */
int smi_predictive_latents_f32(const SmiModelDesc* md, const float* psi, int m, const uint32_t* key,
                               int num_samples, int32_t* assign, float* latents, void* stream);
size_t smi_predictive_cls_layer1_workspace_bytes(const SmiModelDesc* md, int num_samples, int n);
int smi_predictive_cls_layer1_f32(const SmiModelDesc* md, const float* latents, int num_samples, const float* x,
                                  int n, float* h1, void* ws, size_t ws_bytes, void* stream);
int smi_predictive_cls_head_f32(const SmiModelDesc* md, const float* latents, const float* h1,
                                const uint32_t* key, int num_samples, int n, float* logits, int32_t* y_draw,
                                void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SMI_H_ */
