/* dlas_b200.h -- C ABI of libdlas_b200.so: the B200 (sm_100a) data-parallel hot path of
 * OmarZOS/deep-learning-at-scale (mini-batch training / inference of the flow regressors and the
 * Gilbert physical baseline).
 *
 * The reference has NO native interface: its boundary is the Keras call sequence in
 * "Artificial intelligence models/Static neural network models/cnn.py" (cited below as cnn.py).
 * Each entry point names the reference lines it stands in for.  All pointers are DEVICE pointers
 * unless a name ends in _host; `stream` is a cudaStream_t passed as void* (NULL = default
 * stream).  Functions are stream-ordered and asynchronous; none of them allocates device memory
 * (scratch is passed in by the caller).  Return 0 on success, a negative DLAS_ERR_* otherwise;
 * the message is available from dlas_last_error() (thread-local).  There is no CPU fallback: on a
 * device that is not compute capability 10.x every compute entry point returns
 * DLAS_ERR_UNSUPPORTED_DEVICE.
 */
#ifndef DLAS_B200_H
#define DLAS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLAS_OK 0
#define DLAS_ERR_INVALID (-1)
#define DLAS_ERR_CUDA (-2)
#define DLAS_ERR_UNSUPPORTED_DEVICE (-3)
#define DLAS_ERR_WORKSPACE (-4)
#define DLAS_ERR_UNSUPPORTED_SHAPE (-5)
#define DLAS_ERR_TIMEOUT (-6)   /* a bounded wait inside a kernel expired (dlas_mlp_status)                 */
#define DLAS_ERR_RANGE (-7)     /* fp16 operand range of DLAS_MODE_TC exceeded / non-finite gradient sums   */

/* losses: cnn.py:29-32 (mae_clip); MSE from BASELINE.json config 1 */
#define DLAS_LOSS_MSE 0
#define DLAS_LOSS_MAE_CLIP 1
/* optimisers: cnn.py:117 (SGD lr/momentum/decay/nesterov); Adam from BASELINE.json north_star */
#define DLAS_OPT_SGD 0
#define DLAS_OPT_ADAM 1
/* arithmetic mode of the MLP kernels */
#define DLAS_MODE_FP32 0   /* CUDA-core fp32, any layer widths                                   */
#define DLAS_MODE_TC 1     /* tcgen05 tensor cores: fp16 operands, fp32 accumulation in TMEM; fixed shapes,
                              see dlas_mlp_mode_supported (32-256-256-1; d0-64-64-1 with d0 <= 15;
                              d0-256-256-1 with 64 <= d0 <= 240, d0 % 8 == 0: at most 65,536 rows per training
                              call; inference through dlas_mlp_predict_ws).  A REDUCED-RANGE mode: rows, hidden activations
                              and back-propagated terms are rounded to fp16 -- TF32's 11-bit significand but
                              a 5-bit exponent: |value| <= 65504, full precision down to 6.1e-5, flushed to
                              zero below 6e-8.  Feed it standardised columns (raw wellhead pressures of a
                              few thousand psig fit; cumulative volumes or timestamps do not).  Overflow is
                              detected, never silent: dlas_mlp_status returns DLAS_ERR_RANGE.            */
/* Gilbert kernel variants (same results within the stated 1e-6; differ in instruction mix) */
#define DLAS_GILBERT_POLY 0  /* polynomial log2 for all three inputs                              */
#define DLAS_GILBERT_MUFU 7  /* bit i set: input i (P,S,R) takes its mantissa log2 from MUFU.LG2  */

typedef struct dlas_opt_cfg {
  int32_t kind;      /* DLAS_OPT_*                                  */
  float lr;          /* cnn.py:117 lr=0.001                          */
  float momentum;    /* cnn.py:117 momentum=0.99                     */
  float decay;       /* cnn.py:117 decay=1e-6 (lr_t = lr/(1+decay*it)) */
  int32_t nesterov;  /* cnn.py:117 nesterov=True                     */
  float beta1, beta2, eps; /* Adam */
} dlas_opt_cfg;

/* ---- library ---------------------------------------------------------------------------- */
int dlas_version(void);
const char* dlas_last_error(void);
/* 0 if the current device is a B200-class (cc 10.x) GPU, DLAS_ERR_UNSUPPORTED_DEVICE otherwise. */
int dlas_device_check(void);
/* number of SMs of the current device (grid sizing is a multiple of it) */
int dlas_sm_count(void);

/* Kernel timing for the benchmark (no reference counterpart; cnn.py:126-129 only takes wall-clock
 * time around fit).  While enabled, the dominant kernel of each entry point is bracketed by CUDA
 * events on its own stream.  Slots: 0 = Gilbert kernel, 1 = MLP forward/backward kernel.
 * dlas_prof_read synchronises, returns the summed milliseconds and the launch count of a slot
 * (host pointers) and clears it. */
#define DLAS_PROF_GILBERT 0
#define DLAS_PROF_MLP_GRAD 1
int dlas_prof_enable(int on);
int dlas_prof_read(int slot, float* ms_sum_host, int* count_host);
/* Average SM clock (MHz) over the launches of the tensor-core MLP kernel since the last reset, measured by the kernel
 * itself (sum of clock64 deltas / sum of globaltimer deltas of one CTA); 0 if it has not run.  Synchronises the
 * device.  Lets the benchmark report the clock under load without NVML calls inside the timed region. */
int dlas_prof_sm_clock(double* mhz_host, int reset);

/* ---- Gilbert's equation: q = P*S^1.89/(435*R^0.546) ---------------------------------------
 * Replaces the "Physical model" script named at Readme.md:7-8 (no code in the snapshot; formula from
 * BASELINE.json).  SoA fp32 arrays of n rows; 16 algorithmic bytes per row.  <=1e-6 relative to
 * the fp64 formula on the fp32 inputs for positive finite normal inputs; P==0 gives 0;
 * negative, NaN, infinite or subnormal inputs take an IEEE powf slow path. */
int dlas_gilbert_f32(const float* P, const float* S, const float* R, float* q, size_t n,
                     int variant, void* stream);

/* ---- synthetic inputs (stand in for the Spark CSV ingest cnn.py:49-103) ------ */
int dlas_synth_gilbert_f32(float* P, float* S, float* R, uint64_t seed, uint64_t start, size_t n,
                           void* stream);
int dlas_synth_rows_f32(float* X, float* y, uint64_t seed, uint64_t start, size_t n_rows,
                        int n_features, void* stream);
int dlas_synth_series_f32(float* X, float* y, uint64_t seed, uint64_t start, size_t n_seq, int T,
                          int d_in, void* stream);
/* cnn.py:68 randomSplit([0.64,0.16,0.2]): code[i] = 0 train / 1 validation / 2 test for row start+i */
int dlas_split_codes(int8_t* code, uint64_t seed, uint64_t start, size_t n, void* stream);
/* dynamic ANN (Readme.md:19): out[i] = rows i..i+n_lags-1 of series (n_rows, n_features) flattened
 * lag-major; writes n_rows-n_lags+1 windows. */
int dlas_lag_windows_f32(const float* series, size_t n_rows, int n_features, int n_lags, float* out,
                         void* stream);

/* ---- MLP (static / dynamic ANN): Sequential of Dense layers, cnn.py:110-118 ---------------
 * dims[0..n_layers] are the layer widths (n_layers Dense layers); hidden ReLU, linear output.
 * Flat fp32 parameter vector: per layer W[in][out] row-major then b[out]. */
size_t dlas_mlp_n_params(const int* dims, int n_layers);
int dlas_mlp_mode_supported(const int* dims, int n_layers, int mode);
/* bytes of device scratch dlas_mlp_grad / dlas_mlp_fit_steps need for these shapes and mode */
size_t dlas_mlp_workspace_bytes(const int* dims, int n_layers, int mode);

/* Forward + loss + backward over n_rows rows: the body of one `model.fit` batch, cnn.py:127-128,
 * without the optimiser update.  Writes grad_sum[0..n_params) = SUM over rows of per-row
 * gradients (no 1/batch factor) and grad_sum[n_params] = SUM of per-row losses, so that
 * data-parallel shards can be added before the step.  Deterministic reduction order. */
/* DLAS_MODE_TC only: the launch may start while the preceding kernel of the stream is still draining
 * (programmatic dependent launch); X, y and params are read only after that kernel has completed, so a
 * producer kernel enqueued just before this call is safe.  (The library's own multi-step loops load
 * the next batch early; they know their predecessor.) */
int dlas_mlp_grad(const int* dims, int n_layers, int loss, int mode, const float* params,
                  const float* X, const float* y, size_t n_rows, float* grad_sum, void* workspace,
                  size_t workspace_bytes, void* stream);

/* Completion and health of the steps enqueued so far on `stream` with this workspace (no reference counterpart: Keras
 * raises from inside fit).  Synchronises the stream, then returns DLAS_OK, DLAS_ERR_TIMEOUT (a bounded wait inside the
 * kernels expired; that step's results are NaN) or DLAS_ERR_RANGE (DLAS_MODE_TC: a gradient sum was not finite -- an
 * fp16 operand overflowed, see DLAS_MODE_TC, or the training diverged).  DLAS_MODE_FP32 only synchronises. */
int dlas_mlp_status(const int* dims, int n_layers, int mode, const void* workspace, size_t workspace_bytes,
                    void* stream);

/* Optimiser update (cnn.py:117 SGD; Adam): g = grad_sum*grad_scale; `iteration` counts batches
 * already applied (Keras `iterations`).  m is the SGD velocity / Adam first moment, v the Adam
 * second moment (ignored for SGD).  If loss_out != NULL writes *loss_out = grad_sum[n]*grad_scale. */
int dlas_opt_step(const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* grad_sum,
                  size_t n, float grad_scale, uint64_t iteration, float* loss_out, void* stream);

/* Replayable steps (no reference counterpart: cnn.py:127-128 drives every batch from the host).  The batch
 * cursor and the optimiser's iteration counter live in DEVICE memory, so that one step
 *   dlas_mlp_grad_at -> [all-reduce of grad_sum] -> dlas_opt_step_at -> dlas_step_advance
 * can be captured once in a CUDA graph and replayed with no host work per step.
 *   dlas_mlp_grad_at  : dlas_mlp_grad on rows [cursor*batch_rows, (cursor+1)*batch_rows) of X_base / y_base; with
 *                       parity_stride > 0 the sums go to grad_sum + (iteration & 1) * parity_stride (the
 *                       double-buffered bucket dlas_dp_allreduce_update reads from its peers)
 *   dlas_opt_step_at  : dlas_opt_step with the learning-rate schedule / Adam bias taken from state->iteration;
 *                       the batch-mean loss goes to loss_ring[iteration % ring_len] (may be NULL)
 *   dlas_step_advance : cursor = (cursor + 1) % n_batches; iteration += 1 */
typedef struct dlas_step_state {
  int64_t cursor;
  int64_t iteration;
} dlas_step_state;
int dlas_mlp_grad_at(const int* dims, int n_layers, int loss, int mode, const float* params,
                     const float* X_base, const float* y_base, size_t batch_rows,
                     const dlas_step_state* state_dev, float* grad_sum, size_t parity_stride,
                     void* workspace, size_t workspace_bytes, void* stream);
int dlas_opt_step_at(const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* grad_sum,
                     size_t n, float grad_scale, const dlas_step_state* state_dev, float* loss_ring,
                     int64_t ring_len, void* stream);
int dlas_step_advance(dlas_step_state* state_dev, int64_t n_batches, void* stream);

/* Gradient all-reduce FUSED with the optimiser update over NVLink peer memory, in one kernel: what BASELINE.json
 * calls "the Spark driver's gradient/parameter averaging" (the snapshot has no code for it).  Every rank owns a
 * bucket of the same layout in memory its peers can load from (CUDA IPC / symmetric memory; the caller maps it and
 * passes the `world` base addresses in device array peer_bases_dev):
 *     floats [0, n+1)                              gradient sums + loss sum, even iterations
 *     floats [parity_stride, parity_stride+n+1)    the same, odd iterations (double buffer)
 *     uint32 [flags_offset, flags_offset+world)    arrival flags, zero-initialised by the caller, one per source rank
 * The kernel announces "my bucket of iteration it is complete" to every peer (release store of it+1 into
 * flag[rank] of each peer), waits until all `world` flags of its own bucket have reached it+1, then every rank adds
 * the `world` buckets in rank order (so all ranks compute bit-identical sums) and applies dlas_opt_step_at's update
 * with g = sum * grad_scale.  No host work, no NCCL call; replayable in a CUDA graph.  The stream order must be
 * dlas_mlp_grad_at(..., grad_sum = own bucket, parity_stride) -> this call -> dlas_step_advance.  A rank that does
 * not see its peers within a bounded spin writes NaN into its parameters and the loss ring instead of hanging. */
int dlas_dp_allreduce_update(const dlas_opt_cfg* cfg, float* params, float* m, float* v,
                             const uint64_t* peer_bases_dev, int rank, int world, size_t n,
                             size_t parity_stride, size_t flags_offset, float grad_scale,
                             const dlas_step_state* state_dev, float* loss_ring, int64_t ring_len,
                             void* stream);

/* n_steps whole data-parallel steps of DLAS_MODE_TC, TWO launches each (no reference counterpart; same averaging semantics as
 * dlas_dp_allreduce_update): the forward/backward kernel on rows [cursor*batch_rows, ...) of this rank's shard, then ONE
 * kernel that sums the per-CTA partials, PUSHES the sums into every peer's receive buffer as 4-byte floats (all-zero bits
 * mean "not arrived": +0.0 travels as -0.0, the reader clears what it has consumed), polls its own receive buffer until the
 * peers' words have arrived, adds
 * the `world` values in rank order (bit-identical on every rank), applies the optimiser update with
 * g = sum * grad_scale, writes the batch-mean loss to loss_ring[iteration % ring_len] and advances the state
 * (cursor = (cursor + 1) % n_batches, iteration += 1).  One NVLink one-way latency per step; no flags, no NCCL call,
 * the call only enqueues 2*n_steps launches chained by programmatic dependent launch.  With state_dev != NULL the
 * cursor and iteration are read from (and advanced in) device memory, so the call can be captured in a CUDA graph and
 * replayed; with state_dev == NULL the call is host-driven: step s works on batch batch_order_host[s] (model.fit's
 * shuffled epoch order, n_steps entries) or, with batch_order_host == NULL, on batch (cursor0 + s) % n_batches, with the
 * schedule of iteration iteration0 + s (from the second step of a call on, the forward/backward kernel loads its first
 * rows before the previous step has finished).  world == 1 is allowed (no peer, same two launches).  Every rank owns a receive buffer of dlas_mlp_dp_recv_bytes(...) bytes,
 * ZEROED once by the caller BEFORE any rank's first step (synchronise the ranks after clearing), in memory its peers can
 * store to (CUDA IPC / VMM / symmetric memory);
 * recv_bases_dev[r] = the address under which THIS process reaches rank r's buffer.  A rank whose peers' words do not
 * arrive within a bounded spin writes NaN into its parameters and loss ring instead of hanging. */
size_t dlas_mlp_dp_recv_bytes(const int* dims, int n_layers, int mode, int world);   /* 0: not supported */
int dlas_mlp_dp_step_at(const int* dims, int n_layers, int loss, int mode, const dlas_opt_cfg* cfg, float* params,
                        float* m, float* v, const float* X_base, const float* y_base, size_t batch_rows,
                        dlas_step_state* state_dev, int64_t cursor0, int64_t iteration0,
                        const int64_t* batch_order_host, const uint64_t* recv_bases_dev, int rank, int world,
                        float grad_scale, int64_t n_batches, float* loss_ring, int64_t ring_len, int64_t n_steps,
                        void* workspace, size_t workspace_bytes, void* stream);

/* n_steps consecutive single-GPU mini-batch steps (the `model.fit` inner loop, cnn.py:127-128):
 * step s uses rows [b*batch, (b+1)*batch) with b = batch_order_host ? batch_order_host[s]
 * : first_batch + s.  step_losses (device, n_steps floats, may be NULL) receives the batch-mean
 * loss of each step.  iteration0 = batches applied before this call. */
int dlas_mlp_fit_steps(const int* dims, int n_layers, int loss, int mode, const dlas_opt_cfg* cfg,
                       float* params, float* m, float* v, const float* X, const float* y,
                       size_t n_rows, int batch, const int64_t* batch_order_host,
                       int64_t first_batch, int n_steps, uint64_t iteration0, float* step_losses,
                       float* grad_sum, void* workspace, size_t workspace_bytes, void* stream);

/* Row-strided variants (dynamic ANN, Readme.md:19; no code in the snapshot): row i of the input is the dims[0] contiguous
 * floats starting at X + i * x_row_stride.  x_row_stride == dims[0] is the dense matrix of the plain entry points (which
 * call these); x_row_stride = F < dims[0] lays n_lags = dims[0] / F overlapping LAG WINDOWS over a raw series of
 * F-feature rows -- window i = rows i .. i + n_lags - 1, lag-major, exactly what dlas_lag_windows_f32 would materialise
 * (bit-identical results), without the n_lags-fold copy: a 24-lag window over 8 features costs 36 bytes of HBM traffic
 * per sample instead of 772.  The series must hold n_rows + n_lags - 1 rows; y[i] belongs to window i; batch b of the
 * step loop starts b * batch rows further.  DLAS_MODE_TC needs x_row_stride % 4 == 0. */
int dlas_mlp_grad_strided(const int* dims, int n_layers, int loss, int mode, const float* params,
                          const float* X, size_t x_row_stride, const float* y, size_t n_rows, float* grad_sum,
                          void* workspace, size_t workspace_bytes, void* stream);
int dlas_mlp_fit_steps_strided(const int* dims, int n_layers, int loss, int mode, const dlas_opt_cfg* cfg,
                               float* params, float* m, float* v, const float* X, size_t x_row_stride,
                               const float* y, size_t n_rows, int batch, const int64_t* batch_order_host,
                               int64_t first_batch, int n_steps, uint64_t iteration0, float* step_losses,
                               float* grad_sum, void* workspace, size_t workspace_bytes, void* stream);
int dlas_mlp_predict_strided(const int* dims, int n_layers, int mode, const float* params, const float* X,
                             size_t x_row_stride, size_t n_rows, float* yhat, void* stream);
int dlas_mlp_eval_loss_strided(const int* dims, int n_layers, int loss, int mode, const float* params,
                               const float* X, size_t x_row_stride, const float* y, size_t n_rows,
                               float* loss_sum, void* workspace, size_t workspace_bytes, void* stream);

/* Inference (model.predict / test_on_batch, cnn.py:132): yhat (n_rows, dims[n_layers]). */
int dlas_mlp_predict(const int* dims, int n_layers, int mode, const float* params, const float* X,
                     size_t n_rows, float* yhat, void* stream);
/* The same with the workspace of dlas_mlp_workspace_bytes: DLAS_MODE_TC on d0-256-256-1 needs scratch for its forward
 * pass (layer 1 runs in its own kernel and hands 64 KB tile images to the pair kernel); without one, dlas_mlp_predict
 * serves that shape from the fp32 kernels.  Any number of rows (65,536 per pass).  Other shapes ignore the workspace. */
int dlas_mlp_predict_ws(const int* dims, int n_layers, int mode, const float* params, const float* X,
                        size_t n_rows, float* yhat, void* workspace, size_t workspace_bytes, void* stream);
int dlas_mlp_predict_ws_strided(const int* dims, int n_layers, int mode, const float* params, const float* X,
                                size_t x_row_stride, size_t n_rows, float* yhat, void* workspace,
                                size_t workspace_bytes, void* stream);
/* loss_sum[0] = SUM of per-row losses (validation / test loss numerator, cnn.py:121,:132) */
int dlas_mlp_eval_loss(const int* dims, int n_layers, int loss, int mode, const float* params,
                       const float* X, const float* y, size_t n_rows, float* loss_sum,
                       void* workspace, size_t workspace_bytes, void* stream);

/* dlas_mlp_grad without a caller-provided workspace.  This entry point has no counterpart in the
 * reference; it exists for hosts that may not call cudaMalloc at the point where they train (a
 * plug-in inside an executor that owns and has locked all device memory, a serving loop that bans
 * allocator calls because they stall the device).  Same arguments and results as dlas_mlp_grad in
 * DLAS_MODE_FP32; the per-CTA gradient partials live in a device array built into the library:
 * DLAS_NOALLOC_SCRATCH_BYTES per GPU, held until the process exits.  Layer widths that need more
 * scratch than that are refused with DLAS_ERR_WORKSPACE.  It calls no allocating routine.
 * The FIRST call on each device is a warm-up: it looks up the array's device address, and -- CUDA
 * loads modules lazily (the default since CUDA 12.2, and always with the static runtime this library
 * links) -- it is also the moment the driver places the library's kernels and this array on that GPU.
 * Make that first call (n_rows = 0 is enough) before the host's pool takes the rest of device memory
 * and before any stream capture starts, or run the process with CUDA_MODULE_LOADING=EAGER; from the
 * second call on a device onwards nothing is loaded and free device memory does not change.
 * All calls share the array: calls that overlap on the GPU (two streams, two threads) race -- keep
 * them on one stream. */
#define DLAS_NOALLOC_SCRATCH_BYTES (4u << 20)
int dlas_mlp_grad_noalloc(const int* dims, int n_layers, int loss, const float* params,
                          const float* X, const float* y, size_t n_rows, float* grad_sum,
                          void* stream);

/* ---- LSTM regressor (Readme.md:21; BASELINE.json config 5) ---------------------------------
 * n_layers stacked LSTM layers of width hidden over X (n_seq, T, d_in) batch-major, last hidden
 * state -> Dense(hidden -> 1).  Flat layout: per layer K[(d_l+hidden)][4*hidden] (input kernel
 * rows then recurrent kernel rows; gate column blocks i,f,c,o), b[4*hidden]; then head w[hidden],
 * b[1]. */
size_t dlas_lstm_n_params(int d_in, int hidden, int n_layers);
size_t dlas_lstm_workspace_bytes(int d_in, int hidden, int n_layers, int T, size_t n_seq);
/* forward + loss + full BPTT over n_seq sequences: grad_sum[0..n_params) = SUM of per-sequence
 * gradients, grad_sum[n_params] = SUM of per-sequence losses (the body of one fit batch,
 * cnn.py:127-128; the update is dlas_opt_step). */
int dlas_lstm_grad(int d_in, int hidden, int n_layers, int T, int loss, const float* params,
                   const float* X, const float* y, size_t n_seq, float* grad_sum, void* workspace,
                   size_t workspace_bytes, void* stream);
/* The same step with an arithmetic mode (dlas_lstm_grad is mode DLAS_MODE_FP32).  DLAS_MODE_TC (hidden = 128, d_in <= 16;
 * dlas_lstm_mode_supported) runs the stages that have a tcgen05 kernel on the tensor cores -- fp16 operands, fp32
 * accumulation in TMEM, fp32 cell state -- and the others in fp32; stash and parameter layout are the same, the
 * workspace size does not depend on the mode.  Stated tolerance of the mode: tests/test_lstm_gpu.py. */
int dlas_lstm_mode_supported(int d_in, int hidden, int n_layers, int mode);
int dlas_lstm_grad_mode(int mode, int d_in, int hidden, int n_layers, int T, int loss, const float* params,
                        const float* X, const float* y, size_t n_seq, float* grad_sum, void* workspace,
                        size_t workspace_bytes, void* stream);
int dlas_lstm_predict_mode(int mode, int d_in, int hidden, int n_layers, int T, const float* params,
                           const float* X, size_t n_seq, float* yhat, void* workspace,
                           size_t workspace_bytes, void* stream);
/* forward only: yhat (n_seq) -- model.predict / test_on_batch, cnn.py:132 */
int dlas_lstm_predict(int d_in, int hidden, int n_layers, int T, const float* params,
                      const float* X, size_t n_seq, float* yhat, void* workspace,
                      size_t workspace_bytes, void* stream);

/* ---- the script's own CNN (cnn.py:110-114):
 *   Convolution1D(1, F, K, activation='relu', border_mode='valid') -> Dropout(1 - keep) -> Flatten()
 *   -> Dense((L-K+1)*F, n_out)            [cnn.py: L=48, K=13, F=100, n_out=12 -> 44,612 parameters]
 * flat parameters: Wc[K][F], bc[F], Wd[(L-K+1)*F][n_out] (Flatten is time-major: t*F + f), bd[n_out].
 * X is (n_rows, L) fp32 (one input channel), y (n_rows, n_out).  n_out <= 16. */
size_t dlas_cnn_n_params(int L, int K, int F, int n_out);                 /* 0 on a bad shape */
size_t dlas_cnn_workspace_bytes(int L, int K, int F, int n_out, size_t n_rows);
/* replaces the forward+backward half of model.fit's train-on-batch (cnn.py:124-128):
 * grad_sum[0..n_params) = gradient SUMS over the rows, grad_sum[n_params] = sum of per-row losses.
 * Dropout is inverted (a * mask / keep); mask bit of (row r, unit j) is the counter hash of
 * (sample0 + r) * T*F + j on stream 300 of dropout_seed, so every rank/step draws its own bits when
 * the caller passes sample0 = iteration * global_batch + first row of this rank.  keep >= 1: no dropout.
 * Follow with dlas_opt_step (cnn.py:117 SGD). */
int dlas_cnn_grad(int L, int K, int F, int n_out, int loss, const float* params, const float* X,
                  const float* y, size_t n_rows, float keep, uint64_t dropout_seed, uint64_t sample0,
                  float* grad_sum, void* workspace, size_t workspace_bytes, void* stream);
/* n_steps consecutive single-GPU steps of the script's loop (model.fit, cnn.py:127-128, batch_size=20): step s trains on
 * rows [b*batch, (b+1)*batch) with b = batch_order_host ? batch_order_host[s] : first_batch + s; the reduction of the
 * gradient partials applies the optimiser update in the same kernel (six launches per step, no host work between
 * steps); the dropout mask of step s is drawn for samples [(iteration0 + s)*batch, ...), as dlas_cnn_grad would with
 * that sample0.  step_losses (device, n_steps floats, may be NULL) receives the batch-mean losses. */
int dlas_cnn_fit_steps(int L, int K, int F, int n_out, int loss, const dlas_opt_cfg* cfg, float* params, float* m,
                       float* v, const float* X, const float* y, size_t n_rows, int batch,
                       const int64_t* batch_order_host, int64_t first_batch, int n_steps, uint64_t iteration0,
                       float keep, uint64_t dropout_seed, float* step_losses, void* workspace,
                       size_t workspace_bytes, void* stream);
/* model.predict / test_on_batch (cnn.py:132): test-time forward, dropout off; yhat (n_rows, n_out) */
int dlas_cnn_predict(int L, int K, int F, int n_out, const float* params, const float* X,
                     size_t n_rows, float* yhat, void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DLAS_B200_H */
