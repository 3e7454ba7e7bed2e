// C-ABI entry points of the MLP path: mode dispatch (fp32 CUDA-core kernels / tcgen05 kernels)
// and the mini-batch step loop that stands in for Keras `model.fit` (cnn.py:127-128).
#include "common.cuh"

namespace dlas {
// mlp_simt.cu
size_t simt_workspace_bytes(const int* dims, int n_layers);
int simt_grad(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y,
              size_t n_rows, float* grad_sum, void* workspace, size_t workspace_bytes, cudaStream_t st,
              const long long* cursor = nullptr, size_t parity_stride = 0, size_t ldx = 0);
int simt_eval_loss(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y,
                   size_t n_rows, float* loss_sum, void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx = 0);
int simt_predict(const int* dims, int n_layers, const float* params, const float* X, size_t n_rows, float* yhat,
                 cudaStream_t st, size_t ldx = 0);
// optim.cu
int opt_step_impl(const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* grad_sum, size_t n,
                  float grad_scale, uint64_t iteration, float* loss_out, cudaStream_t st);
// mlp_tc.cu
bool tc_supported(const int* dims, int n_layers);
size_t tc_workspace_bytes(const int* dims, int n_layers);
int tc_grad(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y,
            size_t n_rows, float* grad_sum, void* workspace, size_t workspace_bytes, cudaStream_t st,
            const long long* cursor = nullptr, size_t parity_stride = 0, size_t ldx = 0);
int tc_fit_step(const int* dims, int n_layers, int loss, const dlas_opt_cfg* cfg, float* params, float* m, float* v,
                const float* X, const float* y, size_t n_rows, uint64_t iteration, float* loss_out, void* workspace,
                size_t workspace_bytes, cudaStream_t st, bool chained, size_t ldx = 0, bool more_follow = false);
int tc_predict(const int* dims, int n_layers, const float* params, const float* X, size_t n_rows, float* yhat, cudaStream_t st,
               size_t ldx);
int tc_eval_loss(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y, size_t n_rows,
                 float* loss_sum, void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx);
int tc_status(const void* workspace, size_t workspace_bytes, cudaStream_t st);
// mlp_tc64.cu: the d0-64-64-1 network (d0 <= 15) on tcgen05, one CTA per tile
bool tc64_supported(const int* dims, int n_layers);
size_t tc64_workspace_bytes(const int* dims);
int tc64_status(const int* dims, const void* workspace, size_t workspace_bytes, cudaStream_t st);
int tc64_grad(const int* dims, int loss, const float* params, const float* X, const float* y, size_t n_rows, float* grad_sum,
              void* workspace, size_t workspace_bytes, cudaStream_t st, const long long* cursor, size_t parity_stride, size_t ldx);
int tc64_fit_step(const int* dims, int loss, const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* X,
                  const float* y, size_t n_rows, uint64_t iteration, float* loss_out, void* workspace, size_t workspace_bytes,
                  cudaStream_t st, size_t ldx);
int tc64_predict(const int* dims, const float* params, const float* X, size_t n_rows, float* yhat, cudaStream_t st, size_t ldx);
int tc64_eval_loss(const int* dims, int loss, const float* params, const float* X, const float* y, size_t n_rows, float* loss_sum,
                   void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx);
// mlp_tc.cu + mlp_tc_l1.cu: d0-256-256-1 with 64 <= d0 <= 240 (the forward-only kinds need a workspace: dlas_mlp_predict_ws)
bool tcx_supported(const int* dims, int n_layers);
size_t tcx_workspace_bytes(const int* dims, int n_layers);
int tcx_grad(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y, size_t n_rows,
             float* grad_sum, void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx);
int tcx_fit_step(const int* dims, int n_layers, int loss, const dlas_opt_cfg* cfg, float* params, float* m, float* v, const float* X,
                 const float* y, size_t n_rows, uint64_t iteration, float* loss_out, void* workspace, size_t workspace_bytes,
                 cudaStream_t st, bool chained, size_t ldx, bool more_follow);
int tcx_predict(const int* dims, int n_layers, const float* params, const float* X, size_t n_rows, float* yhat, void* workspace,
                size_t workspace_bytes, cudaStream_t st, size_t ldx);
int tcx_eval_loss(const int* dims, int n_layers, int loss, const float* params, const float* X, const float* y, size_t n_rows,
                  float* loss_sum, void* workspace, size_t workspace_bytes, cudaStream_t st, size_t ldx);
inline bool tc_any(const int* dims, int n_layers) {
  return tc_supported(dims, n_layers) || tc64_supported(dims, n_layers) || tcx_supported(dims, n_layers);
}
size_t tc_dp_recv_words(int world);
int tc_dp_step(const int* dims, int n_layers, int loss, const dlas_opt_cfg* cfg, float* params, float* m, float* v,
               const float* X_base, const float* y_base, size_t batch_rows, dlas_step_state* state,
               int64_t cursor0, int64_t iteration0, const int64_t* batch_order_host, const uint64_t* recv_bases_dev, int rank,
               int world, float grad_scale, int64_t n_batches, float* loss_ring, int64_t ring_len, int64_t n_steps,
               void* workspace, size_t workspace_bytes, cudaStream_t st);

namespace {
int check_common(const char* fn, const int* dims, int n_layers, int loss, int mode) {
  DLAS_REQUIRE(dims != nullptr, "%s: dims is null", fn);
  DLAS_REQUIRE(n_layers >= 1 && n_layers <= 8, "%s: n_layers %d not in 1..8", fn, n_layers);
  DLAS_REQUIRE(loss == DLAS_LOSS_MSE || loss == DLAS_LOSS_MAE_CLIP, "%s: unknown loss %d", fn, loss);
  DLAS_REQUIRE(mode == DLAS_MODE_FP32 || mode == DLAS_MODE_TC, "%s: unknown mode %d", fn, mode);
  if (mode == DLAS_MODE_TC && !tc_any(dims, n_layers)) {
    set_error("%s: DLAS_MODE_TC does not support these layer widths (see dlas_mlp_mode_supported)", fn);
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  return DLAS_OK;
}
}  // namespace
}  // namespace dlas

using namespace dlas;

extern "C" size_t dlas_mlp_n_params(const int* dims, int n_layers) {
  if (!dims || n_layers < 1) return 0;
  size_t n = 0;
  for (int l = 0; l < n_layers; ++l) n += (size_t)dims[l] * dims[l + 1] + dims[l + 1];
  return n;
}

extern "C" int dlas_mlp_mode_supported(const int* dims, int n_layers, int mode) {
  if (!dims || n_layers < 1 || n_layers > 8) return 0;
  if (mode == DLAS_MODE_FP32) return simt_workspace_bytes(dims, n_layers) > 0 ? 1 : 0;
  if (mode == DLAS_MODE_TC) return tc_any(dims, n_layers) ? 1 : 0;
  return 0;
}

extern "C" size_t dlas_mlp_workspace_bytes(const int* dims, int n_layers, int mode) {
  if (!dims || n_layers < 1 || n_layers > 8) return 0;
  if (mode == DLAS_MODE_TC && tc64_supported(dims, n_layers)) {   // eval / predict in fp32 mode may share the workspace
    size_t a = tc64_workspace_bytes(dims), b = simt_workspace_bytes(dims, n_layers);
    return a > b ? a : b;
  }
  if (mode == DLAS_MODE_TC && tcx_supported(dims, n_layers)) return tcx_workspace_bytes(dims, n_layers);
  if (mode == DLAS_MODE_TC) return tc_supported(dims, n_layers) ? tc_workspace_bytes(dims, n_layers) : 0;
  return simt_workspace_bytes(dims, n_layers);
}

namespace {
int check_stride(const char* fn, const int* dims, size_t x_row_stride) {
  DLAS_REQUIRE(x_row_stride >= 1 && x_row_stride <= (size_t)dims[0],
               "%s: x_row_stride %zu not in 1..dims[0] = %d (dims[0] = dense rows; less = overlapping lag windows)", fn,
               x_row_stride, dims[0]);
  return DLAS_OK;
}
}  // namespace

extern "C" int dlas_mlp_grad_strided(const int* dims, int n_layers, int loss, int mode, const float* params,
                                     const float* X, size_t x_row_stride, const float* y, size_t n_rows, float* grad_sum,
                                     void* workspace, size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  int rc = check_common("dlas_mlp_grad", dims, n_layers, loss, mode);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(params && grad_sum && (n_rows == 0 || (X && y)), "dlas_mlp_grad: null pointer");
  rc = check_stride("dlas_mlp_grad_strided", dims, x_row_stride);
  if (rc != DLAS_OK) return rc;
  if (mode == DLAS_MODE_TC && tc64_supported(dims, n_layers))
    return tc64_grad(dims, loss, params, X, y, n_rows, grad_sum, workspace, workspace_bytes, (cudaStream_t)stream, nullptr, 0, x_row_stride);
  if (mode == DLAS_MODE_TC && tcx_supported(dims, n_layers))
    return tcx_grad(dims, n_layers, loss, params, X, y, n_rows, grad_sum, workspace, workspace_bytes, (cudaStream_t)stream, x_row_stride);
  if (mode == DLAS_MODE_TC)
    return tc_grad(dims, n_layers, loss, params, X, y, n_rows, grad_sum, workspace, workspace_bytes,
                   (cudaStream_t)stream, nullptr, 0, x_row_stride);
  return simt_grad(dims, n_layers, loss, params, X, y, n_rows, grad_sum, workspace, workspace_bytes,
                   (cudaStream_t)stream, nullptr, 0, x_row_stride);
}

extern "C" int dlas_mlp_grad(const int* dims, int n_layers, int loss, int mode, const float* params,
                             const float* X, const float* y, size_t n_rows, float* grad_sum,
                             void* workspace, size_t workspace_bytes, void* stream) {
  if (!dims) { set_error("dlas_mlp_grad: dims is null"); return DLAS_ERR_INVALID; }
  return dlas_mlp_grad_strided(dims, n_layers, loss, mode, params, X, (size_t)dims[0], y, n_rows, grad_sum, workspace,
                               workspace_bytes, stream);
}

extern "C" int dlas_mlp_grad_at(const int* dims, int n_layers, int loss, int mode, const float* params,
                                const float* X_base, const float* y_base, size_t batch_rows,
                                const dlas_step_state* state_dev, float* grad_sum, size_t parity_stride,
                                void* workspace, size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  int rc = check_common("dlas_mlp_grad_at", dims, n_layers, loss, mode);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(params && grad_sum && X_base && y_base && state_dev, "dlas_mlp_grad_at: null pointer");
  DLAS_REQUIRE(batch_rows >= 1, "dlas_mlp_grad_at: batch_rows is 0");
  const long long* cursor = reinterpret_cast<const long long*>(&state_dev->cursor);
  if (mode == DLAS_MODE_TC && tc64_supported(dims, n_layers))
    return tc64_grad(dims, loss, params, X_base, y_base, batch_rows, grad_sum, workspace, workspace_bytes, (cudaStream_t)stream, cursor,
                     parity_stride, 0);
  if (mode == DLAS_MODE_TC && tcx_supported(dims, n_layers)) {
    set_error("dlas_mlp_grad_at: the tensor-core path of %d-256-256-1 has no device-cursor flavour (use DLAS_MODE_FP32)", dims[0]);
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  if (mode == DLAS_MODE_TC)
    return tc_grad(dims, n_layers, loss, params, X_base, y_base, batch_rows, grad_sum, workspace, workspace_bytes,
                   (cudaStream_t)stream, cursor, parity_stride);
  return simt_grad(dims, n_layers, loss, params, X_base, y_base, batch_rows, grad_sum, workspace, workspace_bytes,
                   (cudaStream_t)stream, cursor, parity_stride);
}

extern "C" int dlas_mlp_fit_steps(const int* dims, int n_layers, int loss, int mode,
                                  const dlas_opt_cfg* cfg, float* params, float* m, float* v,
                                  const float* X, const float* y, size_t n_rows, int batch,
                                  const int64_t* batch_order_host, int64_t first_batch, int n_steps,
                                  uint64_t iteration0, float* step_losses, float* grad_sum,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  if (!dims) { set_error("dlas_mlp_fit_steps: dims is null"); return DLAS_ERR_INVALID; }
  return dlas_mlp_fit_steps_strided(dims, n_layers, loss, mode, cfg, params, m, v, X, (size_t)dims[0], y, n_rows, batch,
                                    batch_order_host, first_batch, n_steps, iteration0, step_losses, grad_sum, workspace,
                                    workspace_bytes, stream);
}

extern "C" int dlas_mlp_fit_steps_strided(const int* dims, int n_layers, int loss, int mode,
                                          const dlas_opt_cfg* cfg, float* params, float* m, float* v,
                                          const float* X, size_t x_row_stride, const float* y, size_t n_rows, int batch,
                                          const int64_t* batch_order_host, int64_t first_batch, int n_steps,
                                          uint64_t iteration0, float* step_losses, float* grad_sum,
                                          void* workspace, size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  int rc = check_common("dlas_mlp_fit_steps", dims, n_layers, loss, mode);
  if (rc != DLAS_OK) return rc;
  rc = check_stride("dlas_mlp_fit_steps_strided", dims, x_row_stride);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(cfg && params && m && X && y && grad_sum, "dlas_mlp_fit_steps: null pointer");
  DLAS_REQUIRE(batch >= 1, "dlas_mlp_fit_steps: batch %d < 1", batch);
  DLAS_REQUIRE(n_steps >= 0, "dlas_mlp_fit_steps: n_steps %d < 0", n_steps);
  const size_t n_batches = n_rows / (size_t)batch;  // Keras would also run a ragged last batch; callers
  DLAS_REQUIRE(n_batches >= 1 || n_steps == 0,      // pass it as its own call with batch = remainder
               "dlas_mlp_fit_steps: n_rows %zu < batch %d", n_rows, batch);
  const size_t n_params = dlas_mlp_n_params(dims, n_layers);
  const size_t d0 = x_row_stride, dL = dims[n_layers];   // batch b starts b*batch rows further: b*batch*stride floats
  cudaStream_t st = (cudaStream_t)stream;
  for (int s = 0; s < n_steps; ++s) {
    int64_t b = batch_order_host ? batch_order_host[s] : first_batch + s;
    DLAS_REQUIRE(b >= 0 && (size_t)b < n_batches, "dlas_mlp_fit_steps: batch index %lld out of range (%zu batches)",
                 (long long)b, n_batches);
    const float* Xb = X + (size_t)b * batch * d0;
    const float* yb = y + (size_t)b * batch * dL;
    if (mode == DLAS_MODE_TC && tc64_supported(dims, n_layers)) {
      rc = tc64_fit_step(dims, loss, cfg, params, m, v, Xb, yb, batch, iteration0 + (uint64_t)s, step_losses ? step_losses + s : nullptr,
                         workspace, workspace_bytes, st, x_row_stride);
      if (rc != DLAS_OK) return rc;
      continue;
    }
    if (mode == DLAS_MODE_TC && tcx_supported(dims, n_layers)) {
      rc = tcx_fit_step(dims, n_layers, loss, cfg, params, m, v, Xb, yb, batch, iteration0 + (uint64_t)s,
                        step_losses ? step_losses + s : nullptr, workspace, workspace_bytes, st, /*chained=*/s > 0, x_row_stride,
                        /*more_follow=*/s + 1 < n_steps);
      if (rc != DLAS_OK) return rc;
      continue;
    }
    if (mode == DLAS_MODE_TC) {   // gradient kernel + (reduction fused with the update): two launches per step
      rc = tc_fit_step(dims, n_layers, loss, cfg, params, m, v, Xb, yb, batch, iteration0 + (uint64_t)s,
                       step_losses ? step_losses + s : nullptr, workspace, workspace_bytes, st, /*chained=*/s > 0, x_row_stride,
                       /*more_follow=*/s + 1 < n_steps);
      if (rc != DLAS_OK) return rc;
      continue;
    }
    rc = simt_grad(dims, n_layers, loss, params, Xb, yb, batch, grad_sum, workspace, workspace_bytes, st, nullptr, 0, x_row_stride);
    if (rc != DLAS_OK) return rc;
    rc = opt_step_impl(cfg, params, m, v, grad_sum, n_params, 1.0f / (float)batch, iteration0 + (uint64_t)s,
                       step_losses ? step_losses + s : nullptr, st);
    if (rc != DLAS_OK) return rc;
  }
  return DLAS_OK;
}

extern "C" int dlas_mlp_predict(const int* dims, int n_layers, int mode, const float* params,
                                const float* X, size_t n_rows, float* yhat, void* stream) {
  if (!dims) { set_error("dlas_mlp_predict: dims is null"); return DLAS_ERR_INVALID; }
  return dlas_mlp_predict_strided(dims, n_layers, mode, params, X, (size_t)dims[0], n_rows, yhat, stream);
}

extern "C" int dlas_mlp_predict_strided(const int* dims, int n_layers, int mode, const float* params,
                                        const float* X, size_t x_row_stride, size_t n_rows, float* yhat, void* stream) {
  DLAS_CHECK_DEVICE();
  int rc = check_common("dlas_mlp_predict", dims, n_layers, DLAS_LOSS_MSE, DLAS_MODE_FP32);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(mode == DLAS_MODE_FP32 || mode == DLAS_MODE_TC, "dlas_mlp_predict: unknown mode %d", mode);
  DLAS_REQUIRE(params && (n_rows == 0 || (X && yhat)), "dlas_mlp_predict: null pointer");
  rc = check_stride("dlas_mlp_predict_strided", dims, x_row_stride);
  if (rc != DLAS_OK) return rc;
  if (mode == DLAS_MODE_TC) {   // forward-only kind of the tcgen05 kernel
    if (tc64_supported(dims, n_layers)) return tc64_predict(dims, params, X, n_rows, yhat, (cudaStream_t)stream, x_row_stride);
    if (tcx_supported(dims, n_layers)) return simt_predict(dims, n_layers, params, X, n_rows, yhat, (cudaStream_t)stream, x_row_stride);
    if (!tc_supported(dims, n_layers)) { set_error("dlas_mlp_predict: DLAS_MODE_TC does not support these layer widths"); return DLAS_ERR_UNSUPPORTED_SHAPE; }
    return tc_predict(dims, n_layers, params, X, n_rows, yhat, (cudaStream_t)stream, x_row_stride);
  }
  return simt_predict(dims, n_layers, params, X, n_rows, yhat, (cudaStream_t)stream, x_row_stride);
}

extern "C" int dlas_mlp_predict_ws(const int* dims, int n_layers, int mode, const float* params, const float* X, size_t n_rows,
                                   float* yhat, void* workspace, size_t workspace_bytes, void* stream) {
  if (!dims) { set_error("dlas_mlp_predict_ws: dims is null"); return DLAS_ERR_INVALID; }
  return dlas_mlp_predict_ws_strided(dims, n_layers, mode, params, X, (size_t)dims[0], n_rows, yhat, workspace, workspace_bytes, stream);
}

extern "C" int dlas_mlp_predict_ws_strided(const int* dims, int n_layers, int mode, const float* params, const float* X,
                                           size_t x_row_stride, size_t n_rows, float* yhat, void* workspace, size_t workspace_bytes,
                                           void* stream) {
  DLAS_CHECK_DEVICE();
  if (dims && n_layers >= 1 && n_layers <= 8 && mode == DLAS_MODE_TC && tcx_supported(dims, n_layers)) {
    DLAS_REQUIRE(params && (n_rows == 0 || (X && yhat)), "dlas_mlp_predict_ws: null pointer");
    int rc = check_stride("dlas_mlp_predict_ws_strided", dims, x_row_stride);
    if (rc != DLAS_OK) return rc;
    if (n_rows == 0) return DLAS_OK;
    return tcx_predict(dims, n_layers, params, X, n_rows, yhat, workspace, workspace_bytes, (cudaStream_t)stream, x_row_stride);
  }
  return dlas_mlp_predict_strided(dims, n_layers, mode, params, X, x_row_stride, n_rows, yhat, stream);   // no scratch needed
}

extern "C" int dlas_mlp_eval_loss(const int* dims, int n_layers, int loss, int mode, const float* params,
                                  const float* X, const float* y, size_t n_rows, float* loss_sum,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  if (!dims) { set_error("dlas_mlp_eval_loss: dims is null"); return DLAS_ERR_INVALID; }
  return dlas_mlp_eval_loss_strided(dims, n_layers, loss, mode, params, X, (size_t)dims[0], y, n_rows, loss_sum, workspace,
                                    workspace_bytes, stream);
}

extern "C" int dlas_mlp_eval_loss_strided(const int* dims, int n_layers, int loss, int mode, const float* params,
                                          const float* X, size_t x_row_stride, const float* y, size_t n_rows, float* loss_sum,
                                          void* workspace, size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  int rc = check_common("dlas_mlp_eval_loss", dims, n_layers, loss, DLAS_MODE_FP32);
  if (rc != DLAS_OK) return rc;
  DLAS_REQUIRE(mode == DLAS_MODE_FP32 || mode == DLAS_MODE_TC, "dlas_mlp_eval_loss: unknown mode %d", mode);
  DLAS_REQUIRE(params && loss_sum && (n_rows == 0 || (X && y)), "dlas_mlp_eval_loss: null pointer");
  rc = check_stride("dlas_mlp_eval_loss_strided", dims, x_row_stride);
  if (rc != DLAS_OK) return rc;
  if (mode == DLAS_MODE_TC) {
    if (tc64_supported(dims, n_layers))
      return tc64_eval_loss(dims, loss, params, X, y, n_rows, loss_sum, workspace, workspace_bytes, (cudaStream_t)stream, x_row_stride);
    if (tcx_supported(dims, n_layers))
      return tcx_eval_loss(dims, n_layers, loss, params, X, y, n_rows, loss_sum, workspace, workspace_bytes, (cudaStream_t)stream, x_row_stride);
    if (!tc_supported(dims, n_layers)) { set_error("dlas_mlp_eval_loss: DLAS_MODE_TC does not support these layer widths"); return DLAS_ERR_UNSUPPORTED_SHAPE; }
    return tc_eval_loss(dims, n_layers, loss, params, X, y, n_rows, loss_sum, workspace, workspace_bytes, (cudaStream_t)stream, x_row_stride);
  }
  return simt_eval_loss(dims, n_layers, loss, params, X, y, n_rows, loss_sum, workspace, workspace_bytes,
                        (cudaStream_t)stream, x_row_stride);
}

extern "C" size_t dlas_mlp_dp_recv_bytes(const int* dims, int n_layers, int mode, int world) {
  if (!dims || mode != DLAS_MODE_TC || !tc_supported(dims, n_layers) || world < 1 || world > 8) return 0;
  return tc_dp_recv_words(world) * 4;
}

extern "C" int dlas_mlp_dp_step_at(const int* dims, int n_layers, int loss, int mode, const dlas_opt_cfg* cfg, float* params,
                                   float* m, float* v, const float* X_base, const float* y_base, size_t batch_rows,
                                   dlas_step_state* state_dev, int64_t cursor0, int64_t iteration0,
                                   const int64_t* batch_order_host, const uint64_t* recv_bases_dev, int rank, int world,
                                   float grad_scale, int64_t n_batches, float* loss_ring, int64_t ring_len, int64_t n_steps,
                                   void* workspace, size_t workspace_bytes, void* stream) {
  DLAS_CHECK_DEVICE();
  int rc = check_common("dlas_mlp_dp_step_at", dims, n_layers, loss, mode);
  if (rc != DLAS_OK) return rc;
  if (mode != DLAS_MODE_TC) {
    set_error("dlas_mlp_dp_step_at: DLAS_MODE_TC only (fp32 mode: dlas_mlp_grad_at -> dlas_dp_allreduce_update -> dlas_step_advance)");
    return DLAS_ERR_UNSUPPORTED_SHAPE;
  }
  return tc_dp_step(dims, n_layers, loss, cfg, params, m, v, X_base, y_base, batch_rows, state_dev, cursor0, iteration0, batch_order_host,
                    recv_bases_dev, rank, world,
                    grad_scale, n_batches, loss_ring, ring_len, n_steps, workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int dlas_mlp_status(const int* dims, int n_layers, int mode, const void* workspace, size_t workspace_bytes,
                               void* stream) {
  DLAS_CHECK_DEVICE();
  int rc = check_common("dlas_mlp_status", dims, n_layers, DLAS_LOSS_MSE, mode);
  if (rc != DLAS_OK) return rc;
  if (mode != DLAS_MODE_TC) {   // the fp32 kernels have no device-side failure mode to report; still a completion point
    DLAS_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return DLAS_OK;
  }
  if (tc64_supported(dims, n_layers)) return tc64_status(dims, workspace, workspace_bytes, (cudaStream_t)stream);
  return tc_status(workspace, workspace_bytes, (cudaStream_t)stream);
}
