"""Thin tensor-level wrappers over the C ABI.  PyTorch is used for device memory and streams only;
every computation below happens inside libdlas_b200.so."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import check, int_array, ptr, stream_ptr

BASE_SEED = 20261106


def _dev_f32(t, name):
    if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
        raise TypeError("%s must be a contiguous float32 CUDA tensor" % name)
    return t


# ---- Gilbert --------------------------------------------------------------------------------
def gilbert(P, S, R, out=None, variant=_lib.GILBERT_MUFU):
    """q = P * S^1.89 / (435 * R^0.546), element-wise over fp32 CUDA vectors (Readme.md:7-8)."""
    _dev_f32(P, "P"), _dev_f32(S, "S"), _dev_f32(R, "R")
    if not (P.numel() == S.numel() == R.numel()):
        raise ValueError("P, S, R must have the same number of elements")
    if out is None:
        out = torch.empty_like(P)
    _dev_f32(out, "out")
    check(_lib.load().dlas_gilbert_f32(ptr(P), ptr(S), ptr(R), ptr(out), P.numel(), int(variant), stream_ptr()))
    return out


# ---- synthetic inputs -------------------------------------------------------------------------
def synth_gilbert(n, seed=BASE_SEED, start=0, device="cuda"):
    P, S, R = (torch.empty(n, dtype=torch.float32, device=device) for _ in range(3))
    check(_lib.load().dlas_synth_gilbert_f32(ptr(P), ptr(S), ptr(R), seed, start, n, stream_ptr()))
    return P, S, R


def synth_rows(n_rows, n_features, seed=BASE_SEED, start=0, device="cuda", out=None):
    if out is None:
        X = torch.empty((n_rows, n_features), dtype=torch.float32, device=device)
        y = torch.empty((n_rows, 1), dtype=torch.float32, device=device)
    else:
        X, y = out
    check(_lib.load().dlas_synth_rows_f32(ptr(X), ptr(y), seed, start, n_rows, n_features, stream_ptr()))
    return X, y


def synth_series(n_seq, T, d_in, seed=BASE_SEED, start=0, device="cuda"):
    X = torch.empty((n_seq, T, d_in), dtype=torch.float32, device=device)
    y = torch.empty((n_seq, 1), dtype=torch.float32, device=device)
    check(_lib.load().dlas_synth_series_f32(ptr(X), ptr(y), seed, start, n_seq, T, d_in, stream_ptr()))
    return X, y


def split_codes(n, seed=BASE_SEED, start=0, device="cuda"):
    """cnn.py:68 randomSplit([0.64, 0.16, 0.2]) -> int8 codes 0/1/2 per row."""
    code = torch.empty(n, dtype=torch.int8, device=device)
    check(_lib.load().dlas_split_codes(ptr(code), seed, start, n, stream_ptr()))
    return code


def lag_windows(series, n_lags):
    """(n, F) raw series -> (n - n_lags + 1, n_lags * F) lag-major windows (dynamic ANN input)."""
    _dev_f32(series, "series")
    n, f = series.shape
    if n < n_lags:
        return torch.empty((0, n_lags * f), dtype=torch.float32, device=series.device)
    out = torch.empty((n - n_lags + 1, n_lags * f), dtype=torch.float32, device=series.device)
    check(_lib.load().dlas_lag_windows_f32(ptr(series), n, f, n_lags, ptr(out), stream_ptr()))
    return out


# ---- MLP ------------------------------------------------------------------------------------
def mlp_n_params(dims):
    return int(_lib.load().dlas_mlp_n_params(int_array(dims), len(dims) - 1))


def mlp_mode_supported(dims, mode):
    return bool(_lib.load().dlas_mlp_mode_supported(int_array(dims), len(dims) - 1, int(mode)))


def mlp_workspace(dims, mode=_lib.MODE_FP32, device="cuda"):
    nbytes = int(_lib.load().dlas_mlp_workspace_bytes(int_array(dims), len(dims) - 1, int(mode)))
    if nbytes == 0:
        raise _lib.DlasError(-5, "layer widths %s are not supported in mode %d" % (list(dims), mode))
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def make_opt_cfg(kind, lr, momentum=0.0, decay=0.0, nesterov=False, beta1=0.9, beta2=0.999, eps=1e-7):
    return _lib.OptCfg(int(kind), float(lr), float(momentum), float(decay), int(bool(nesterov)),
                       float(beta1), float(beta2), float(eps))


class LagWindows:
    """Lag windows laid over a raw series WITHOUT materialising them (dynamic ANN, Readme.md:19): window i is rows
    i .. i + n_lags - 1 of `series` (n, F) flattened lag-major = the n_lags * F contiguous floats starting at row i, so
    the window matrix is just the series read with a row stride of F floats (dlas_mlp_*_strided).  shape mirrors the
    materialised (n - n_lags + 1, n_lags * F) matrix ops.lag_windows would build; rows [a, b) by slicing."""

    def __init__(self, series, n_lags, start=0, count=None):
        _dev_f32(series, "series")
        self.series, self.n_lags, self.start = series, int(n_lags), int(start)
        total = series.shape[0] - self.n_lags + 1
        self.count = total - self.start if count is None else int(count)
        if self.count < 0 or self.start + self.count > max(total, 0):
            raise ValueError("window range [%d, %d) outside the %d windows of this series" % (self.start, self.start + self.count, total))
        self.row_stride = int(series.shape[1])
        self.shape = (self.count, self.n_lags * self.row_stride)
        self.device = series.device

    def __getitem__(self, sl):
        if not isinstance(sl, slice) or sl.step not in (None, 1):
            raise TypeError("LagWindows supports contiguous row slices only")
        a, b, _ = sl.indices(self.count)
        return LagWindows(self.series, self.n_lags, self.start + a, max(0, b - a))

    def data_ptr(self):
        return self.series.data_ptr() + self.start * self.row_stride * 4

    def materialize(self):
        return lag_windows(self.series, self.n_lags)[self.start:self.start + self.count]


def _x_stride(X, dims):
    return X.row_stride if isinstance(X, LagWindows) else int(dims[0])


def mlp_grad(dims, loss, mode, params, X, y, grad_sum, workspace, n_rows=None):
    n_rows = X.shape[0] if n_rows is None else n_rows
    check(_lib.load().dlas_mlp_grad_strided(int_array(dims), len(dims) - 1, int(loss), int(mode), ptr(params), ptr(X),
                                            _x_stride(X, dims), ptr(y), n_rows, ptr(grad_sum), ptr(workspace), workspace.numel(),
                                            stream_ptr()))
    return grad_sum


def mlp_status(dims, mode, workspace):
    """Synchronise the current stream and raise DlasError if a step enqueued with this workspace failed on the device
    (DLAS_ERR_TIMEOUT = -6, DLAS_ERR_RANGE = -7: the fp16 operand range of the tensor-core mode was exceeded)."""
    check(_lib.load().dlas_mlp_status(int_array(dims), len(dims) - 1, int(mode), ptr(workspace), workspace.numel(), stream_ptr()))


def opt_step(cfg, params, m, v, grad_sum, grad_scale, iteration, loss_out=None):
    check(_lib.load().dlas_opt_step(C.byref(cfg), ptr(params), ptr(m), ptr(v), ptr(grad_sum), params.numel(),
                                    float(grad_scale), int(iteration), ptr(loss_out), stream_ptr()))


def mlp_fit_steps(dims, loss, mode, cfg, params, m, v, X, y, batch, n_steps, iteration0, step_losses, grad_sum,
                  workspace, batch_order=None, first_batch=0):
    order = None
    if batch_order is not None:
        order = (C.c_int64 * len(batch_order))(*[int(b) for b in batch_order])
    check(_lib.load().dlas_mlp_fit_steps_strided(int_array(dims), len(dims) - 1, int(loss), int(mode), C.byref(cfg),
                                         ptr(params), ptr(m), ptr(v), ptr(X), _x_stride(X, dims), ptr(y), X.shape[0], int(batch),
                                         order, int(first_batch), int(n_steps), int(iteration0),
                                         ptr(step_losses), ptr(grad_sum), ptr(workspace), workspace.numel(),
                                         stream_ptr()))


def mlp_predict(dims, mode, params, X, out=None, workspace=None):
    """yhat of every row.  With the model's workspace the call is dlas_mlp_predict_ws: the shapes whose tensor-core forward pass
    needs scratch (d0-256-256-1) then run on the tensor cores too."""
    if not isinstance(X, LagWindows):
        _dev_f32(X, "X")
    if out is None:
        out = torch.empty((X.shape[0], dims[-1]), dtype=torch.float32, device=X.device)
    if workspace is not None:
        check(_lib.load().dlas_mlp_predict_ws_strided(int_array(dims), len(dims) - 1, int(mode), ptr(params), ptr(X), _x_stride(X, dims),
                                                      X.shape[0], ptr(out), ptr(workspace), workspace.numel(), stream_ptr()))
        return out
    check(_lib.load().dlas_mlp_predict_strided(int_array(dims), len(dims) - 1, int(mode), ptr(params), ptr(X), _x_stride(X, dims),
                                               X.shape[0], ptr(out), stream_ptr()))
    return out


def mlp_eval_loss(dims, loss, mode, params, X, y, workspace, loss_sum=None):
    if loss_sum is None:
        loss_sum = torch.empty(1, dtype=torch.float32, device=X.device)
    check(_lib.load().dlas_mlp_eval_loss_strided(int_array(dims), len(dims) - 1, int(loss), int(mode), ptr(params), ptr(X),
                                                 _x_stride(X, dims), ptr(y), X.shape[0], ptr(loss_sum), ptr(workspace),
                                                 workspace.numel(), stream_ptr()))
    return loss_sum


# ---- kernel timing (bench.py roofline) --------------------------------------------------------
PROF_GILBERT, PROF_MLP_GRAD = 0, 1


def prof_enable(on=True):
    check(_lib.load().dlas_prof_enable(int(bool(on))))


def prof_read(slot):
    """(summed kernel milliseconds, launches) recorded for `slot` since the last read."""
    ms, cnt = C.c_float(0.0), C.c_int(0)
    check(_lib.load().dlas_prof_read(int(slot), C.byref(ms), C.byref(cnt)))
    return float(ms.value), int(cnt.value)


# ---- LSTM -------------------------------------------------------------------------------------
def lstm_n_params(d_in, hidden, n_layers):
    return int(_lib.load().dlas_lstm_n_params(d_in, hidden, n_layers))


def lstm_workspace(d_in, hidden, n_layers, T, n_seq, device="cuda"):
    nbytes = int(_lib.load().dlas_lstm_workspace_bytes(d_in, hidden, n_layers, T, n_seq))
    if nbytes == 0:
        raise _lib.DlasError(-5, "LSTM shape d_in=%d hidden=%d layers=%d is not supported" % (d_in, hidden, n_layers))
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def lstm_mode_supported(d_in, hidden, n_layers, mode):
    return bool(_lib.load().dlas_lstm_mode_supported(int(d_in), int(hidden), int(n_layers), int(mode)))


def lstm_grad(d_in, hidden, n_layers, T, loss, params, X, y, grad_sum, workspace, mode=_lib.MODE_FP32):
    check(_lib.load().dlas_lstm_grad_mode(int(mode), d_in, hidden, n_layers, T, int(loss), ptr(params), ptr(X), ptr(y), X.shape[0],
                                          ptr(grad_sum), ptr(workspace), workspace.numel(), stream_ptr()))
    return grad_sum


def lstm_predict(d_in, hidden, n_layers, T, params, X, workspace, out=None, mode=_lib.MODE_FP32):
    if out is None:
        out = torch.empty((X.shape[0], 1), dtype=torch.float32, device=X.device)
    check(_lib.load().dlas_lstm_predict_mode(int(mode), d_in, hidden, n_layers, T, ptr(params), ptr(X), X.shape[0], ptr(out),
                                        ptr(workspace), workspace.numel(), stream_ptr()))
    return out


# ---- the CNN of cnn.py:110-114 -----------------------------------------------------------------------
def cnn_n_params(L, K, F, n_out):
    n = int(_lib.load().dlas_cnn_n_params(L, K, F, n_out))
    if n == 0:
        raise _lib.DlasError(-1, "CNN shape L=%d K=%d F=%d n_out=%d is not supported" % (L, K, F, n_out))
    return n


def cnn_workspace(L, K, F, n_out, n_rows, device="cuda"):
    nbytes = int(_lib.load().dlas_cnn_workspace_bytes(L, K, F, n_out, int(n_rows)))
    return torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)


def cnn_grad(L, K, F, n_out, loss, params, X, y, grad_sum, workspace, keep=1.0, dropout_seed=0, sample0=0):
    check(_lib.load().dlas_cnn_grad(L, K, F, n_out, int(loss), ptr(params), ptr(X), ptr(y), X.shape[0], float(keep),
                                    int(dropout_seed), int(sample0), ptr(grad_sum), ptr(workspace), workspace.numel(),
                                    stream_ptr()))
    return grad_sum


def cnn_fit_steps(L, K, F, n_out, loss, cfg, params, m, v, X, y, batch, batch_order, iteration0, keep, dropout_seed, step_losses,
                  workspace, first_batch=0, n_steps=None):
    """n_steps single-GPU steps (gradient kernels + reduction fused with the update), batch order from the host."""
    order = None
    if batch_order is not None:
        order = (C.c_int64 * len(batch_order))(*[int(b) for b in batch_order])
        n_steps = len(batch_order)
    check(_lib.load().dlas_cnn_fit_steps(L, K, F, n_out, int(loss), C.byref(cfg), ptr(params), ptr(m), ptr(v), ptr(X), ptr(y),
                                         X.shape[0], int(batch), order, int(first_batch), int(n_steps), int(iteration0), float(keep),
                                         int(dropout_seed), ptr(step_losses), ptr(workspace), workspace.numel(), stream_ptr()))


def cnn_predict(L, K, F, n_out, params, X, workspace, out=None):
    if out is None:
        out = torch.empty((X.shape[0], n_out), dtype=torch.float32, device=X.device)
    check(_lib.load().dlas_cnn_predict(L, K, F, n_out, ptr(params), ptr(X), X.shape[0], ptr(out), ptr(workspace),
                                       workspace.numel(), stream_ptr()))
    return out


# ---- replayable steps: cursor + iteration counter in device memory (CUDA-graph friendly) --------------
def mlp_grad_at(dims, loss, mode, params, X_base, y_base, batch_rows, state, grad_sum, workspace, parity_stride=0):
    check(_lib.load().dlas_mlp_grad_at(int_array(dims), len(dims) - 1, int(loss), int(mode), ptr(params), ptr(X_base),
                                       ptr(y_base), int(batch_rows), ptr(state), ptr(grad_sum), int(parity_stride),
                                       ptr(workspace), workspace.numel(), stream_ptr()))


def dp_allreduce_update(cfg, params, m, v, peer_bases, rank, world, parity_stride, flags_offset, grad_scale, state,
                        loss_ring=None):
    check(_lib.load().dlas_dp_allreduce_update(C.byref(cfg), ptr(params), ptr(m), ptr(v), ptr(peer_bases), int(rank), int(world),
                                               params.numel(), int(parity_stride), int(flags_offset), float(grad_scale),
                                               ptr(state), ptr(loss_ring), 0 if loss_ring is None else loss_ring.numel(),
                                               stream_ptr()))


def prof_sm_clock(reset=True):
    """Average SM MHz seen by the tensor-core MLP kernel since the last reset (0.0 if it has not run)."""
    mhz = C.c_double(0.0)
    check(_lib.load().dlas_prof_sm_clock(C.byref(mhz), 1 if reset else 0))
    return float(mhz.value)


def mlp_dp_recv_bytes(dims, mode, world):
    return int(_lib.load().dlas_mlp_dp_recv_bytes(int_array(dims), len(dims) - 1, int(mode), int(world)))


def mlp_dp_step_at(dims, loss, mode, cfg, params, m, v, X_base, y_base, batch_rows, state, recv_bases, rank, world, grad_scale,
                   n_batches, loss_ring, workspace, n_steps=1, cursor0=0, iteration0=0, batch_order=None):
    """n_steps whole data-parallel steps (DLAS_MODE_TC): forward/backward kernel + the reduce/push/poll/update/advance kernel.
    batch_order (host-driven flavour only): the batch index of every step, e.g. model.fit's shuffled epoch order."""
    order = None
    if batch_order is not None:
        order = (C.c_int64 * len(batch_order))(*[int(b) for b in batch_order])
        n_steps = len(batch_order)
    check(_lib.load().dlas_mlp_dp_step_at(int_array(dims), len(dims) - 1, int(loss), int(mode), C.byref(cfg), ptr(params), ptr(m),
                                          ptr(v), ptr(X_base), ptr(y_base), int(batch_rows), ptr(state), int(cursor0),
                                          int(iteration0), order, ptr(recv_bases), int(rank), int(world), float(grad_scale), int(n_batches), ptr(loss_ring),
                                          0 if loss_ring is None else loss_ring.numel(), int(n_steps), ptr(workspace),
                                          workspace.numel(), stream_ptr()))


def opt_step_at(cfg, params, m, v, grad_sum, grad_scale, state, loss_ring=None):
    check(_lib.load().dlas_opt_step_at(C.byref(cfg), ptr(params), ptr(m), ptr(v), ptr(grad_sum), params.numel(),
                                       float(grad_scale), ptr(state), ptr(loss_ring),
                                       0 if loss_ring is None else loss_ring.numel(), stream_ptr()))


def step_advance(state, n_batches):
    check(_lib.load().dlas_step_advance(ptr(state), int(n_batches), stream_ptr()))
