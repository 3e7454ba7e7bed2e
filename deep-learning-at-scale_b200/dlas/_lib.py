"""ctypes binding of libdlas_b200.so (include/dlas_b200.h).

The library is the product: if it is missing or a call fails this module raises -- there is no
PyTorch / NumPy fallback path anywhere in the package."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# DLAS_B200_LIB: debugging only (tools/tc_trace.py points it at the library built with `build.py --trace`)
LIB_PATH = os.environ.get("DLAS_B200_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libdlas_b200.so")

LOSS_MSE, LOSS_MAE_CLIP = 0, 1
OPT_SGD, OPT_ADAM = 0, 1
MODE_FP32, MODE_TC = 0, 1
GILBERT_POLY, GILBERT_MUFU = 0, 7


class DlasError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libdlas_b200 error %d: %s" % (code, msg))
        self.code = code


class OptCfg(C.Structure):
    _fields_ = [("kind", C.c_int32), ("lr", C.c_float), ("momentum", C.c_float), ("decay", C.c_float),
                ("nesterov", C.c_int32), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float)]


_p, _sz, _u64, _i, _f, _i64 = C.c_void_p, C.c_size_t, C.c_uint64, C.c_int, C.c_float, C.c_int64
_ip = C.POINTER(C.c_int)

# name -> (restype, argtypes); every symbol include/dlas_b200.h declares
SIGNATURES = {
    "dlas_version": (_i, []),
    "dlas_last_error": (C.c_char_p, []),
    "dlas_device_check": (_i, []),
    "dlas_sm_count": (_i, []),
    "dlas_prof_enable": (_i, [_i]),
    "dlas_prof_read": (_i, [_i, C.POINTER(C.c_float), C.POINTER(C.c_int)]),
    "dlas_gilbert_f32": (_i, [_p, _p, _p, _p, _sz, _i, _p]),
    "dlas_synth_gilbert_f32": (_i, [_p, _p, _p, _u64, _u64, _sz, _p]),
    "dlas_synth_rows_f32": (_i, [_p, _p, _u64, _u64, _sz, _i, _p]),
    "dlas_synth_series_f32": (_i, [_p, _p, _u64, _u64, _sz, _i, _i, _p]),
    "dlas_split_codes": (_i, [_p, _u64, _u64, _sz, _p]),
    "dlas_lag_windows_f32": (_i, [_p, _sz, _i, _i, _p, _p]),
    "dlas_mlp_n_params": (_sz, [_ip, _i]),
    "dlas_mlp_mode_supported": (_i, [_ip, _i, _i]),
    "dlas_mlp_workspace_bytes": (_sz, [_ip, _i, _i]),
    "dlas_mlp_grad": (_i, [_ip, _i, _i, _i, _p, _p, _p, _sz, _p, _p, _sz, _p]),
    "dlas_mlp_status": (_i, [_ip, _i, _i, _p, _sz, _p]),
    "dlas_opt_step": (_i, [C.POINTER(OptCfg), _p, _p, _p, _p, _sz, _f, _u64, _p, _p]),
    "dlas_mlp_grad_at": (_i, [_ip, _i, _i, _i, _p, _p, _p, _sz, _p, _p, _sz, _p, _sz, _p]),
    "dlas_dp_allreduce_update": (_i, [C.POINTER(OptCfg), _p, _p, _p, _p, _i, _i, _sz, _sz, _sz, _f, _p, _p, C.c_int64, _p]),
    "dlas_opt_step_at": (_i, [C.POINTER(OptCfg), _p, _p, _p, _p, _sz, _f, _p, _p, C.c_int64, _p]),
    "dlas_step_advance": (_i, [_p, C.c_int64, _p]),
    "dlas_mlp_fit_steps": (_i, [_ip, _i, _i, _i, C.POINTER(OptCfg), _p, _p, _p, _p, _p, _sz, _i,
                                C.POINTER(C.c_int64), C.c_int64, _i, _u64, _p, _p, _p, _sz, _p]),
    "dlas_mlp_grad_strided": (_i, [_ip, _i, _i, _i, _p, _p, _sz, _p, _sz, _p, _p, _sz, _p]),
    "dlas_mlp_fit_steps_strided": (_i, [_ip, _i, _i, _i, C.POINTER(OptCfg), _p, _p, _p, _p, _sz, _p, _sz, _i,
                                        C.POINTER(C.c_int64), C.c_int64, _i, _u64, _p, _p, _p, _sz, _p]),
    "dlas_mlp_predict_strided": (_i, [_ip, _i, _i, _p, _p, _sz, _sz, _p, _p]),
    "dlas_mlp_predict_ws": (_i, [_ip, _i, _i, _p, _p, _sz, _p, _p, _sz, _p]),
    "dlas_mlp_predict_ws_strided": (_i, [_ip, _i, _i, _p, _p, _sz, _sz, _p, _p, _sz, _p]),
    "dlas_mlp_eval_loss_strided": (_i, [_ip, _i, _i, _i, _p, _p, _sz, _p, _sz, _p, _p, _sz, _p]),
    "dlas_mlp_predict": (_i, [_ip, _i, _i, _p, _p, _sz, _p, _p]),
    "dlas_mlp_eval_loss": (_i, [_ip, _i, _i, _i, _p, _p, _p, _sz, _p, _p, _sz, _p]),
    "dlas_mlp_grad_noalloc": (_i, [_ip, _i, _i, _p, _p, _p, _sz, _p, _p]),
    "dlas_prof_sm_clock": (_i, [C.POINTER(C.c_double), _i]),
    "dlas_mlp_dp_recv_bytes": (_sz, [_ip, _i, _i, _i]),
    "dlas_mlp_dp_step_at": (_i, [_ip, _i, _i, _i, _p, _p, _p, _p, _p, _p, _sz, _p, _i64, _i64, C.POINTER(C.c_int64), _p, _i, _i, _f, _i64, _p, _i64, _i64, _p, _sz, _p]),
    "dlas_lstm_n_params": (_sz, [_i, _i, _i]),
    "dlas_lstm_workspace_bytes": (_sz, [_i, _i, _i, _i, _sz]),
    "dlas_lstm_grad": (_i, [_i, _i, _i, _i, _i, _p, _p, _p, _sz, _p, _p, _sz, _p]),
    "dlas_lstm_mode_supported": (_i, [_i, _i, _i, _i]),
    "dlas_lstm_grad_mode": (_i, [_i, _i, _i, _i, _i, _i, _p, _p, _p, _sz, _p, _p, _sz, _p]),
    "dlas_lstm_predict_mode": (_i, [_i, _i, _i, _i, _i, _p, _p, _sz, _p, _p, _sz, _p]),
    "dlas_lstm_predict": (_i, [_i, _i, _i, _i, _p, _p, _sz, _p, _p, _sz, _p]),
    "dlas_cnn_n_params": (_sz, [_i, _i, _i, _i]),
    "dlas_cnn_workspace_bytes": (_sz, [_i, _i, _i, _i, _sz]),
    "dlas_cnn_grad": (_i, [_i, _i, _i, _i, _i, _p, _p, _p, _sz, _f, _u64, _u64, _p, _p, _sz, _p]),
    "dlas_cnn_fit_steps": (_i, [_i, _i, _i, _i, _i, _p, _p, _p, _p, _p, _p, _sz, _i, C.POINTER(C.c_int64), _i64, _i, _u64, _f, _u64,
                                _p, _p, _sz, _p]),
    "dlas_cnn_predict": (_i, [_i, _i, _i, _i, _p, _p, _sz, _p, _p, _sz, _p]),
}

_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built: build it with
    `python deep-learning-at-scale_b200/build.py` (or `__graft_entry__.build()`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libdlas_b200.so not found at %s -- the CUDA extension is the product; "
                          "build it with `python deep-learning-at-scale_b200/build.py`" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise DlasError(rc, load().dlas_last_error().decode("utf-8", "replace"))


def int_array(vals):
    return (C.c_int * len(vals))(*[int(v) for v in vals])


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
