"""CPU checks of the drop-in boundary: the C-ABI library builds for sm_100a, loads, and exports
exactly what include/dlas_b200.h declares (no compute calls without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dlas_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dlas_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as G
    G.build()
    from dlas import _lib
    return _lib.load()


def test_header_declares_the_path():
    syms = declared_symbols()
    for must in ("dlas_gilbert_f32", "dlas_mlp_grad", "dlas_opt_step", "dlas_mlp_fit_steps", "dlas_mlp_predict",
                 "dlas_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol(lib):
    for s in declared_symbols():
        assert hasattr(lib, s), "libdlas_b200.so does not export %s" % s


def test_binding_table_matches_header(lib):
    from dlas import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()


def test_version_and_error_string(lib):
    assert lib.dlas_version() >= 100
    assert isinstance(lib.dlas_last_error(), bytes)


def test_host_only_queries(lib):
    from dlas import _lib
    dims = _lib.int_array([32, 256, 256, 1])
    assert lib.dlas_mlp_n_params(dims, 3) == 32 * 256 + 256 + 256 * 256 + 256 + 256 + 1 == 74497
    dims = _lib.int_array([8, 64, 64, 1])
    assert lib.dlas_mlp_n_params(dims, 3) == 4801


def test_mode_table_is_a_host_only_query(lib):
    """Which layer widths the tensor-core mode takes (no device needed): 32-256-256-1, d0-64-64-1 with d0 <= 15, d0-256-256-1 with
    64 <= d0 <= 240 and d0 % 8 == 0 (the dynamic ANN's 24 x 8 = 192 inputs); LSTM: hidden 128, d_in <= 16."""
    from dlas import _lib
    tc = lambda d: lib.dlas_mlp_mode_supported(_lib.int_array(d), len(d) - 1, _lib.MODE_TC)
    for d in ([32, 256, 256, 1], [8, 64, 64, 1], [1, 64, 64, 1], [15, 64, 64, 1], [192, 256, 256, 1], [64, 256, 256, 1], [240, 256, 256, 1]):
        assert tc(d) == 1, d
    for d in ([16, 64, 64, 1], [32, 256, 256, 2], [32, 256, 1], [100, 256, 256, 1], [248, 256, 256, 1], [48, 256, 256, 1],
              [192, 256, 128, 1], [32, 128, 256, 1]):
        assert tc(d) == 0, d
    assert lib.dlas_lstm_mode_supported(8, 128, 2, _lib.MODE_TC) == 1 and lib.dlas_lstm_mode_supported(16, 128, 4, _lib.MODE_TC) == 1
    assert lib.dlas_lstm_mode_supported(8, 64, 2, _lib.MODE_TC) == 0 and lib.dlas_lstm_mode_supported(17, 128, 2, _lib.MODE_TC) == 0
    assert lib.dlas_lstm_mode_supported(8, 64, 2, _lib.MODE_FP32) == 1


def test_no_cpu_fallback_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    rc = lib.dlas_gilbert_f32(None, None, None, None, 16, 0, None)
    assert rc == -3  # DLAS_ERR_UNSUPPORTED_DEVICE: fails loudly, computes nothing
    assert b"CUDA" in lib.dlas_last_error() or b"device" in lib.dlas_last_error()
    import dlas
    m = dlas.Sequential([dlas.Dense(8, 64, activation="relu"), dlas.Dense(64, 1)])
    with pytest.raises(RuntimeError):
        m.compile(loss="mse", optimizer=dlas.SGD())


def test_product_never_imports_oracle():
    """The product path must not import, call or link anything under oracle/ (comments may cite it)."""
    pkg = os.path.join(ROOT, "deep-learning-at-scale_b200")
    pat = re.compile(r"^\s*(import|from)\s+[^#\n]*\b(dlas_oracle|oracle)\b", re.M)
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                assert not pat.search(open(os.path.join(d, f), errors="replace").read()), f
            if f.endswith((".cu", ".cuh")):
                assert "#include \"../../oracle" not in open(os.path.join(d, f), errors="replace").read(), f


def declared_arity():
    """name -> number of parameters of every prototype in the header (0 for `(void)`)."""
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for name, params in re.findall(r"\b(dlas_[a-z0-9_]+)\s*\(([^;{}]*?)\)\s*;", src, flags=re.S):
        p = params.strip()
        out[name] = 0 if p in ("", "void") else p.count(",") + 1
    return out


def test_binding_arity_matches_header():
    """A ctypes table that drifts from the header corrupts the call silently: same parameter count on both sides."""
    from dlas import _lib
    ar = declared_arity()
    assert sorted(ar) == declared_symbols()
    for name, (_res, args) in _lib.SIGNATURES.items():
        assert len(args) == ar[name], "%s: header declares %d parameters, the binding passes %d" % (name, ar[name], len(args))


def test_tensor_core_kernel_resource_budget(lib):
    """ptxas -v of the tcgen05 kernel (kept by build.py): it launches with 168 registers x 384 threads (the whole register
    file; setmaxnreg then moves registers between the warpgroups) and must not spill more than a few words -- a 300-byte
    spill in the issuer once cost 40 % of the throughput."""
    log = os.path.join(ROOT, "deep-learning-at-scale_b200", "build", "ptxas_mlp_tc.log")
    if not os.path.exists(log):
        pytest.skip("no ptxas log (library was not built by build.py in this tree)")
    txt = open(log).read()
    m = re.search(r"Function properties for \S*mlp_tc_kernel\S*\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\s+"
                  r"ptxas info\s+: Used (\d+) registers", txt)
    assert m, "mlp_tc_kernel not found in the ptxas log"
    stack, st, ld, regs = map(int, m.groups())
    assert regs == 168 and stack <= 64 and st <= 64 and ld <= 64, (stack, st, ld, regs)
