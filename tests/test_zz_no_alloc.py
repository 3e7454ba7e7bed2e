"""dlas_mlp_grad_noalloc -- the entry point for hosts that are not allowed to allocate device
memory where they train (an executor plug-in whose pool owns and has locked the GPU's memory, a
serving loop that bans cudaMalloc because it stalls the device).  It keeps its scratch in an array
built into the library, so this test calls it through the C ABI with inputs and outputs only and
checks it against dlas_mlp_grad (which takes a caller-allocated workspace), twice in a row: the
built-in array keeps the first call's contents, and the second call must not see them.

Collected last (file name), so a run that stops at the first failure has already run every other
test; it passes on its own in a fresh process as well."""
import ctypes as C

import numpy as np
import pytest
import torch

import dlas_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-4  # the tolerance the MLP parity tests use (BASELINE.json, fp32 path)


@pytest.fixture(scope="module")
def env():
    import __graft_entry__ as G
    G.build()
    from dlas import _lib, ops
    return _lib, ops


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()


def _relerr(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


def _both(_lib, ops, dims, loss, p, X, y):
    lib = _lib.load()
    n = ops.mlp_n_params(dims)
    ref = torch.zeros(n + 1, device="cuda")
    ops.mlp_grad(dims, loss, _lib.MODE_FP32, p, X, y, ref, ops.mlp_workspace(dims, _lib.MODE_FP32))
    got = torch.full((n + 1,), float("nan"), device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = lib.dlas_mlp_grad_noalloc(_lib.int_array(dims), len(dims) - 1, loss, _lib.ptr(p), _lib.ptr(X), _lib.ptr(y),
                                   X.shape[0], _lib.ptr(got), _lib.stream_ptr())
    assert rc == 0, lib.dlas_last_error()
    torch.cuda.synchronize()
    taken = free0 - torch.cuda.mem_get_info()[0]     # device memory the call itself took (cudaMemGetInfo)
    return got.cpu().numpy(), ref.cpu().numpy(), taken


def test_two_calls_in_a_row_match_the_workspace_entry_point(env):
    _lib, ops = env
    dims = [8, 64, 64, 1]
    p = _cuda(O.glorot_init(dims, 11))
    # first call: 6000 rows (all CTAs busy); second call: 70 rows, other data, other loss -- only three
    # CTAs get a tile, the rest of the array still holds the first call's partials
    for call, (n_rows, start, loss) in enumerate(((6000, 0, 0), (70, 9000, 1))):
        Xh, yh = O.synth_rows(O.BASE_SEED, start, n_rows, dims[0])
        got, ref, taken = _both(_lib, ops, dims, loss, p, _cuda(Xh), _cuda(yh * 8))
        # The first call on a device may be the one that makes the driver load the library's module (kernels + the
        # built-in array: CUDA loads lazily); from the second call on, free device memory must not move at all.
        if call > 0:
            assert taken == 0, "dlas_mlp_grad_noalloc took %d bytes of device memory on its second call" % taken
        assert np.isfinite(got).all()
        assert got[-1] == pytest.approx(ref[-1], rel=TOL)          # loss sum
        assert _relerr(got[:-1], ref[:-1]) < TOL                    # gradient sums
        lr, gref = O.mlp_loss_grad_sums(p.cpu().numpy().astype(np.float64), dims, Xh.astype(np.float64),
                                        (yh * 8).astype(np.float64), loss)
        assert got[-1] == pytest.approx(lr, rel=TOL) and _relerr(got[:-1], gref) < TOL


def test_empty_batch_and_oversize_network(env):
    _lib, ops = env
    lib = _lib.load()
    dims = [8, 64, 64, 1]
    p = _cuda(O.glorot_init(dims, 11))
    out = torch.ones(ops.mlp_n_params(dims) + 1, device="cuda")
    rc = lib.dlas_mlp_grad_noalloc(_lib.int_array(dims), 3, 0, _lib.ptr(p), None, None, 0, _lib.ptr(out), _lib.stream_ptr())
    assert rc == 0 and float(out.abs().max()) == 0.0
    # 32-256-256-1 needs 44 MB of per-CTA partials: beyond the built-in 4 MiB, refused with a clear message
    big = [32, 256, 256, 1]
    pb = _cuda(O.glorot_init(big, 1))
    outb = torch.zeros(ops.mlp_n_params(big) + 1, device="cuda")
    X, y = torch.zeros((4, 32), device="cuda"), torch.zeros((4, 1), device="cuda")
    rc = lib.dlas_mlp_grad_noalloc(_lib.int_array(big), 3, 0, _lib.ptr(pb), _lib.ptr(X), _lib.ptr(y), 4, _lib.ptr(outb),
                                   _lib.stream_ptr())
    assert rc == -4
    msg = lib.dlas_last_error().decode()
    assert "DLAS_NOALLOC_SCRATCH_BYTES" in msg and "dlas_mlp_grad" in msg
