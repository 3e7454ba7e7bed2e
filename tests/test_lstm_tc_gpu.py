"""GPU parity of the tensor-core stages of the LSTM path (DLAS_MODE_TC: hidden 128, d_in <= 16; fp16 operands, fp32
accumulation, fp32 cell state).  Every stage can be exchanged for its fp32 counterpart (mode 0x100 | stage bits: 1 forward
recurrence, 2 backward recurrence, 4 weight gradient), so each is first compared with the fp32 kernels on the same stash,
then the whole mode with the fp64 oracle.  Stated tolerances: a stage against fp32 2e-3 on gradient blocks (operand
rounding 2^-11, averaged over the rows of the batch), the mode against the oracle 5e-3 on gradients and 1e-3 on the loss."""
import numpy as np
import pytest
import torch

import dlas_oracle as O

pytestmark = pytest.mark.gpu
TC_FWD, TC_BWD, TC_DW = 1, 2, 4


@pytest.fixture(scope="module")
def dl():
    import __graft_entry__ as G
    G.build()
    import dlas
    return dlas


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()


def _relerr(a, b):
    return float(np.linalg.norm(np.asarray(a, np.float64) - b) / max(np.linalg.norm(b), 1e-300))


def _setup(d, h, nl, T, B, seed):
    rng = np.random.default_rng(seed)
    p = O.lstm_init(d, h, nl, 100 + seed)
    p = p + 0.02 * rng.standard_normal(p.size).astype(np.float32)      # non-zero biases
    X = rng.standard_normal((B, T, d)).astype(np.float32)
    y = rng.standard_normal((B, 1)).astype(np.float32)
    return p, X, y


def _grad(mode, d, h, nl, T, loss, p, X, y):
    from dlas import ops
    ws = ops.lstm_workspace(d, h, nl, T, X.shape[0])
    gs = torch.zeros(ops.lstm_n_params(d, h, nl) + 1, device="cuda")
    ops.lstm_grad(d, h, nl, T, loss, _cuda(p), _cuda(X), _cuda(y), gs, ws, mode=mode)
    torch.cuda.synchronize()
    g = gs.cpu().numpy()
    assert np.isfinite(g).all()
    return g


def _blocks(d, h, nl, g, ref):
    offs, head, n = O.lstm_offsets(d, h, nl)
    out = {}
    for li, (k, b, dl_) in enumerate(offs):
        nk = (dl_ + h) * 4 * h
        out["K%d" % li] = _relerr(g[k:k + nk], ref[k:k + nk])
        out["b%d" % li] = _relerr(g[b:b + 4 * h], ref[b:b + 4 * h])
    return out


@pytest.mark.parametrize("shape", [(8, 128, 2, 48, 300), (8, 128, 2, 48, 37), (5, 128, 1, 3, 1), (16, 128, 3, 6, 700)])
def test_weight_gradient_stage_vs_fp32(dl, shape):
    """dK, db from the tcgen05 GEMM over all (sequence, t) rows vs the fp32 kernel on the same stash: the other entries
    (head, loss) come from the same fp32 kernels and must be bit-identical."""
    from dlas import ops
    d, h, nl, T, B = shape
    assert ops.lstm_mode_supported(d, h, nl, dl.MODE_TC) and not ops.lstm_mode_supported(d, 64, nl, dl.MODE_TC)
    p, X, y = _setup(d, h, nl, T, B, B)
    ref = _grad(dl.MODE_FP32, d, h, nl, T, 0, p, X, y)
    got = _grad(0x100 | TC_DW, d, h, nl, T, 0, p, X, y)
    errs = _blocks(d, h, nl, got[:-1], ref[:-1].astype(np.float64))
    print(shape, errs)
    _, (hw, hb), n = O.lstm_offsets(d, h, nl)
    assert np.array_equal(got[hw:], ref[hw:])          # head weights, head bias, loss sum
    tol = 2e-3 if B * T >= 1000 else 1e-2
    assert max(errs.values()) < tol
    again = _grad(0x100 | TC_DW, d, h, nl, T, 0, p, X, y)
    assert np.array_equal(got, again)                  # deterministic


def test_mode_tc_vs_oracle(dl):
    d, h, nl, T, B = 8, 128, 2, 24, 256
    p, X, y = _setup(d, h, nl, T, B, 7)
    for loss in (O.LOSS_MSE, O.LOSS_MAE_CLIP):
        yy = y * (8.0 if loss == O.LOSS_MAE_CLIP else 1.0)
        got = _grad(dl.MODE_TC, d, h, nl, T, loss, p, X, yy)
        lr, gref = O.lstm_loss_grad_sums(p.astype(np.float64), d, h, nl, X.astype(np.float64), yy.astype(np.float64), loss)
        print("loss", loss, abs(got[-1] / lr - 1), _relerr(got[:-1], gref))
        assert got[-1] == pytest.approx(lr, rel=1e-3)
        assert _relerr(got[:-1], gref) < 5e-3 + (2.0 / B if loss else 0.0)


@pytest.mark.parametrize("shape", [(8, 128, 2, 48, 300), (8, 128, 2, 48, 37), (5, 128, 1, 3, 1), (16, 128, 3, 6, 700), (8, 128, 2, 5, 9472 + 128 + 3)])
def test_forward_recurrence_stage_vs_fp32(dl, shape):
    """The persistent tcgen05 recurrence (state on chip, CTA pair exchanging h through distributed shared memory) against
    the fp32 forward: predictions within 2e-3 of max|yhat| (fp16 operands, recurrence of T steps), and the gradient of the
    step whose BPTT / weight gradient stay fp32 but consume the tensor-core stash within 1e-2."""
    from dlas import ops
    d, h, nl, T, B = shape
    p, X, y = _setup(d, h, nl, T, B, B + 1)
    ws = ops.lstm_workspace(d, h, nl, T, B)
    ref = ops.lstm_predict(d, h, nl, T, _cuda(p), _cuda(X), ws).cpu().numpy()
    got = ops.lstm_predict(d, h, nl, T, _cuda(p), _cuda(X), ws, mode=0x100 | TC_FWD).cpu().numpy()
    assert np.isfinite(got).all()
    print(shape, "max |dyhat| / max|yhat| = %.3g" % (np.abs(got - ref).max() / np.abs(ref).max()))
    assert np.abs(got - ref).max() < 2e-3 * max(1.0, np.abs(ref).max())
    again = ops.lstm_predict(d, h, nl, T, _cuda(p), _cuda(X), ws, mode=0x100 | TC_FWD).cpu().numpy()
    assert np.array_equal(got, again)
    if B <= 1000:
        g32 = _grad(dl.MODE_FP32, d, h, nl, T, 0, p, X, y)
        gtc = _grad(0x100 | TC_FWD, d, h, nl, T, 0, p, X, y)
        assert gtc[-1] == pytest.approx(g32[-1], rel=2e-3)
        assert _relerr(gtc[:-1], g32[:-1].astype(np.float64)) < (1e-2 if B * T >= 1000 else 5e-2)


@pytest.mark.parametrize("shape", [(8, 128, 2, 48, 300), (8, 128, 2, 48, 37), (5, 128, 1, 3, 1), (16, 128, 3, 6, 700), (8, 128, 4, 5, 130)])
@pytest.mark.parametrize("loss", [O.LOSS_MSE, O.LOSS_MAE_CLIP])
def test_backward_recurrence_stage_vs_fp32(dl, shape, loss):
    """BPTT on the tensor cores (dz K^T per step, partial dh exchanged inside the CTA pair) on the fp32 forward's stash,
    weight gradient in fp32: the loss sum is computed in fp32 from the same predictions (bit-identical), gradients
    within 5e-3 per block (fp16 operands through T steps of recurrence)."""
    d, h, nl, T, B = shape
    p, X, y = _setup(d, h, nl, T, B, B + 2)
    yy = y * (8.0 if loss == O.LOSS_MAE_CLIP else 1.0)
    ref = _grad(dl.MODE_FP32, d, h, nl, T, loss, p, X, yy)
    got = _grad(0x100 | TC_BWD, d, h, nl, T, loss, p, X, yy)
    errs = _blocks(d, h, nl, got[:-1], ref[:-1].astype(np.float64))
    _, (hw, hb), n = O.lstm_offsets(d, h, nl)
    print(shape, loss, errs, "head", _relerr(got[hw:hw + h], ref[hw:hw + h].astype(np.float64)))
    assert got[-1] == pytest.approx(ref[-1], rel=1e-6)
    assert _relerr(got[hw:hw + h + 1], ref[hw:hw + h + 1].astype(np.float64)) < 1e-5      # head: fp32, another summation order
    assert max(errs.values()) < (5e-3 if B * T >= 1000 else 3e-2)
    again = _grad(0x100 | TC_BWD, d, h, nl, T, loss, p, X, yy)
    assert np.array_equal(got, again)


# the last two shapes hold more 128-sequence tiles than the layer pipeline has lanes (74 SM pairs / layers): tiles queue up per lane
@pytest.mark.parametrize("shape", [(8, 128, 2, 48, 300), (16, 128, 3, 6, 77), (3, 128, 1, 9, 1), (8, 128, 2, 4, 9472 + 128 + 45), (8, 128, 3, 3, 3072 + 300)])
def test_whole_mode_on_ragged_batches_vs_fp32(dl, shape):
    """DLAS_MODE_TC end to end (all three stages, the tensor-core kernels' own stash layout) against the fp32 path on
    batches that are no multiple of the 128-sequence tile / the 32-sequence layout block."""
    from dlas import ops
    d, h, nl, T, B = shape
    p, X, y = _setup(d, h, nl, T, B, B + 3)
    ref = _grad(dl.MODE_FP32, d, h, nl, T, 0, p, X, y)
    got = _grad(dl.MODE_TC, d, h, nl, T, 0, p, X, y)
    errs = _blocks(d, h, nl, got[:-1], ref[:-1].astype(np.float64))
    print(shape, errs)
    assert got[-1] == pytest.approx(ref[-1], rel=2e-3)
    assert max(errs.values()) < (1e-2 if B * T >= 1000 else 5e-2)
    ws = ops.lstm_workspace(d, h, nl, T, B)
    y32 = ops.lstm_predict(d, h, nl, T, _cuda(p), _cuda(X), ws).cpu().numpy()
    ytc = ops.lstm_predict(d, h, nl, T, _cuda(p), _cuda(X), ws, mode=dl.MODE_TC).cpu().numpy()
    assert np.abs(ytc - y32).max() < 2e-3 * max(1.0, np.abs(y32).max())
