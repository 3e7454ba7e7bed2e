"""GPU parity of the tcgen05 kernel for the small static ANN d0-64-64-1 (BASELINE.json configs[0]: 8-64-64-1) through the
C ABI, DLAS_MODE_TC: fp16 operands, fp32 accumulation.  Same stated tolerances as the 32-256-256-1 kernel
(tests/test_mlp_tc_gpu.py): final weights 1e-4 (SGD), per-step loss 3e-4 from 4096 rows up, a single raw gradient 1e-2."""
import os

import numpy as np
import pytest
import torch

import dlas_oracle as O

pytestmark = pytest.mark.gpu
TOL, LOSS_TOL, GRAD_TOL = 1e-4, 3e-4, 1e-2


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


def _params(dims, seed):
    p = O.glorot_init(dims, seed)
    offs, _ = O.mlp_offsets(dims)
    for li, (w, b) in enumerate(offs):      # non-zero biases so that every term is exercised
        p[b:b + dims[li + 1]] = 0.05 * np.sin(1.0 + np.arange(dims[li + 1]) + li)
    return p


@pytest.mark.parametrize("d0", [8, 5, 12, 1, 15])
@pytest.mark.parametrize("n,loss", [(1, 0), (128, 1), (1000, 0), (37888 + 5, 0), (37888 + 5, 1), (65536, 0)])
def test_grad_vs_oracle(dl, d0, n, loss):
    from dlas import ops
    dims = [d0, 64, 64, 1]
    assert ops.mlp_mode_supported(dims, dl.MODE_TC)
    X, y = O.synth_rows(O.BASE_SEED, 31, n, d0)
    y = y * (8.0 if loss == O.LOSS_MAE_CLIP else 1.0)
    p = _params(dims, 21)
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    gs = torch.zeros(p.size + 1, device="cuda")
    ops.mlp_grad(dims, loss, dl.MODE_TC, _cuda(p), _cuda(X), _cuda(y), gs, ws)
    ops.mlp_status(dims, dl.MODE_TC, ws)
    g = gs.cpu().numpy()
    assert np.isfinite(g).all()
    lr, gref = O.mlp_loss_grad_sums(p.astype(np.float64), dims, X.astype(np.float64), y.astype(np.float64).reshape(-1, 1), loss)
    offs, _ = O.mlp_offsets(dims)
    errs = {}
    for li, (w, b) in enumerate(offs):
        nw = dims[li] * dims[li + 1]
        errs["W%d" % (li + 1)] = _relerr(g[w:w + nw], gref[w:w + nw])
        errs["b%d" % (li + 1)] = _relerr(g[b:b + dims[li + 1]], gref[b:b + dims[li + 1]])
    print("d0=%d n=%d loss=%d loss_rel=%.3g" % (d0, n, loss, abs(g[-1] / lr - 1)), errs)
    gtol = GRAD_TOL + (2.0 / n if loss == O.LOSS_MAE_CLIP else 0.0)
    assert g[-1] == pytest.approx(lr, rel=(LOSS_TOL if n >= 4096 else 1e-2) + (2.0 / n if loss else 0.0))
    if n >= 128:
        assert max(errs.values()) < gtol and _relerr(g[:-1], gref) < gtol
    # deterministic, and consistent with the fp32 kernel
    gs2 = torch.zeros_like(gs)
    ops.mlp_grad(dims, loss, dl.MODE_TC, _cuda(p), _cuda(X), _cuda(y), gs2, ws)
    assert torch.equal(gs, gs2)


@pytest.mark.parametrize("name,loss", [("sgd_mse", "mse"), ("adam_mse", "mse"), ("sgd_maeclip", "mae_clip")])
def test_fit_steps_golden(dl, golden_dir, name, loss):
    from dlas import ops
    g = np.load(os.path.join(golden_dir, "mlp_8_64_64_1_steps.npz"))
    dims = [8, 64, 64, 1]
    b, ns = int(g["batch"]), int(g["n_steps"])
    opt = dl.Adam() if name.startswith("adam") else dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True)
    p = _cuda(g["params0"])
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    gs = torch.zeros(p.numel() + 1, device="cuda")
    losses = torch.zeros(ns, device="cuda")
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    ops.mlp_fit_steps(dims, dl.LOSS_MSE if loss == "mse" else dl.LOSS_MAE_CLIP, dl.MODE_TC, opt.cfg(), p, m, v, _cuda(g["X"]),
                      _cuda(g["y"]), b, ns, 0, losses, gs, ws)
    ops.mlp_status(dims, dl.MODE_TC, ws)
    # batch 32: single rows' operand rounding is not averaged away (1e-2 on the loss below 4096 rows, as for predictions);
    # Adam's stated tolerance on weights is 2e-3 (tests/test_mlp_tc_gpu.py)
    np.testing.assert_allclose(losses.cpu().numpy(), g[name + "_losses"], rtol=1e-2)
    assert _relerr(p.cpu().numpy(), g[name + "_final"]) < (2e-3 if name.startswith("adam") else TOL)


def test_facade_epoch_vs_oracle(dl):
    """Sequential(mode='tc') on 8-64-64-1: 40 SGD-Nesterov steps (cnn.py:117) of batch 4096 through model.fit, the validation
    loss and predictions from the forward-only kinds of the same kernel."""
    dims, batch, nb = [8, 64, 64, 1], 4096, 40
    X, y = O.synth_rows(O.BASE_SEED, 0, batch * nb, 8)
    p0 = O.glorot_init(dims, O.BASE_SEED)
    model = dl.Sequential([dl.Dense(8, 64, activation="relu"), dl.Dense(64, 64, activation="relu"), dl.Dense(64, 1)], mode="tc")
    model.compile(loss="mse", optimizer=dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    model.set_weights_flat(p0)
    Xv, yv = O.synth_rows(O.BASE_SEED, 10_000_000, 20000, 8)
    hist = model.fit(X, y, validation_data=(Xv, yv), nb_epoch=1, batch_size=batch, shuffle=False, verbose=0)
    pr = p0.astype(np.float64)
    ref = O.mlp_fit_steps(pr, dims, X.astype(np.float64), y.astype(np.float64), batch, nb, O.LOSS_MSE, O.sgd_config(), O.opt_state(pr.size))
    assert hist["loss"][0] == pytest.approx(ref.mean(), rel=LOSS_TOL)
    assert _relerr(model.get_weights_flat(), pr) < TOL
    lv, _ = O.mlp_loss_grad_sums(pr, dims, Xv.astype(np.float64), yv.astype(np.float64), O.LOSS_MSE)
    assert hist["val_loss"][0] == pytest.approx(lv / 20000, rel=LOSS_TOL)
    yh = model.predict(Xv).cpu().numpy()
    yr = O.mlp_forward(pr, dims, Xv.astype(np.float64))[-1]
    assert np.abs(yh - yr).max() < 2e-3 * max(1.0, np.abs(yr).max())


def test_lag_windows_in_place(dl):
    """3-lag windows over a 4-feature series = a 12-64-64-1 dynamic ANN read in place (row stride 4 floats)."""
    from dlas import ops
    dims = [12, 64, 64, 1]
    s, ys = O.synth_rows(O.BASE_SEED, 0, 5000, 4)
    series, yw = _cuda(s), _cuda(ys[2:])
    W, V = ops.lag_windows(series, 3), dl.LagWindows(series, 3)
    p = _cuda(O.glorot_init(dims, 5))
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    ga, gb = (torch.zeros(p.numel() + 1, device="cuda") for _ in range(2))
    ops.mlp_grad(dims, 0, dl.MODE_TC, p, W, yw, ga, ws)
    ops.mlp_grad(dims, 0, dl.MODE_TC, p, V, yw, gb, ws)
    torch.cuda.synchronize()
    assert torch.isfinite(ga).all() and torch.equal(ga, gb)
