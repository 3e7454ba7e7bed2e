"""GPU parity of the tensor-core path for the wide-input ANN d0-256-256-1 (BASELINE.json configs[3], Readme.md:19: the dynamic
ANN on 24 lags x 8 features = 192 inputs) through the C ABI, DLAS_MODE_TC: layer 1 in mlp_tc_l1.cu (H1 tile images in, partial
dZ1 tile images out), layers 2 and 3 in the CTA-pair kernel of mlp_tc.cu.  fp16 operands, fp32 accumulation; the stated
tolerances are those of the 32-256-256-1 kernel (tests/test_mlp_tc_gpu.py): final weights 1e-4 (SGD) / 3e-3 (Adam), per-step
loss 3e-4 (SGD) / 5e-4 (Adam) from 4096 rows up, a single raw MSE gradient 1e-2 from 4096 rows up.  Below 4096 rows, and for the
+-1 output gradients of the clipped MAE, a raw gradient is held to 2e-2: the signal in dW1 is the correlation of one of 192
input columns with the back-propagated term, a sum that largely cancels, while every row's fp16 rounding (2^-11, one more than in
the fused kernel: the two CTAs' partial dZ1 tiles are added as packed halves) adds to it -- measured 1.0e-2 at 1000 rows,
5.7e-3 at 9477, 2.1e-3 at 65,536 (1 / sqrt(rows)), and 1.4e-2 for the clipped MAE at 4099 rows."""
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


def test_supported_shapes(dl):
    from dlas import ops
    for d0 in (64, 72, 192, 240):
        assert ops.mlp_mode_supported([d0, 256, 256, 1], dl.MODE_TC)
    for dims in ([60, 256, 256, 1], [248, 256, 256, 1], [192, 256, 128, 1], [192, 128, 256, 1], [192, 256, 256, 2]):
        assert not ops.mlp_mode_supported(dims, dl.MODE_TC)


@pytest.mark.parametrize("d0,n,loss", [(192, 1, 0), (192, 128, 1), (192, 1000, 0), (192, 9472 + 5, 0), (192, 9472 + 5, 1),
                                       (192, 65536, 0), (64, 4099, 0), (240, 4099, 1), (72, 19000, 0)])
def test_grad_vs_oracle(dl, d0, n, loss):
    from dlas import ops
    dims = [d0, 256, 256, 1]
    X, y = O.synth_rows(O.BASE_SEED, 31, n, d0)
    y = y * (8.0 if loss == O.LOSS_MAE_CLIP else 1.0)
    p = _params(dims, 21)
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    gs = torch.zeros(p.size + 1, device="cuda")
    Xd, yd, pd = _cuda(X), _cuda(y), _cuda(p)
    ops.mlp_grad(dims, loss, dl.MODE_TC, pd, Xd, yd, gs, ws)
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
    gtol = (GRAD_TOL if (n >= 4096 and loss == O.LOSS_MSE) else 2 * GRAD_TOL) + (2.0 / n if loss == O.LOSS_MAE_CLIP else 0.0)
    assert g[-1] == pytest.approx(lr, rel=(LOSS_TOL if n >= 4096 else 1e-2) + (2.0 / n if loss else 0.0))
    if n >= 128:
        assert max(errs.values()) < gtol and _relerr(g[:-1], gref) < gtol
    # deterministic, and close to the fp32 kernel
    gs2 = torch.zeros_like(gs)
    ops.mlp_grad(dims, loss, dl.MODE_TC, pd, Xd, yd, gs2, ws)
    assert torch.equal(gs, gs2)
    if loss == 0 and n >= 128:
        gf = torch.zeros_like(gs)
        ops.mlp_grad(dims, loss, dl.MODE_FP32, pd, Xd, yd, gf, ops.mlp_workspace(dims, dl.MODE_FP32))
        assert _relerr(g[:-1], gf.cpu().numpy()[:-1].astype(np.float64)) < gtol


@pytest.mark.parametrize("opt_name", ["sgd", "adam"])
def test_fit_steps_vs_oracle(dl, opt_name):
    """24 steps of batch 4096 on the 192-256-256-1 network through dlas_mlp_fit_steps (five launches per step, the weight image
    chained between steps) against the float64 oracle."""
    from dlas import ops
    dims, batch, ns = [192, 256, 256, 1], 4096, 24
    X, y = O.synth_rows(O.BASE_SEED, 0, batch * ns, 192)
    p0 = O.glorot_init(dims, O.BASE_SEED)
    opt = dl.Adam() if opt_name == "adam" else dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True)
    p = _cuda(p0)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    gs = torch.zeros(p.numel() + 1, device="cuda")
    losses = torch.zeros(ns, device="cuda")
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    ops.mlp_fit_steps(dims, dl.LOSS_MSE, dl.MODE_TC, opt.cfg(), p, m, v, _cuda(X), _cuda(y), batch, ns, 0, losses, gs, ws)
    ops.mlp_status(dims, dl.MODE_TC, ws)
    pr = p0.astype(np.float64)
    cfg = O.adam_config() if opt_name == "adam" else O.sgd_config()
    ref = O.mlp_fit_steps(pr, dims, X.astype(np.float64), y.astype(np.float64), batch, ns, O.LOSS_MSE, cfg, O.opt_state(pr.size))
    got = losses.cpu().numpy()
    print("%s: max per-step loss rel err %.3g; weights rel err %.3g" % (opt_name, np.abs(got / ref - 1).max(), _relerr(p.cpu().numpy(), pr)))
    np.testing.assert_allclose(got, ref, rtol=5e-4 if opt_name == "adam" else LOSS_TOL)
    # Adam: measured 1.97e-3 on the weights of this shape (3.3e-4 on the losses); see test_adam_batch_4096 of the 32-input kernel
    assert _relerr(p.cpu().numpy(), pr) < (3e-3 if opt_name == "adam" else TOL)


def test_lag_windows_in_place(dl):
    """The dynamic ANN's input proper: 24-lag windows over an 8-feature series read in place (row stride 8 floats); the
    gradient is bit-identical to the one of the materialised window matrix, and training through the facade matches the oracle."""
    from dlas import ops
    dims = [192, 256, 256, 1]
    s, ys = O.synth_rows(O.BASE_SEED, 0, 20000 + 23, 8)
    series, yw = _cuda(s), _cuda(ys[23:])
    W, V = ops.lag_windows(series, 24), dl.LagWindows(series, 24)
    p = _cuda(O.glorot_init(dims, 5))
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    ga, gb = (torch.zeros(p.numel() + 1, device="cuda") for _ in range(2))
    ops.mlp_grad(dims, 0, dl.MODE_TC, p, W, yw, ga, ws)
    ops.mlp_grad(dims, 0, dl.MODE_TC, p, V, yw, gb, ws)
    torch.cuda.synchronize()
    assert torch.isfinite(ga).all() and torch.equal(ga, gb)


def test_facade_epoch_vs_oracle(dl):
    """Sequential(mode='tc') on 192-256-256-1: one epoch of 10 steps of batch 4096 through model.fit; the validation loss and
    the predictions run on the tensor cores too (dlas_mlp_eval_loss / dlas_mlp_predict_ws: l1_fwd + the forward-only kinds of
    the pair kernel)."""
    dims, batch, nb = [192, 256, 256, 1], 4096, 10
    X, y = O.synth_rows(O.BASE_SEED, 0, batch * nb, 192)
    p0 = O.glorot_init(dims, O.BASE_SEED)
    model = dl.Sequential([dl.Dense(192, 256, activation="relu"), dl.Dense(256, 256, activation="relu"), dl.Dense(256, 1)], mode="tc")
    model.compile(loss="mse", optimizer=dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    model.set_weights_flat(p0)
    Xv, yv = O.synth_rows(O.BASE_SEED, 10_000_000, 5000, 192)
    hist = model.fit(X, y, validation_data=(Xv, yv), nb_epoch=1, batch_size=batch, shuffle=False, verbose=0)
    pr = p0.astype(np.float64)
    ref = O.mlp_fit_steps(pr, dims, X.astype(np.float64), y.astype(np.float64), batch, nb, O.LOSS_MSE, O.sgd_config(), O.opt_state(pr.size))
    assert hist["loss"][0] == pytest.approx(ref.mean(), rel=LOSS_TOL)
    assert _relerr(model.get_weights_flat(), pr) < TOL
    lv, _ = O.mlp_loss_grad_sums(pr, dims, Xv.astype(np.float64), yv.astype(np.float64), O.LOSS_MSE)
    assert hist["val_loss"][0] == pytest.approx(lv / 5000, rel=LOSS_TOL)
    yh = model.predict(Xv).cpu().numpy()
    yr = O.mlp_forward(pr, dims, Xv.astype(np.float64))[-1]
    assert np.abs(yh - yr).max() < 2e-3 * max(1.0, np.abs(yr).max())


@pytest.mark.parametrize("d0,n", [(192, 1), (192, 127), (64, 9472 + 77), (192, 65536 + 300), (240, 2 * 65536 + 5)])
def test_predict_and_eval_loss_on_the_tensor_cores(dl, d0, n):
    """Forward-only kinds for d0-256-256-1 through the C ABI: any number of rows (65,536 per pass), ragged tails, lag windows
    in place.  Stated tolerance as for 32-256-256-1: predictions within 2e-3 of max|yhat| of the fp64 oracle, loss sums within
    3e-4 from 4096 rows up (1e-2 below); without a workspace dlas_mlp_predict serves the shape from the fp32 kernels."""
    from dlas import ops
    dims = [d0, 256, 256, 1]
    X, y = O.synth_rows(O.BASE_SEED, 77, n, d0)
    p = _params(dims, 21)
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    Xd, yd, pd = _cuda(X), _cuda(y), _cuda(p)
    yh = ops.mlp_predict(dims, dl.MODE_TC, pd, Xd, workspace=ws).cpu().numpy().ravel()
    ls = float(ops.mlp_eval_loss(dims, dl.LOSS_MSE, dl.MODE_TC, pd, Xd, yd, ws).item())
    ops.mlp_status(dims, dl.MODE_TC, ws)
    yr = O.mlp_forward(p.astype(np.float64), dims, X.astype(np.float64))[-1].ravel()
    lr, _ = O.mlp_loss_grad_sums(p.astype(np.float64), dims, X.astype(np.float64), y.astype(np.float64).reshape(-1, 1), O.LOSS_MSE)
    assert np.isfinite(yh).all() and np.abs(yh - yr).max() < 2e-3 * max(1.0, np.abs(yr).max())
    if n >= 100:     # (the squared error of one well-predicted row is ill-conditioned: its prediction is checked above)
        assert ls == pytest.approx(lr, rel=LOSS_TOL if n >= 4096 else 1e-2)
    # deterministic; and the entry point without a workspace still answers (fp32 kernels)
    yh2 = ops.mlp_predict(dims, dl.MODE_TC, pd, Xd, workspace=ws).cpu().numpy().ravel()
    assert np.array_equal(yh, yh2)
    yf = ops.mlp_predict(dims, dl.MODE_TC, pd, Xd).cpu().numpy().ravel()
    assert np.abs(yf - yr).max() < 1e-4 * max(1.0, np.abs(yr).max())


def test_predict_lag_windows_in_place(dl):
    from dlas import ops
    dims = [192, 256, 256, 1]
    s, _ = O.synth_rows(O.BASE_SEED, 0, 70000 + 23, 8)
    series = _cuda(s)
    W, V = ops.lag_windows(series, 24), dl.LagWindows(series, 24)
    p = _cuda(O.glorot_init(dims, 5))
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    ya = ops.mlp_predict(dims, dl.MODE_TC, p, W, workspace=ws)
    yb = ops.mlp_predict(dims, dl.MODE_TC, p, V, workspace=ws)
    torch.cuda.synchronize()
    assert torch.isfinite(ya).all() and torch.equal(ya, yb)


def test_rows_per_call_limit(dl):
    from dlas import ops
    dims = [192, 256, 256, 1]
    n = 65536 + 128
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    gs = torch.zeros(int(ops.mlp_n_params(dims)) + 1, device="cuda")
    with pytest.raises(dl.DlasError) as e:
        ops.mlp_grad(dims, 0, dl.MODE_TC, torch.zeros(gs.numel() - 1, device="cuda"), torch.zeros(n, 192, device="cuda"),
                     torch.zeros(n, device="cuda"), gs, ws)
    assert "at most" in str(e.value)


@pytest.mark.parametrize("mode", ["fp32", "tc"])
@pytest.mark.parametrize("name", ["sgd_mse", "adam_mse"])
def test_dynamic_ann_golden(dl, golden_dir, mode, name):
    """The committed configs[3] fixture (tests/golden/make_golden.py: raw (n + 23, 8) series, 24-lag windows, 4 steps of 128):
    windows read in place from the series through dlas_mlp_fit_steps_strided.  fp32 mode: loss and weights within 1e-4;
    tensor-core mode at this batch of 128: loss 1e-2 (single rows' operand rounding is not averaged away below 4096 rows),
    weights 1e-4 (SGD) / 3e-3 (Adam)."""
    import os
    from dlas import ops
    g = np.load(os.path.join(golden_dir, "dynamic_ann_192_256_256_1_steps.npz"))
    dims, lags, b, ns = g["dims"].tolist(), int(g["lags"]), int(g["batch"]), int(g["n_steps"])
    md = dl.MODE_TC if mode == "tc" else dl.MODE_FP32
    opt = dl.Adam() if name.startswith("adam") else dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True)
    V = dl.LagWindows(_cuda(g["series"]), lags)
    p = _cuda(g["params0"])
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    gs = torch.zeros(p.numel() + 1, device="cuda")
    losses = torch.zeros(ns, device="cuda")
    ws = ops.mlp_workspace(dims, md)
    ops.mlp_fit_steps(dims, dl.LOSS_MSE, md, opt.cfg(), p, m, v, V, _cuda(g["y"]), b, ns, 0, losses, gs, ws)
    ops.mlp_status(dims, md, ws)
    ltol, wtol = (1e-4, 1e-4) if mode == "fp32" else (1e-2, 3e-3 if name.startswith("adam") else 1e-4)
    np.testing.assert_allclose(losses.cpu().numpy(), g[name + "_losses"], rtol=ltol)
    assert _relerr(p.cpu().numpy(), g[name + "_final"].astype(np.float64)) < wtol
    # first-batch gradient of the fixture
    ops.mlp_grad(dims, dl.LOSS_MSE, md, _cuda(g["params0"]), dl.LagWindows(_cuda(g["series"][:b + lags - 1]), lags), _cuda(g["y"][:b]), gs, ws)
    gg = gs.cpu().numpy()
    assert gg[-1] == pytest.approx(float(g["grad0_loss_sum"]), rel=ltol)
    # a raw gradient of 128 rows in tensor-core mode: 1 / sqrt(rows) from the 1.0e-2 measured at 1000 rows (module docstring)
    assert _relerr(gg[:-1], g["grad0"].astype(np.float64)) < (1e-4 if mode == "fp32" else 3e-2)
