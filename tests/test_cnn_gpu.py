"""GPU parity of the script's own CNN (cnn.py:110-114: Convolution1D -> Dropout -> Flatten -> Dense) through
the C ABI against the fp64 oracle.  Tolerance (BASELINE.json): 1e-4 relative, fp32 path.  The dropout mask
is a counter hash both sides evaluate, so train-mode gradients are compared with dropout ON."""
import numpy as np
import pytest
import torch

import dlas_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-4
CNN = (48, 13, 100, 12)      # cnn.py:111-114


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


def _data(shape, B, seed):
    L, K, F, n_out = shape
    rng = np.random.default_rng(seed)
    p = O.cnn_init(L, K, F, n_out, seed, dtype=np.float32)
    (wc, bc, wd, bd), n = O.cnn_offsets(L, K, F, n_out)
    p[bc:bc + F] = rng.normal(0, 0.05, F)
    p[bd:] = rng.normal(0, 0.05, n_out)
    X = rng.normal(0, 1, (B, L)).astype(np.float32)
    y = rng.normal(0, 3, (B, n_out)).astype(np.float32)
    return p, X, y


def _grad(shape, loss, p, X, y, keep=1.0, seed=0, sample0=0):
    from dlas import ops
    ws = ops.cnn_workspace(*shape, X.shape[0])
    gs = torch.full((ops.cnn_n_params(*shape) + 1,), 7.0, device="cuda")
    ops.cnn_grad(*shape, loss, _cuda(p), _cuda(X), _cuda(y), gs, ws, keep=keep, dropout_seed=seed, sample0=sample0)
    torch.cuda.synchronize()
    g = gs.cpu().numpy()
    return g[-1], g[:-1]


def test_n_params(dl):
    from dlas import ops
    assert ops.cnn_n_params(*CNN) == 44612 == O.cnn_offsets(*CNN)[1]
    with pytest.raises(Exception):
        ops.cnn_n_params(10, 13, 4, 2)        # kernel longer than the series
    with pytest.raises(Exception):
        ops.cnn_n_params(48, 13, 100, 17)     # more outputs than the kernels hold


@pytest.mark.parametrize("shape,B", [(CNN, 20), (CNN, 1), (CNN, 333), ((9, 3, 5, 2), 7), ((16, 16, 7, 1), 65), ((20, 1, 3, 16), 130)])
@pytest.mark.parametrize("loss", [O.LOSS_MSE, O.LOSS_MAE_CLIP])
def test_grad_vs_oracle_no_dropout(dl, shape, B, loss):
    p, X, y = _data(shape, B, 11 + B)
    ls, gr = _grad(shape, loss, p, X, y)
    ls64, g64 = O.cnn_loss_grad_sums(p.astype(np.float64), *shape, X.astype(np.float64), y, loss)
    assert ls == pytest.approx(ls64, rel=TOL)
    assert _relerr(gr, g64) < TOL


@pytest.mark.parametrize("shape,B,sample0", [(CNN, 20, 0), (CNN, 20, 20 * 12345), ((9, 3, 5, 2), 33, 7)])
@pytest.mark.parametrize("keep", [0.5, 0.8])
def test_grad_vs_oracle_dropout(dl, shape, B, sample0, keep):
    L, K, F, n_out = shape
    p, X, y = _data(shape, B, 5)
    mask = O.dropout_mask(99, sample0, B, (L - K + 1) * F, keep=np.float32(keep))
    assert abs(mask.mean() - keep) < 0.05
    ls, gr = _grad(shape, O.LOSS_MAE_CLIP, p, X, y, keep=keep, seed=99, sample0=sample0)
    ls64, g64 = O.cnn_loss_grad_sums(p.astype(np.float64), *shape, X.astype(np.float64), y, O.LOSS_MAE_CLIP, mask=mask, keep=float(np.float32(keep)))
    assert ls == pytest.approx(ls64, rel=TOL)
    assert _relerr(gr, g64) < TOL
    # a different step draws a different mask
    _, gr2 = _grad(shape, O.LOSS_MAE_CLIP, p, X, y, keep=keep, seed=99, sample0=sample0 + B)
    assert _relerr(gr2, g64) > 1e-2


def test_grad_is_deterministic_and_empty_batch(dl):
    p, X, y = _data(CNN, 100, 3)
    a = _grad(CNN, O.LOSS_MSE, p, X, y, keep=0.5, seed=1)
    b = _grad(CNN, O.LOSS_MSE, p, X, y, keep=0.5, seed=1)
    assert a[0] == b[0] and np.array_equal(a[1], b[1])
    ls, gr = _grad(CNN, O.LOSS_MSE, p, X[:0], y[:0])
    assert ls == 0.0 and not gr.any()


def test_predict_is_test_mode(dl):
    from dlas import ops
    p, X, _ = _data(CNN, 50, 2)
    ws = ops.cnn_workspace(*CNN, 50)
    yh = ops.cnn_predict(*CNN, _cuda(p), _cuda(X), ws).cpu().numpy()
    ref = O.cnn_forward(p.astype(np.float64), *CNN, X.astype(np.float64))[0]
    assert _relerr(yh, ref) < TOL


def test_workspace_too_small_is_an_error(dl):
    from dlas import ops, _lib
    p, X, y = _data(CNN, 20, 2)
    gs = torch.zeros(44613, device="cuda")
    with pytest.raises(_lib.DlasError):
        ops.cnn_grad(*CNN, 0, _cuda(p), _cuda(X), _cuda(y), gs, torch.empty(1024, dtype=torch.uint8, device="cuda"))


def test_sgd_steps_match_oracle(dl):
    """cnn.py:117-128 at its own batch (20): ten SGD steps with dropout against the fp64 oracle loop."""
    from dlas import ops
    L, K, F, n_out = CNN
    B, steps = 20, 10
    p0, X, y = _data(CNN, B * steps, 8)
    ocfg = O.sgd_config(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True)      # cnn.py:117
    cfg = ops.make_opt_cfg(O.OPT_SGD, 0.001, momentum=0.99, decay=1e-6, nesterov=True)
    loss_out = torch.zeros(steps, device="cuda")
    p = _cuda(p0); m = torch.zeros_like(p); v = torch.zeros_like(p)
    gs = torch.zeros(p.numel() + 1, device="cuda")
    ws = ops.cnn_workspace(*CNN, B)
    Xd, yd = _cuda(X), _cuda(y)
    p64 = p0.astype(np.float64); st = O.opt_state(p64.size)
    losses64 = []
    for it in range(steps):
        sl = slice(it * B, (it + 1) * B)
        ops.cnn_grad(*CNN, O.LOSS_MAE_CLIP, p, Xd[sl], yd[sl], gs, ws, keep=0.5, dropout_seed=4, sample0=it * B)
        ops.opt_step(cfg, p, m, v, gs, 1.0 / B, it, loss_out[it:it + 1])
        mask = O.dropout_mask(4, it * B, B, (L - K + 1) * F, keep=0.5)
        ls, g = O.cnn_loss_grad_sums(p64, *CNN, X[sl].astype(np.float64), y[sl], O.LOSS_MAE_CLIP, mask=mask, keep=0.5)
        p64 = O.opt_step(p64, g / B, ocfg, st)
        losses64.append(ls / B)
    torch.cuda.synchronize()
    np.testing.assert_allclose(loss_out.cpu().numpy(), losses64, rtol=TOL)
    assert _relerr(p.cpu().numpy(), p64) < TOL


def test_keras_facade_builds_the_script_model(dl):
    """cnn.py:110-118 verbatim through the facade: shape inferred from Dense(3600, 12); training lowers the
    loss on a learnable target; predict/evaluate run in test mode (dropout off)."""
    import dlas
    model = dlas.Sequential(seed=5)
    model.add(dlas.Convolution1D(1, 100, 13, activation="relu"))
    model.add(dlas.Dropout(0.5))
    model.add(dlas.Flatten())
    model.add(dlas.Dense(3600, 12))
    model.compile(loss=dlas.mae_clip, optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    assert model._impl.shape == CNN and model._impl.n_params == 44612
    p0 = model.get_weights_flat()
    np.testing.assert_array_equal(p0, O.cnn_init(*CNN, 5, dtype=np.float32))
    rng = np.random.default_rng(0)
    t = np.arange(60)[None, :] * 0.3 + rng.uniform(0, 6.28, (2000, 1))
    W = np.sin(t).astype(np.float32)
    X, y = W[:, :48], W[:, 48:]
    before = model.evaluate(X, y)
    assert before == pytest.approx(model.evaluate(X, y))                       # test mode is deterministic
    ref = O.cnn_loss_grad_sums(p0.astype(np.float64), *CNN, X.astype(np.float64), y, O.LOSS_MAE_CLIP)[0] / 2000
    assert before == pytest.approx(ref, rel=TOL)
    hist = model.fit(X, y, nb_epoch=3, batch_size=20, verbose=0)
    assert len(hist["loss"]) == 3 and model.evaluate(X, y) < 0.5 * before
    assert model.predict(X[:7]).shape == (7, 12)
    with pytest.raises(NotImplementedError):
        m2 = dlas.Sequential(); m2.add(dlas.Convolution1D(1, 4, 3, activation="relu")); m2.add(dlas.Dense(8, 1)); m2.compile()


def test_fit_steps_matches_the_step_by_step_loop(dl):
    """dlas_cnn_fit_steps (reduction fused with the update, batch order from the host) against dlas_cnn_grad +
    dlas_opt_step called batch by batch: same masks (sample0 = iteration * batch), same losses, same weights."""
    from dlas import ops
    B, nb = 20, 6
    p0, X, y = _data(CNN, B * nb, 9)
    cfg = ops.make_opt_cfg(O.OPT_SGD, 0.001, momentum=0.99, decay=1e-6, nesterov=True)
    order = [3, 0, 5, 1, 4, 2]
    it0 = 11
    Xd, yd = _cuda(X), _cuda(y)
    ws = ops.cnn_workspace(*CNN, B)
    # step by step
    p = _cuda(p0); m = torch.zeros_like(p); v = torch.zeros_like(p)
    gs = torch.zeros(p.numel() + 1, device="cuda")
    ref = torch.zeros(nb, device="cuda")
    for s, b in enumerate(order):
        sl = slice(b * B, (b + 1) * B)
        ops.cnn_grad(*CNN, O.LOSS_MAE_CLIP, p, Xd[sl], yd[sl], gs, ws, keep=0.5, dropout_seed=4, sample0=(it0 + s) * B)
        ops.opt_step(cfg, p, m, v, gs, 1.0 / B, it0 + s, ref[s:s + 1])
    # one call
    p1 = _cuda(p0); m1 = torch.zeros_like(p1); v1 = torch.zeros_like(p1)
    got = torch.zeros(nb, device="cuda")
    ops.cnn_fit_steps(*CNN, O.LOSS_MAE_CLIP, cfg, p1, m1, v1, Xd, yd, B, order, it0, 0.5, 4, got, ws)
    torch.cuda.synchronize()
    np.testing.assert_allclose(got.cpu().numpy(), ref.cpu().numpy(), rtol=1e-6)
    assert _relerr(p1.cpu().numpy(), p.cpu().numpy().astype(np.float64)) < 1e-6
    with pytest.raises(Exception):
        ops.cnn_fit_steps(*CNN, O.LOSS_MAE_CLIP, cfg, p1, m1, v1, Xd, yd, B, [nb], it0, 0.5, 4, got, ws)   # batch index out of range
