"""GPU parity of the MLP path (static / dynamic ANN) through the C ABI against the fp64 oracle and
the committed golden vectors.  Tolerance from BASELINE.json: per-step loss and final weights within
1e-4 relative for the fp32/TF32 path (weights: ||w - w*||_2 / ||w*||_2)."""
import os

import numpy as np
import pytest
import torch

import dlas_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-4


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


def _grad(dl, dims, loss, p, X, y, mode=0):
    from dlas import ops
    ws = ops.mlp_workspace(dims, mode)
    gs = torch.zeros(ops.mlp_n_params(dims) + 1, device="cuda")
    ops.mlp_grad(dims, loss, mode, _cuda(p), _cuda(X), _cuda(y), gs, ws)
    torch.cuda.synchronize()
    g = gs.cpu().numpy()
    return g[-1], g[:-1]


@pytest.mark.parametrize("tag", ["8_64_64_1", "32_256_256_1"])
def test_grad_golden(dl, golden_dir, tag):
    g = np.load(os.path.join(golden_dir, "mlp_%s_steps.npz" % tag))
    dims, b = g["dims"].tolist(), int(g["batch"])
    ls, gr = _grad(dl, dims, O.LOSS_MSE, g["params0"], g["X"][:b], g["y"][:b])
    assert ls == pytest.approx(float(g["grad0_loss_sum"]), rel=TOL)
    assert _relerr(gr, g["grad0"]) < TOL


@pytest.mark.parametrize("dims,n", [([8, 64, 64, 1], 1000), ([32, 256, 256, 1], 777), ([192, 256, 256, 1], 300),
                                    ([5, 7, 3], 65), ([3, 1], 31), ([16, 300, 12], 100), ([8, 64, 64, 1], 1)])
@pytest.mark.parametrize("loss", [O.LOSS_MSE, O.LOSS_MAE_CLIP])
def test_grad_vs_oracle_ragged_shapes(dl, dims, n, loss):
    rng = np.random.default_rng(n)
    p = (rng.normal(size=O.mlp_n_params(dims)) * 0.2).astype(np.float32)
    X = rng.normal(size=(n, dims[0])).astype(np.float32)
    y = (rng.normal(size=(n, dims[-1])) * 5).astype(np.float32)
    ls, gr = _grad(dl, dims, loss, p, X, y)
    lr, gref = O.mlp_loss_grad_sums(p.astype(np.float64), dims, X.astype(np.float64), y.astype(np.float64), loss)
    assert ls == pytest.approx(lr, rel=TOL)
    assert _relerr(gr, gref) < TOL


def test_grad_empty_batch_is_zero(dl):
    from dlas import ops
    dims = [8, 64, 64, 1]
    ws = ops.mlp_workspace(dims, 0)
    gs = torch.ones(ops.mlp_n_params(dims) + 1, device="cuda")
    ops.mlp_grad(dims, 0, 0, _cuda(O.glorot_init(dims, 1)), torch.empty((0, 8), device="cuda"),
                 torch.empty((0, 1), device="cuda"), gs, ws, n_rows=0)
    assert float(gs.abs().max()) == 0.0


def test_grad_is_deterministic(dl):
    dims = [32, 256, 256, 1]
    X, y = O.synth_rows(O.BASE_SEED, 0, 5000, 32)
    p = O.glorot_init(dims, 3)
    a = _grad(dl, dims, 0, p, X, y)
    b = _grad(dl, dims, 0, p, X, y)
    assert a[0] == b[0] and np.array_equal(a[1], b[1])


def test_workspace_too_small_is_an_error(dl):
    from dlas import ops, _lib
    dims = [8, 64, 64, 1]
    with pytest.raises(_lib.DlasError):
        ops.mlp_grad(dims, 0, 0, _cuda(O.glorot_init(dims, 1)), torch.zeros((4, 8), device="cuda"),
                     torch.zeros((4, 1), device="cuda"), torch.zeros(4802, device="cuda"),
                     torch.empty(16, dtype=torch.uint8, device="cuda"))


@pytest.mark.parametrize("tag", ["8_64_64_1", "32_256_256_1"])
@pytest.mark.parametrize("name,loss", [("sgd_mse", "mse"), ("adam_mse", "mse"), ("sgd_maeclip", "mae_clip")])
def test_fit_steps_golden(dl, golden_dir, tag, name, loss):
    """Per-step loss and final weights after the golden steps, reference hyper-parameters (cnn.py:117)."""
    from dlas import ops
    g = np.load(os.path.join(golden_dir, "mlp_%s_steps.npz" % tag))
    dims, b, ns = g["dims"].tolist(), int(g["batch"]), int(g["n_steps"])
    opt = dl.Adam() if name.startswith("adam") else dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True)
    cfg = opt.cfg()
    p = _cuda(g["params0"])
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    gs = torch.zeros(p.numel() + 1, device="cuda")
    losses = torch.zeros(ns, device="cuda")
    ws = ops.mlp_workspace(dims, 0)
    ops.mlp_fit_steps(dims, dl.LOSS_MSE if loss == "mse" else dl.LOSS_MAE_CLIP, 0, cfg, p, m, v, _cuda(g["X"]),
                      _cuda(g["y"]), b, ns, 0, losses, gs, ws)
    torch.cuda.synchronize()
    np.testing.assert_allclose(losses.cpu().numpy(), g[name + "_losses"], rtol=TOL)
    assert _relerr(p.cpu().numpy(), g[name + "_final"]) < TOL


def test_facade_fit_200_steps_vs_oracle(dl):
    """Keras-shaped call sequence, 2 epochs x 100 batches with per-epoch batch shuffle, vs the oracle."""
    dims, batch, nb = [8, 64, 64, 1], 128, 100
    X, y = O.synth_rows(O.BASE_SEED, 0, batch * nb, 8)
    p0 = O.glorot_init(dims, O.BASE_SEED)
    model = dl.Sequential()
    model.add(dl.Dense(8, 64, activation="relu"))
    model.add(dl.Dense(64, 64, activation="relu"))
    model.add(dl.Dense(64, 1))
    model.compile(loss="mse", optimizer=dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    model.set_weights_flat(p0)
    hist = model.fit(X, y, nb_epoch=2, batch_size=batch, verbose=0)
    pr = p0.astype(np.float64)
    st, cfg = O.opt_state(pr.size), O.sgd_config()
    ep_loss = []
    for ep in range(2):
        order = O.batch_order(O.BASE_SEED, ep, nb)
        l = O.mlp_fit_steps(pr, dims, X.astype(np.float64), y.astype(np.float64), batch, nb, O.LOSS_MSE, cfg, st, order=order)
        ep_loss.append(l.mean())
    np.testing.assert_allclose(hist["loss"], ep_loss, rtol=TOL)
    assert _relerr(model.get_weights_flat(), pr) < TOL
    # predict / evaluate (cnn.py:132)
    yh = model.predict(X[:1000]).cpu().numpy()
    ref = O.mlp_forward(pr, dims, X[:1000].astype(np.float64))[-1]
    np.testing.assert_allclose(yh, ref, rtol=1e-3, atol=1e-5)
    ev = model.evaluate(X[:1000], y[:1000])
    assert ev == pytest.approx(float(((ref.ravel() - y[:1000]) ** 2).mean()), rel=TOL)


def test_facade_ragged_last_batch_callbacks_and_checkpoint(dl, tmp_path):
    dims = [8, 64, 64, 1]
    X, y = O.synth_rows(O.BASE_SEED, 0, 1050, 8)       # 8 full batches of 128 + 26 rows
    Xv, yv = O.synth_rows(O.BASE_SEED, 5000, 500, 8)
    model = dl.Sequential([dl.Dense(8, 64, activation="relu"), dl.Dense(64, 64, activation="relu"), dl.Dense(64, 1)])
    model.compile(loss=dl.mae_clip, optimizer=dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    p0 = model.get_weights_flat().copy()
    path = str(tmp_path) + "/models/static_ann.mdl"     # cnn.py:122 path shape
    hist = model.fit(X, y, validation_data=(Xv, yv), nb_epoch=3, batch_size=128, verbose=0, shuffle=False,
                     callbacks=[dl.EarlyStopping(monitor="val_loss", patience=10),
                                dl.ModelCheckpoint(path, save_best_only=True)])
    assert model.iterations == 27 and os.path.exists(path)
    pr = p0.astype(np.float64)
    st, cfg = O.opt_state(pr.size), O.sgd_config()
    X64, y64 = X.astype(np.float64), y.astype(np.float64)
    for ep in range(3):
        O.mlp_fit_steps(pr, dims, X64[:1024], y64[:1024], 128, 8, O.LOSS_MAE_CLIP, cfg, st)
        O.mlp_fit_steps(pr, dims, X64[1024:], y64[1024:], 26, 1, O.LOSS_MAE_CLIP, cfg, st)
    assert _relerr(model.get_weights_flat(), pr) < TOL
    lv, _ = O.mlp_loss_grad_sums(pr, dims, Xv.astype(np.float64), yv.astype(np.float64), O.LOSS_MAE_CLIP)
    assert hist["val_loss"][-1] == pytest.approx(lv / 500, rel=TOL)


def test_dynamic_ann_windows_train(dl):
    """Dynamic ANN (Readme.md:19): 24-lag windows of an 8-feature series -> 192-256-256-1."""
    from dlas import ops
    dims = [192, 256, 256, 1]
    s, ys = O.synth_rows(O.BASE_SEED, 0, 600, 8)
    W = ops.lag_windows(_cuda(s), 24)
    yw = ys[23:]
    p = O.glorot_init(dims, 5)
    ws = ops.mlp_workspace(dims, 0)
    gs = torch.zeros(p.size + 1, device="cuda")
    ops.mlp_grad(dims, 0, 0, _cuda(p), W, _cuda(yw), gs, ws)
    lr, gref = O.mlp_loss_grad_sums(p.astype(np.float64), dims, O.lag_windows(s, 24).astype(np.float64),
                                    yw.astype(np.float64), O.LOSS_MSE)
    g = gs.cpu().numpy()
    assert g[-1] == pytest.approx(lr, rel=TOL) and _relerr(g[:-1], gref) < TOL


def test_full_size_batch_properties(dl):
    """One 65,536-row batch of the 32-256-256-1 net (the bench shape): gradient of a batch equals
    the sum of the gradients of its two halves (linearity of the per-row sum), and the row-sharded
    sums reproduce what 2 data-parallel ranks would all-reduce."""
    from dlas import ops
    dims = [32, 256, 256, 1]
    X, y = ops.synth_rows(65536, 32)
    p = _cuda(O.glorot_init(dims, 9))
    ws = ops.mlp_workspace(dims, 0)
    full, a, b = (torch.zeros(p.numel() + 1, device="cuda") for _ in range(3))
    ops.mlp_grad(dims, 0, 0, p, X, y, full, ws)
    ops.mlp_grad(dims, 0, 0, p, X[:32768], y[:32768], a, ws)
    ops.mlp_grad(dims, 0, 0, p, X[32768:], y[32768:], b, ws)
    torch.cuda.synchronize()
    f, s = full.cpu().numpy().astype(np.float64), (a + b).cpu().numpy().astype(np.float64)
    assert _relerr(s, f) < 1e-5
    Xh, yh = O.synth_rows(O.BASE_SEED, 0, 4096, 32)     # oracle on the first 4096 rows
    ops.mlp_grad(dims, 0, 0, p, X[:4096], y[:4096], a, ws)
    lr, gref = O.mlp_loss_grad_sums(p.cpu().numpy().astype(np.float64), dims, Xh.astype(np.float64), yh.astype(np.float64), 0)
    g = a.cpu().numpy()
    assert g[-1] == pytest.approx(lr, rel=TOL) and _relerr(g[:-1], gref) < TOL


@pytest.mark.parametrize("mode,dims", [("fp32", [8, 64, 64, 1]), ("tc", [32, 256, 256, 1])])
def test_graph_replayed_steps_match_host_driven_steps(dl, mode, dims):
    """A step captured once in a CUDA graph (device-resident batch cursor and iteration counter) and
    replayed must give what the host-driven step loop gives -- same batches, same lr schedule."""
    batch, nb = 1024, 12
    X, y = O.synth_rows(O.BASE_SEED, 0, batch * nb, dims[0])
    p0 = O.glorot_init(dims, 3)

    def make():
        m = dl.Sequential(mode=mode)
        for a, b in zip(dims[:-2], dims[1:-1]):
            m.add(dl.Dense(a, b, activation="relu"))
        m.add(dl.Dense(dims[-2], dims[-1]))
        m.compile(loss="mse", optimizer=dl.SGD(lr=0.01, momentum=0.9, decay=1e-3, nesterov=True))   # visible decay
        m.set_weights_flat(p0)
        return m

    Xd, yd = _cuda(X), _cuda(y).reshape(-1, 1)
    a = make()
    st = a.graph_stepper(Xd, yd, batch, warmup_steps=3)     # steps 0..2 run eagerly, then the step is captured
    st.run(nb + 4 - 3)                                       # steps 3..15: the cursor wraps after 12 batches
    torch.cuda.synchronize()
    b = make()
    for s in range(nb + 4):
        k = s % nb
        b.train_on_batch(Xd[k * batch:(k + 1) * batch], yd[k * batch:(k + 1) * batch])
    torch.cuda.synchronize()
    assert a.iterations == b.iterations == nb + 4
    np.testing.assert_allclose(a.get_weights_flat(), b.get_weights_flat(), rtol=1e-5, atol=1e-7)
    pr = p0.astype(np.float64)
    cfg, stt = O.sgd_config(lr=0.01, momentum=0.9, decay=1e-3), O.opt_state(pr.size)
    ref = np.concatenate([O.mlp_fit_steps(pr, dims, X.astype(np.float64), y.astype(np.float64), batch, nb, O.LOSS_MSE, cfg, stt),
                          O.mlp_fit_steps(pr, dims, X.astype(np.float64), y.astype(np.float64), batch, 4, O.LOSS_MSE, cfg, stt)])
    np.testing.assert_allclose(st.recent_losses(8), ref[-8:], rtol=3e-4 if mode == "tc" else 1e-4)
    assert _relerr(a.get_weights_flat(), pr) < 1e-4


def test_lag_windows_read_in_place_match_the_materialised_windows(dl):
    """dlas_mlp_*_strided: 24-lag windows laid over the raw (n, 8) series (row stride 8 floats, nothing copied) give the
    SAME BITS as the materialised (n-23, 192) window matrix -- gradient sums, predictions, loss sum and three fit steps."""
    from dlas import ops
    dims = [192, 256, 256, 1]
    n = 3000
    s, ys = O.synth_rows(O.BASE_SEED, 0, n, 8)
    series = _cuda(s)
    yw = _cuda(ys[23:])
    W = ops.lag_windows(series, 24)                 # materialised: (n - 23, 192)
    V = dl.LagWindows(series, 24)                   # in place
    assert V.shape == tuple(W.shape)
    p = _cuda(O.glorot_init(dims, 5))
    ws = ops.mlp_workspace(dims, 0)
    ga, gb = (torch.zeros(p.numel() + 1, device="cuda") for _ in range(2))
    ops.mlp_grad(dims, 0, 0, p, W, yw, ga, ws)
    ops.mlp_grad(dims, 0, 0, p, V, yw, gb, ws)
    assert torch.equal(ga, gb)
    assert torch.equal(ops.mlp_predict(dims, 0, p, W), ops.mlp_predict(dims, 0, p, V))
    assert torch.equal(ops.mlp_eval_loss(dims, 1, 0, p, W, yw, ws), ops.mlp_eval_loss(dims, 1, 0, p, V, yw, ws))
    # a slice of the windows (rows 100..1124) and the batch loop (3 steps of 512 windows, shuffled order)
    ops.mlp_grad(dims, 0, 0, p, W[100:1124].contiguous(), yw[100:1124].contiguous(), ga, ws)
    ops.mlp_grad(dims, 0, 0, p, V[100:1124], yw[100:1124].contiguous(), gb, ws)
    assert torch.equal(ga, gb)
    cfg = dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True).cfg()
    res = []
    for Xin in (W, V):
        q = p.clone()
        m, v = torch.zeros_like(q), torch.zeros_like(q)
        ls = torch.zeros(3, device="cuda")
        ops.mlp_fit_steps(dims, 0, 0, cfg, q, m, v, Xin, yw, 512, 3, 0, ls, ga, ws, batch_order=[4, 0, 2])
        torch.cuda.synchronize()
        res.append((q, ls))
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])
    # through the facade
    model = dl.Sequential([dl.Dense(192, 256, activation="relu"), dl.Dense(256, 256, activation="relu"), dl.Dense(256, 1)])
    model.compile(loss="mse", optimizer=dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    model.set_weights_flat(p.cpu().numpy())
    h = model.fit(V, yw, nb_epoch=1, batch_size=512, verbose=0)
    assert np.isfinite(h["loss"][0])
    assert model.evaluate(V, yw) == pytest.approx(model.evaluate(W, yw), rel=0, abs=0)


def test_lag_windows_in_tensor_core_mode(dl):
    """4-lag windows over an 8-feature series are a 32-256-256-1 dynamic ANN: the tcgen05 kernel reads them in place too."""
    from dlas import ops
    dims = [32, 256, 256, 1]
    s, ys = O.synth_rows(O.BASE_SEED, 0, 5000, 8)
    series, yw = _cuda(s), _cuda(ys[3:])
    W, V = ops.lag_windows(series, 4), dl.LagWindows(series, 4)
    p = _cuda(O.glorot_init(dims, 5))
    ws = ops.mlp_workspace(dims, dl.MODE_TC)
    ga, gb = (torch.zeros(p.numel() + 1, device="cuda") for _ in range(2))
    ops.mlp_grad(dims, 0, dl.MODE_TC, p, W, yw, ga, ws)
    ops.mlp_grad(dims, 0, dl.MODE_TC, p, V, yw, gb, ws)
    torch.cuda.synchronize()
    assert torch.isfinite(ga).all() and torch.equal(ga, gb)
