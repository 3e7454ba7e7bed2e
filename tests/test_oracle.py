"""CPU tests of the oracle itself (the reference has no tests to borrow):
golden vectors, hand-checkable values, and independent cross-checks (torch.autograd, torch.nn.LSTM,
torch.optim.SGD) that pin our restatement of the un-vendored Keras semantics."""
import os

import numpy as np
import pytest
import torch

import dlas_oracle as O


def _g(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


# ---- Gilbert ----------------------------------------------------------------------------------
def test_gilbert_hand_values():
    q = O.gilbert(np.float32([435, 870, 435]), np.float32([1, 1, 2]), np.float32([1, 1, 1]))
    assert q[0] == pytest.approx(1.0, rel=1e-15)
    assert q[1] == pytest.approx(2.0, rel=1e-15)
    assert q[2] == pytest.approx(2.0 ** 1.89, rel=1e-15)
    # textbook form P_wh = 435 R^0.546 q / S^1.89 inverts to the same q
    P, S, R = 1000.0, 32.0, 0.5
    qq = O.gilbert(np.float32([P]), np.float32([S]), np.float32([R]))[0]
    assert 435.0 * R ** 0.546 * qq / S ** 1.89 == pytest.approx(P, rel=1e-14)


def test_gilbert_golden(golden_dir):
    g = _g(golden_dir, "gilbert_kat.npz")
    q = O.gilbert(g["P"], g["S"], g["R"])
    np.testing.assert_allclose(q, g["q64"], rtol=1e-15)
    # fp32 evaluation of the plain formula stays inside the 1e-6 budget on the bench range
    sl = slice(int(g["n_hand"]), int(g["n_hand"]) + 4096)
    q32 = O.gilbert(g["P"][sl], g["S"][sl], g["R"][sl], np.float32)
    assert np.abs(q32 / g["q64"][sl] - 1).max() < 1e-6


def test_gilbert_kernel_algorithm_emulated(golden_dir):
    """The fp32 instruction sequence of csrc/gilbert.cu, emulated in NumPy, meets 1e-6."""
    import gilbert_poly_fit as F
    g = _g(golden_dir, "gilbert_kat.npz")
    QC, _ = F.fit(F.q_fun, F.lo, F.hi, 7)
    EC, _ = F.fit(lambda r: np.exp2(r), -0.5, 0.5, 6)
    ok = g["P"] > 0
    q = F.gilbert32(g["P"][ok], g["S"][ok], g["R"][ok], QC.astype(np.float32), EC.astype(np.float32))
    assert np.abs(q / g["q64"][ok] - 1).max() < 5e-7


# ---- synthetic generator --------------------------------------------------------------------
def test_splitmix_known_answers():
    # splitmix64 reference outputs for the stream seeded with 0 (published test vector)
    s = np.uint64(0)
    outs = []
    for _ in range(3):
        outs.append(int(O.splitmix64(s)))
        with np.errstate(over="ignore"):
            s = s + np.uint64(0x9E3779B97F4A7C15)
    assert outs == [0xE220A8397B1DCDAF, 0x6E789E6AA1B965F4, 0x06C45D188009454F]


def test_synth_ranges_and_determinism():
    P, S, R = O.synth_gilbert(O.BASE_SEED, 12345, 20000)
    assert P.dtype == S.dtype == R.dtype == np.float32
    assert P.min() >= 50 and P.max() <= 5000 and S.min() >= 4 and S.max() <= 128
    assert R.min() >= np.float32(0.02) and R.max() < 10.25
    P2, _, _ = O.synth_gilbert(O.BASE_SEED, 12345 + 100, 50)
    np.testing.assert_array_equal(P[100:150], P2)  # any slice regenerates
    X, y = O.synth_rows(O.BASE_SEED, 7, 1000, 32)
    assert X.min() >= -1 and X.max() < 1 and np.abs(y).max() <= 1.05
    X2, y2 = O.synth_rows(O.BASE_SEED, 507, 10, 32)
    np.testing.assert_array_equal(X[500:510], X2)
    np.testing.assert_array_equal(y[500:510], y2)


def test_split_ratios():
    c = O.split_code(O.BASE_SEED, np.arange(200000))
    fr = [np.mean(c == k) for k in range(3)]
    assert abs(fr[0] - 0.64) < 0.005 and abs(fr[1] - 0.16) < 0.005 and abs(fr[2] - 0.20) < 0.005


def test_batch_order_is_permutation():
    o = O.batch_order(O.BASE_SEED, 3, 1000)
    assert sorted(o.tolist()) == list(range(1000))
    assert not np.array_equal(o, O.batch_order(O.BASE_SEED, 4, 1000))


# ---- losses ---------------------------------------------------------------------------------
def test_mae_clip_matches_cnn_py_definition():
    rng = np.random.default_rng(0)
    y, yh = rng.normal(size=(50, 12)) * 5, rng.normal(size=(50, 12)) * 5
    per, g = O.loss_and_grad(yh, y, O.LOSS_MAE_CLIP)
    np.testing.assert_allclose(per, np.clip(np.abs(y - yh), 0, 6).mean(axis=-1))  # cnn.py:29-32
    yt = torch.tensor(yh, requires_grad=True)
    torch.clamp((torch.tensor(y) - yt).abs(), 0, 6).mean(dim=-1).sum().backward()
    np.testing.assert_allclose(g, yt.grad.numpy(), atol=1e-15)


# ---- MLP backward vs autograd ------------------------------------------------------------------
@pytest.mark.parametrize("dims", [[8, 64, 64, 1], [5, 7, 3], [32, 256, 256, 1]])
@pytest.mark.parametrize("loss", [O.LOSS_MSE, O.LOSS_MAE_CLIP])
def test_mlp_backward_vs_autograd(dims, loss):
    rng = np.random.default_rng(1)
    n = O.mlp_n_params(dims)
    p = rng.normal(size=n) * 0.3
    X = rng.normal(size=(17, dims[0]))
    y = rng.normal(size=(17, dims[-1])) * 4
    ls, g = O.mlp_loss_grad_sums(p, dims, X, y, loss)
    pt = torch.tensor(p, requires_grad=True)
    h = torch.tensor(X)
    offs, _ = O.mlp_offsets(dims)
    for li, (w, b) in enumerate(offs):
        W = pt[w:w + dims[li] * dims[li + 1]].reshape(dims[li], dims[li + 1])
        h = h @ W + pt[b:b + dims[li + 1]]
        if li < len(offs) - 1:
            h = torch.relu(h)
    d = h - torch.tensor(y)
    per = (d * d).mean(dim=1) if loss == O.LOSS_MSE else torch.clamp(d.abs(), 0, 6).mean(dim=1)
    per.sum().backward()
    assert ls == pytest.approx(per.sum().item(), rel=1e-12)
    np.testing.assert_allclose(g, pt.grad.numpy(), rtol=1e-9, atol=1e-12)


def test_mlp_golden(golden_dir):
    for tag in ("8_64_64_1", "32_256_256_1"):
        g = _g(golden_dir, "mlp_%s_steps.npz" % tag)
        dims = g["dims"].tolist()
        for name, cfg, loss in (("sgd_mse", O.sgd_config(), O.LOSS_MSE), ("adam_mse", O.adam_config(), O.LOSS_MSE),
                                ("sgd_maeclip", O.sgd_config(), O.LOSS_MAE_CLIP)):
            p = g["params0"].astype(np.float64)
            st = O.opt_state(p.size)
            losses = O.mlp_fit_steps(p, dims, g["X"].astype(np.float64), g["y"].astype(np.float64), int(g["batch"]),
                                     int(g["n_steps"]), loss, cfg, st)
            np.testing.assert_allclose(losses, g[name + "_losses"], rtol=1e-12)
            np.testing.assert_allclose(p, g[name + "_final"], rtol=1e-10, atol=1e-14)


# ---- optimisers -------------------------------------------------------------------------------
def test_sgd_nesterov_matches_torch_without_decay():
    """Keras' v = m v - lr g; p += m v - lr g  equals torch's Nesterov SGD when lr is constant."""
    rng = np.random.default_rng(2)
    p = rng.normal(size=20)
    pt = torch.tensor(p.copy(), requires_grad=True)
    opt = torch.optim.SGD([pt], lr=0.001, momentum=0.99, nesterov=True)
    cfg, st = O.sgd_config(decay=0.0), O.opt_state(20)
    for _ in range(25):
        g = rng.normal(size=20)
        pt.grad = torch.tensor(g)
        opt.step()
        O.opt_step(p, g, cfg, st)
    np.testing.assert_allclose(p, pt.detach().numpy(), rtol=1e-10, atol=1e-13)


def test_sgd_decay_schedule():
    cfg, st = O.sgd_config(lr=0.1, momentum=0.0, decay=0.5, nesterov=False), O.opt_state(1)
    p = np.zeros(1)
    for it in range(4):  # lr_t = 0.1 / (1 + 0.5 it), applied with the pre-increment counter
        before = p.copy()
        O.opt_step(p, np.ones(1), cfg, st)
        assert before[0] - p[0] == pytest.approx(0.1 / (1 + 0.5 * it), rel=1e-12)


def test_adam_first_step_and_eps_placement():
    cfg, st = O.adam_config(lr=0.001, eps=1e-7), O.opt_state(3)
    p = np.zeros(3)
    g = np.array([1.0, -2.0, 1e-9])
    O.opt_step(p, g, cfg, st)
    m, v = 0.1 * g, 0.001 * g * g
    lr_t = 0.001 * np.sqrt(1 - 0.999) / (1 - 0.9)
    np.testing.assert_allclose(p, -lr_t * m / (np.sqrt(v) + 1e-7), rtol=1e-12)


# ---- data-parallel averaging -------------------------------------------------------------------
def test_dp_replicas_equal_global_batch(golden_dir):
    g = _g(golden_dir, "dp_equivalence.npz")
    for w in (2, 4, 8):
        np.testing.assert_allclose(g["losses_w%d" % w], g["losses_w1"], rtol=1e-12)
        np.testing.assert_allclose(g["final_w%d" % w], g["final_w1"], rtol=1e-9, atol=1e-13)
    dims = g["dims"].tolist()
    p = g["params0"].astype(np.float64)
    st = O.opt_state(p.size)
    losses = O.mlp_fit_steps(p, dims, g["X"].astype(np.float64), g["y"].astype(np.float64), int(g["batch"]),
                             int(g["n_steps"]), O.LOSS_MSE, O.sgd_config(), st, world=4)
    np.testing.assert_allclose(losses, g["losses_w4"], rtol=1e-12)


# ---- LSTM ----------------------------------------------------------------------------------
def _torch_lstm(p, d, h, nl, X):
    offs, (hw, hb), _ = O.lstm_offsets(d, h, nl)
    lstm = torch.nn.LSTM(d, h, nl, batch_first=True).double()
    with torch.no_grad():
        for li, (k, b, dl) in enumerate(offs):
            K = torch.tensor(p[k:k + (dl + h) * 4 * h].reshape(dl + h, 4 * h))
            getattr(lstm, "weight_ih_l%d" % li).copy_(K[:dl].T)   # torch gate order i,f,g,o = Keras i,f,c,o
            getattr(lstm, "weight_hh_l%d" % li).copy_(K[dl:].T)
            getattr(lstm, "bias_ih_l%d" % li).copy_(torch.tensor(p[b:b + 4 * h]))
            getattr(lstm, "bias_hh_l%d" % li).zero_()
    out, _ = lstm(torch.tensor(X))
    return out[:, -1, :] @ torch.tensor(p[hw:hw + h].reshape(h, 1)) + torch.tensor(p[hb:hb + 1]), lstm


def test_lstm_forward_vs_torch_and_bptt_vs_autograd():
    d, h, nl, T, B = 3, 8, 2, 5, 6
    rng = np.random.default_rng(3)
    _, _, n = O.lstm_offsets(d, h, nl)
    p = rng.normal(size=n) * 0.4
    X = rng.normal(size=(B, T, d))
    y = rng.normal(size=(B, 1))
    yh = O.lstm_forward(p, d, h, nl, X)
    yt, lstm = _torch_lstm(p, d, h, nl, X)
    np.testing.assert_allclose(yh, yt.detach().numpy(), rtol=1e-10, atol=1e-12)
    ls, g = O.lstm_loss_grad_sums(p, d, h, nl, X, y, O.LOSS_MSE)
    ((yt - torch.tensor(y)) ** 2).sum().backward()
    offs, (hw, hb), _ = O.lstm_offsets(d, h, nl)
    for li, (k, b, dl) in enumerate(offs):
        K = g[k:k + (dl + h) * 4 * h].reshape(dl + h, 4 * h)
        np.testing.assert_allclose(K[:dl].T, getattr(lstm, "weight_ih_l%d" % li).grad.numpy(), rtol=1e-8, atol=1e-11)
        np.testing.assert_allclose(K[dl:].T, getattr(lstm, "weight_hh_l%d" % li).grad.numpy(), rtol=1e-8, atol=1e-11)
        np.testing.assert_allclose(g[b:b + 4 * h], getattr(lstm, "bias_ih_l%d" % li).grad.numpy(), rtol=1e-8, atol=1e-11)


def test_lstm_golden(golden_dir):
    g = _g(golden_dir, "lstm_vectors.npz")
    for tag in ("small", "h128"):
        d, h, nl, T, B = g[tag + "_shape"].tolist()
        ls, gr = O.lstm_loss_grad_sums(g[tag + "_params"].astype(np.float64), d, h, nl, g[tag + "_X"].astype(np.float64),
                                       g[tag + "_y"].astype(np.float64), O.LOSS_MSE)
        assert ls == pytest.approx(float(g[tag + "_loss_sum"]), rel=1e-12)
        np.testing.assert_allclose(gr, g[tag + "_grad"], rtol=2e-6 if tag == "h128" else 1e-10, atol=1e-9)


# ---- epoch-level host logic (cnn.py:121-123) ------------------------------------------------------
def test_early_stopping_trace():
    v = [5, 4, 3] + [3.5] * 12
    stop, best = O.early_stopping_trace(v, patience=10)
    assert (stop, best) == (12, 2)
    assert O.early_stopping_trace([3, 2, 1], patience=10) == (None, 2)


def test_lag_windows():
    s = np.arange(20, dtype=np.float64).reshape(10, 2)
    w = O.lag_windows(s, 3)
    assert w.shape == (8, 6)
    np.testing.assert_array_equal(w[0], [0, 1, 2, 3, 4, 5])
    np.testing.assert_array_equal(w[-1], [14, 15, 16, 17, 18, 19])


def test_torch_cpu_port_matches_oracle(golden_dir):
    """The timed CPU baseline (oracle/torch_cpu.py) computes what the fp64 oracle computes."""
    import torch_cpu as T
    g = _g(golden_dir, "mlp_8_64_64_1_steps.npz")
    dims, b, ns = g["dims"].tolist(), int(g["batch"]), int(g["n_steps"])
    for name, loss in (("sgd_mse", "mse"), ("sgd_maeclip", "mae_clip")):
        m = T.TorchCpuMlp(dims, g["params0"], loss=loss)
        X, y = torch.from_numpy(g["X"]), torch.from_numpy(g["y"]).reshape(-1, 1)
        losses = [m.step(X[s * b:(s + 1) * b], y[s * b:(s + 1) * b]) for s in range(ns)]
        np.testing.assert_allclose(losses, g[name + "_losses"], rtol=1e-4)
        ref = g[name + "_final"]
        assert np.linalg.norm(m.flat() - ref) / np.linalg.norm(ref) < 1e-5
    P, S, R = O.synth_gilbert(O.BASE_SEED, 0, 10000)
    q = T.gilbert_numpy_f32(P, S, R, threads=3)
    assert np.abs(q / O.gilbert(P, S, R) - 1).max() < 1e-6


def test_dynamic_ann_golden_reproduces(golden_dir):
    """configs[3] fixture (raw series + lag windows, 192-256-256-1): the oracle reproduces its committed losses / weights, and
    the torch-CPU port (the timed CPU baseline) lands on them within the fp32 tolerance."""
    import torch_cpu as T
    g = _g(golden_dir, "dynamic_ann_192_256_256_1_steps.npz")
    dims, lags, b, ns = g["dims"].tolist(), int(g["lags"]), int(g["batch"]), int(g["n_steps"])
    W = O.lag_windows(g["series"].astype(np.float64), lags)
    assert W.shape == (b * ns, dims[0]) and dims[0] == lags * 8
    p = g["params0"].astype(np.float64)
    losses = O.mlp_fit_steps(p, dims, W, g["y"].astype(np.float64), b, ns, O.LOSS_MSE, O.sgd_config(), O.opt_state(p.size))
    np.testing.assert_allclose(losses, g["sgd_mse_losses"], rtol=1e-12)
    assert np.linalg.norm(p - g["sgd_mse_final"]) / np.linalg.norm(p) < 1e-6          # stored as fp32
    m = T.TorchCpuMlp(dims, g["params0"], loss="mse")
    X, y = torch.from_numpy(W.astype(np.float32)), torch.from_numpy(g["y"]).reshape(-1, 1)
    tl = [m.step(X[s * b:(s + 1) * b], y[s * b:(s + 1) * b]) for s in range(ns)]
    np.testing.assert_allclose(tl, g["sgd_mse_losses"], rtol=1e-4)
    assert np.linalg.norm(m.flat() - g["sgd_mse_final"]) / np.linalg.norm(g["sgd_mse_final"]) < 1e-5


def test_parameter_averaging_reduces_to_gradient_averaging_for_plain_sgd():
    """every=1, no momentum: averaging parameters after one local step == one step on the mean gradient."""
    dims, batch, steps = [5, 7, 1], 32, 5
    rng = np.random.default_rng(4)
    p0 = rng.normal(size=O.mlp_n_params(dims)) * 0.3
    X, y = rng.normal(size=(batch * steps, 5)), rng.normal(size=(batch * steps, 1))
    cfg = O.sgd_config(lr=0.05, momentum=0.0, decay=0.0, nesterov=False)
    a = p0.copy()
    la = O.mlp_fit_steps_param_avg(a, dims, X, y, batch, steps, O.LOSS_MSE, cfg, world=4, every=1)
    b = p0.copy()
    lb = O.mlp_fit_steps(b, dims, X, y, batch, steps, O.LOSS_MSE, cfg, O.opt_state(b.size))
    np.testing.assert_allclose(a, b, rtol=1e-10, atol=1e-13)
    np.testing.assert_allclose(la, lb, rtol=1e-10)
    c = p0.copy()
    O.mlp_fit_steps_param_avg(c, dims, X, y, batch, steps, O.LOSS_MSE, cfg, world=4, every=3)   # local SGD drifts
    assert np.abs(c - b).max() > 1e-6


def test_cnn_backward_vs_autograd_and_shapes():
    """cnn.py:110-114: Conv1D(1->100, k=13, relu) -> Dropout(0.5) -> Flatten -> Dense(3600 -> 12) on (48, 1)."""
    L, K, F, n_out, B = 48, 13, 100, 12, 5
    (wc, bc, wd, bd), n = O.cnn_offsets(L, K, F, n_out)
    assert n == 44612 and (L - K + 1) * F == 3600          # 1,400 + 43,212 parameters, Dense(3600, 12)
    rng = np.random.default_rng(8)
    p = rng.normal(size=n) * 0.1
    X, y = rng.normal(size=(B, L)), rng.normal(size=(B, n_out)) * 4
    mask = O.dropout_mask(O.BASE_SEED, 77, B, 3600)
    assert 0.45 < mask.mean() < 0.55
    for loss in (O.LOSS_MSE, O.LOSS_MAE_CLIP):
        ls, g = O.cnn_loss_grad_sums(p, L, K, F, n_out, X, y, loss, mask)
        pt = torch.tensor(p, requires_grad=True)
        conv = torch.nn.functional.conv1d(torch.tensor(X)[:, None, :], pt[wc:wc + K * F].reshape(K, F).T[:, None, :], pt[bc:bc + F])
        a = torch.relu(conv).permute(0, 2, 1).reshape(B, 3600) * torch.tensor(mask) / 0.5      # (B, T, F) time-major flatten
        out = a @ pt[wd:wd + 3600 * n_out].reshape(3600, n_out) + pt[bd:bd + n_out]
        d = out - torch.tensor(y)
        per = (d * d).mean(dim=1) if loss == O.LOSS_MSE else torch.clamp(d.abs(), 0, 6).mean(dim=1)
        per.sum().backward()
        assert ls == pytest.approx(per.sum().item(), rel=1e-12)
        np.testing.assert_allclose(g, pt.grad.numpy(), rtol=1e-9, atol=1e-12)
