"""GPU parity of the LSTM path (forward, full BPTT, step) through the C ABI against the fp64 oracle
and the committed golden vectors.  Tolerance (BASELINE.json): 1e-4 relative, fp32 path."""
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


def _grad(d, h, nl, T, loss, p, X, y):
    from dlas import ops
    ws = ops.lstm_workspace(d, h, nl, T, X.shape[0])
    gs = torch.zeros(ops.lstm_n_params(d, h, nl) + 1, device="cuda")
    ops.lstm_grad(d, h, nl, T, loss, _cuda(p), _cuda(X), _cuda(y), gs, ws)
    torch.cuda.synchronize()
    g = gs.cpu().numpy()
    return g[-1], g[:-1]


@pytest.mark.parametrize("tag", ["small", "h128"])
def test_grad_golden(dl, golden_dir, tag):
    g = np.load(os.path.join(golden_dir, "lstm_vectors.npz"))
    d, h, nl, T, B = g[tag + "_shape"].tolist()
    from dlas import ops
    assert ops.lstm_n_params(d, h, nl) == g[tag + "_params"].size
    ls, gr = _grad(d, h, nl, T, 0, g[tag + "_params"], g[tag + "_X"], g[tag + "_y"])
    assert ls == pytest.approx(float(g[tag + "_loss_sum"]), rel=TOL)
    assert _relerr(gr, g[tag + "_grad"].astype(np.float64)) < TOL
    ws = ops.lstm_workspace(d, h, nl, T, B)
    yh = ops.lstm_predict(d, h, nl, T, _cuda(g[tag + "_params"]), _cuda(g[tag + "_X"]), ws).cpu().numpy()
    np.testing.assert_allclose(yh, g[tag + "_yhat"], rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("shape", [(8, 128, 2, 48, 37), (3, 16, 1, 7, 9), (8, 64, 3, 12, 20), (5, 128, 2, 3, 1)])
@pytest.mark.parametrize("loss", [O.LOSS_MSE, O.LOSS_MAE_CLIP])
def test_grad_vs_oracle_ragged(dl, shape, loss):
    d, h, nl, T, B = shape
    rng = np.random.default_rng(B)
    _, _, n = O.lstm_offsets(d, h, nl)
    p = (rng.normal(size=n) * 0.15).astype(np.float32)
    X = rng.normal(size=(B, T, d)).astype(np.float32)
    y = (rng.normal(size=(B, 1)) * 3).astype(np.float32)
    ls, gr = _grad(d, h, nl, T, loss, p, X, y)
    lr, gref = O.lstm_loss_grad_sums(p.astype(np.float64), d, h, nl, X.astype(np.float64), y.astype(np.float64), loss)
    assert ls == pytest.approx(lr, rel=TOL)
    assert _relerr(gr, gref) < TOL


def test_keras_shaped_fit_vs_oracle(dl):
    """Sequential([LSTM, LSTM, Dense(1)]) for 6 steps of Adam and of the reference SGD (cnn.py:117)."""
    d, h, nl, T, B, steps = 8, 128, 2, 48, 64, 6
    X, y = O.synth_series(O.BASE_SEED, 0, B * steps, T, d)
    p0 = O.lstm_init(d, h, nl, O.BASE_SEED)
    for opt, cfg in ((dl.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True), O.sgd_config()), (dl.Adam(), O.adam_config())):
        model = dl.Sequential()
        model.add(dl.LSTM(h, input_shape=(T, d), return_sequences=True))
        model.add(dl.LSTM(h))
        model.add(dl.Dense(1))
        model.compile(loss="mse", optimizer=opt)
        np.testing.assert_allclose(model.get_weights_flat(), p0, rtol=1e-6, atol=1e-7)   # same init as the oracle
        model.set_weights_flat(p0)
        hist = model.fit(X, y, nb_epoch=1, batch_size=B, shuffle=False, verbose=0)
        pr = p0.astype(np.float64)
        losses = O.lstm_fit_steps(pr, d, h, nl, X.astype(np.float64), y.astype(np.float64), B, steps, O.LOSS_MSE, cfg,
                                  O.opt_state(pr.size))
        assert hist["loss"][0] == pytest.approx(losses.mean(), rel=TOL)
        assert _relerr(model.get_weights_flat(), pr) < TOL
        ev = model.evaluate(X[:100], y[:100])
        ref = O.lstm_forward(pr, d, h, nl, X[:100].astype(np.float64))
        assert ev == pytest.approx(float(((ref.ravel() - y[:100]) ** 2).mean()), rel=1e-3)


def test_batch_4096_shard_linearity(dl):
    """Config 5 batch (4096 sequences): the gradient sum over the batch equals the sum over two
    2048-sequence shards -- what two data-parallel ranks all-reduce."""
    from dlas import ops
    d, h, nl, T = 8, 128, 2, 48
    X, y = ops.synth_series(4096, T, d)
    p = _cuda(O.lstm_init(d, h, nl, 4))
    ws = ops.lstm_workspace(d, h, nl, T, 4096)
    full, a, b = (torch.zeros(p.numel() + 1, device="cuda") for _ in range(3))
    ops.lstm_grad(d, h, nl, T, 0, p, X, y, full, ws)
    ops.lstm_grad(d, h, nl, T, 0, p, X[:2048], y[:2048], a, ws)
    ops.lstm_grad(d, h, nl, T, 0, p, X[2048:], y[2048:], b, ws)
    torch.cuda.synchronize()
    assert _relerr((a + b).cpu().numpy(), full.cpu().numpy().astype(np.float64)) < 1e-5
