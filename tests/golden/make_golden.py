"""Generates the committed golden vectors from the fp64 oracle (fixed seeds).

    python tests/golden/make_golden.py

The reference ships no fixtures (Readme.md:25), so these files pin OUR
restatement: tests/test_oracle.py re-runs the oracle against them, the -m gpu tests compare the
CUDA kernels with them."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import dlas_oracle as O  # noqa: E402

SEED = O.BASE_SEED


def gilbert_kat():
    hand_P = np.array([435.0, 870.0, 435.0, 435.0, 1.0, 1000.0, 50.0, 5000.0, 0.0], dtype=np.float32)
    hand_S = np.array([1.0, 1.0, 2.0, 64.0, 1.0, 32.0, 4.0, 128.0, 16.0], dtype=np.float32)
    hand_R = np.array([1.0, 1.0, 1.0, 1.0, 1.0, 0.5, 0.02, 10.0, 1.0], dtype=np.float32)
    P, S, R = O.synth_gilbert(SEED, 0, 4096)
    rng = np.random.default_rng(7)
    wP = np.exp(rng.uniform(-20, 20, 1024)).astype(np.float32)
    wS = np.exp(rng.uniform(-8, 8, 1024)).astype(np.float32)
    wR = np.exp(rng.uniform(-8, 8, 1024)).astype(np.float32)
    P = np.concatenate([hand_P, P, wP]); S = np.concatenate([hand_S, S, wS]); R = np.concatenate([hand_R, R, wR])
    q = O.gilbert(P, S, R, np.float64)
    np.savez_compressed(os.path.join(HERE, "gilbert_kat.npz"), P=P, S=S, R=R, q64=q, n_hand=len(hand_P))


def mlp_steps(dims, batch, n_steps, tag):
    n = batch * n_steps
    X, y = O.synth_rows(SEED, 0, n, dims[0])
    p0 = O.glorot_init(dims, SEED)
    out = dict(dims=np.array(dims), batch=batch, n_steps=n_steps, X=X, y=y, params0=p0)
    for name, cfg, loss in (("sgd_mse", O.sgd_config(), O.LOSS_MSE), ("adam_mse", O.adam_config(), O.LOSS_MSE),
                            ("sgd_maeclip", O.sgd_config(), O.LOSS_MAE_CLIP)):
        p = p0.astype(np.float64)
        st = O.opt_state(p.size)
        losses = O.mlp_fit_steps(p, dims, X.astype(np.float64), y.astype(np.float64), batch, n_steps, loss, cfg, st)
        out[name + "_losses"] = losses
        out[name + "_final"] = p
    ls, g = O.mlp_loss_grad_sums(p0.astype(np.float64), dims, X[:batch].astype(np.float64), y[:batch].astype(np.float64), O.LOSS_MSE)
    out["grad0_loss_sum"] = ls
    out["grad0"] = g
    np.savez_compressed(os.path.join(HERE, "mlp_%s_steps.npz" % tag), **out)


def dynamic_ann_steps(lags, batch, n_steps):
    """configs[3] shape: an (n + lags - 1, 8) raw series, sample i = rows i .. i + lags - 1 flattened lag-major (Readme.md:17-19),
    192-256-256-1.  The series is stored, not the windows: consumers build them (oracle) or read them in place (kernels)."""
    n = batch * n_steps
    series, ys = O.synth_rows(SEED, 5000, n + lags - 1, 8)
    y = ys[lags - 1:]
    dims = [lags * 8, 256, 256, 1]
    W = O.lag_windows(series.astype(np.float64), lags)
    p0 = O.glorot_init(dims, SEED + 3)
    out = dict(dims=np.array(dims), lags=lags, batch=batch, n_steps=n_steps, series=series, y=y, params0=p0)
    for name, cfg in (("sgd_mse", O.sgd_config()), ("adam_mse", O.adam_config())):
        p = p0.astype(np.float64)
        out[name + "_losses"] = O.mlp_fit_steps(p, dims, W, y.astype(np.float64), batch, n_steps, O.LOSS_MSE, cfg, O.opt_state(p.size))
        out[name + "_final"] = p.astype(np.float32)          # 115,457 parameters: stored as fp32 (tolerances are 1e-4)
    ls, g = O.mlp_loss_grad_sums(p0.astype(np.float64), dims, W[:batch], y[:batch].astype(np.float64), O.LOSS_MSE)
    out["grad0_loss_sum"] = ls
    out["grad0"] = g.astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "dynamic_ann_192_256_256_1_steps.npz"), **out)


def lstm_vectors():
    out = {}
    for tag, (d, h, nl, T, B) in (("small", (3, 8, 2, 5, 6)), ("h128", (8, 128, 2, 48, 4))):
        X, y = O.synth_series(SEED, 0, B, T, d)
        p = O.lstm_init(d, h, nl, SEED)
        ls, g = O.lstm_loss_grad_sums(p.astype(np.float64), d, h, nl, X.astype(np.float64), y.astype(np.float64), O.LOSS_MSE)
        yhat = O.lstm_forward(p.astype(np.float64), d, h, nl, X.astype(np.float64))
        out.update({tag + "_shape": np.array([d, h, nl, T, B]), tag + "_X": X, tag + "_y": y, tag + "_params": p,
                    tag + "_loss_sum": ls, tag + "_grad": g.astype(np.float32 if tag == "h128" else np.float64),
                    tag + "_yhat": yhat})
    np.savez_compressed(os.path.join(HERE, "lstm_vectors.npz"), **out)


def dp_equivalence():
    dims = [8, 64, 64, 1]
    batch, n_steps = 64, 6
    X, y = O.synth_rows(SEED, 1000, batch * n_steps, dims[0])
    p0 = O.glorot_init(dims, SEED + 1)
    out = dict(dims=np.array(dims), batch=batch, n_steps=n_steps, X=X, y=y, params0=p0)
    for w in (1, 2, 4, 8):
        p = p0.astype(np.float64)
        st = O.opt_state(p.size)
        out["losses_w%d" % w] = O.mlp_fit_steps(p, dims, X.astype(np.float64), y.astype(np.float64), batch, n_steps,
                                                 O.LOSS_MSE, O.sgd_config(), st, world=w)
        out["final_w%d" % w] = p
    np.savez_compressed(os.path.join(HERE, "dp_equivalence.npz"), **out)


if __name__ == "__main__":
    gilbert_kat()
    mlp_steps([8, 64, 64, 1], 32, 8, "8_64_64_1")
    mlp_steps([32, 256, 256, 1], 64, 4, "32_256_256_1")
    dynamic_ann_steps(24, 128, 4)
    lstm_vectors()
    dp_equivalence()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
