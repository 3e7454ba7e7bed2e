"""CPU oracle for the dlas-b200 hot path.  TEST INFRASTRUCTURE -- NOT A PRODUCT PATH.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product (the ``dlas`` package and
``libdlas_b200.so``) never routes through it and has no CPU fallback.

PARITY UNPINNED.  The reference snapshot (/root/reference) holds one un-runnable Keras-0.x /
Theano / PySpark sketch (``cnn.py``) and a README of headings; it ships no tests, fixtures,
golden vectors or numbers, and Keras/Theano/Spark are neither vendored nor installable here
(SURVEY.md section 0, section 8c).  This file is therefore a *restatement* of the declared
intent.  What it follows, line by line where a line exists:

* ``mae_clip``              -> cnn.py:29-32  (CLIP_VALUE = 6, mean over the last axis)
* SGD hyper-parameters      -> cnn.py:117    (lr=0.001, momentum=0.99, decay=1e-6, nesterov)
* SGD / Adam update rules   -> Keras 1.x ``optimizers.py`` (un-vendored; restated from the
                               published algorithm: lr_t = lr/(1+decay*it), v = m*v - lr_t*g,
                               Nesterov p += m*v - lr_t*g; Adam with epsilon outside the
                               bias-corrected sqrt)
* batch loop / shuffle      -> cnn.py:127-128 (``fit`` with batch_size, nb_epoch)
* 64/16/20 split            -> cnn.py:68
* EarlyStopping/Checkpoint  -> cnn.py:121-123
* Gilbert's equation        -> BASELINE.json north_star (q = P*S^1.89/(435*R^0.546));
                               Readme.md:7-8 names it, no code exists
* static/dynamic ANN, LSTM  -> BASELINE.json configs; Readme.md:15-21 headings only
* LSTM cell                 -> Keras ``recurrent.LSTM`` gate order i,f,c,o (un-vendored)

Self-made pins (tests/test_oracle.py): manual backward vs ``torch.autograd``; LSTM forward vs
``torch.nn.LSTM``; SGD vs ``torch.optim.SGD`` where the two rules coincide; committed golden
vectors under tests/golden/ produced by ``tests/golden/make_golden.py``.

Everything here is NumPy.  ``dtype=np.float64`` is the truth; ``dtype=np.float32`` gives the
expected-rounding companion.
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------------------
# Counter-based synthetic generator (SURVEY.md section 8d [PIN]); mirrored bit-for-bit in
# csrc/synth.cu so any slice of a device-generated shard can be regenerated here.
# --------------------------------------------------------------------------------------
BASE_SEED = 20261106
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(x):
    """splitmix64 finaliser on uint64 arrays (wrap-around arithmetic)."""
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = x + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def stream_key(seed: int, stream: int) -> np.uint64:
    with np.errstate(over="ignore"):
        s = np.uint64(seed & 0xFFFFFFFFFFFFFFFF)
        t = np.uint64(stream & 0xFFFFFFFFFFFFFFFF) * np.uint64(0xD1342543DE82EF95)
        return splitmix64(s ^ splitmix64(t))


def hash_u24(seed: int, stream: int, idx):
    """24-bit integers n with u = n * 2^-24 in [0,1); idx is a uint64 counter array."""
    key = stream_key(seed, stream)
    with np.errstate(over="ignore"):
        h = splitmix64(key + np.asarray(idx, dtype=np.uint64))
    return (h >> np.uint64(40)).astype(np.uint32)


def hash_uniform(seed: int, stream: int, idx):
    return hash_u24(seed, stream, idx).astype(np.float32) * np.float32(2.0 ** -24)


def fma32(a, b, c):
    """Emulated single-rounded fp32 fma (product exact in fp64; the fp64 add rounds once more,
    a double rounding that differs from a true fma with probability ~2^-29 per call)."""
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(np.float32)


# Stream ids of the synthetic workloads
STREAM_GILBERT_P, STREAM_GILBERT_S, STREAM_GILBERT_R = 1, 2, 3
STREAM_X, STREAM_NOISE, STREAM_TEACHER = 10, 11, 12
STREAM_SPLIT, STREAM_PERM = 20, 21


def synth_gilbert(seed: int, start: int, n: int):
    """Config 2 inputs.  P = 50 + 4950u psig, S = 4 + 124u (64ths of an inch),
    R = 0.02 * 2^j * (1+v): piecewise-linear log-uniform on [0.02, 10.24) Mscf/stb, built only
    from exactly-rounded fp32 operations so that device and host agree bit for bit."""
    idx = np.arange(start, start + n, dtype=np.uint64)
    uP = hash_uniform(seed, STREAM_GILBERT_P, idx)
    uS = hash_uniform(seed, STREAM_GILBERT_S, idx)
    nR = hash_u24(seed, STREAM_GILBERT_R, idx).astype(np.uint64) * np.uint64(9)
    P = fma32(np.float32(4950.0), uP, np.float32(50.0))
    S = fma32(np.float32(124.0), uS, np.float32(4.0))
    j = (nR >> np.uint64(24)).astype(np.int32)
    v = ((nR & np.uint64(0xFFFFFF)) >> np.uint64(1)).astype(np.float32) * np.float32(2.0 ** -23)
    R = (np.float32(0.02) * np.ldexp(np.float32(1.0) + v, j).astype(np.float32)).astype(np.float32)
    return P, S, R


def teacher_weights(seed: int, n_features: int):
    idx = np.arange(n_features, dtype=np.uint64)
    u = hash_uniform(seed, STREAM_TEACHER, idx)
    return (np.float32(2.0) * u - np.float32(1.0)).astype(np.float32)


def synth_rows(seed: int, start: int, n: int, n_features: int):
    """Configs 1/3/4 rows.  X ~ U[-1,1) fp32 (n, F); y = softsign(x.w_t) + 0.05*(2u-1) with the
    dot product accumulated by sequential fp32 fma in feature order (device-reproducible).
    softsign replaces the survey's tanh because IEEE division is bit-reproducible and tanh is not."""
    rows = np.arange(start, start + n, dtype=np.uint64)
    idx = rows[:, None] * np.uint64(n_features) + np.arange(n_features, dtype=np.uint64)[None, :]
    X = (np.float32(2.0) * hash_uniform(seed, STREAM_X, idx) - np.float32(1.0)).astype(np.float32)
    wt = teacher_weights(seed, n_features)
    s = np.zeros(n, dtype=np.float32)
    for k in range(n_features):
        s = fma32(X[:, k], wt[k], s)
    base = (s / (np.float32(1.0) + np.abs(s))).astype(np.float32)
    un = hash_uniform(seed, STREAM_NOISE, rows)
    noise = (np.float32(2.0) * un - np.float32(1.0)).astype(np.float32)
    y = fma32(np.float32(0.05), noise, base)
    return X, y


def split_code(seed: int, rows):
    """cnn.py:68 randomSplit([0.64,0.16,0.2]) made deterministic: 0=train, 1=validation, 2=test."""
    n = hash_u24(seed, STREAM_SPLIT, np.asarray(rows, dtype=np.uint64)) % np.uint32(100)
    return np.where(n < 64, 0, np.where(n < 80, 1, 2)).astype(np.int8)


def batch_order(seed: int, epoch: int, n_batches: int):
    """Per-epoch shuffle (Keras fit(shuffle=True), cnn.py:127) pinned at batch granularity:
    batches are contiguous row blocks; their ORDER is permuted by sorting hash keys."""
    keys = splitmix64(stream_key(seed, STREAM_PERM + 1000 * (epoch + 1)) + np.arange(n_batches, dtype=np.uint64))
    return np.argsort(keys, kind="stable").astype(np.int64)


# --------------------------------------------------------------------------------------
# Gilbert's equation (BASELINE.json north_star; Readme.md:7-8)
# --------------------------------------------------------------------------------------
GILBERT_A, GILBERT_B, GILBERT_C = 1.89, 0.546, 435.0


def gilbert(P, S, R, dtype=np.float64):
    """q = P * S^1.89 / (435 * R^0.546).  Truth is dtype=float64 evaluated on the fp32-rounded
    inputs (input rounding alone moves q by up to 1.9e-7)."""
    P = np.asarray(P).astype(dtype)
    S = np.asarray(S).astype(dtype)
    R = np.asarray(R).astype(dtype)
    return P * np.power(S, dtype(GILBERT_A)) / (dtype(GILBERT_C) * np.power(R, dtype(GILBERT_B)))


# --------------------------------------------------------------------------------------
# Losses (per-sample loss, and d(sum of per-sample losses)/d(yhat))
# --------------------------------------------------------------------------------------
LOSS_MSE, LOSS_MAE_CLIP = 0, 1
MAE_CLIP_VALUE = 6.0  # cnn.py:31


def loss_and_grad(yhat, y, loss: int):
    """Returns (per_sample_loss (B,), dsum/dyhat (B, n_out)).  The batch mean (Keras) is applied by
    the caller as 1/B so that data-parallel shards can exchange plain sums."""
    d = yhat - y
    n_out = yhat.shape[1]
    if loss == LOSS_MSE:
        return (d * d).mean(axis=1), 2.0 * d / n_out
    if loss == LOSS_MAE_CLIP:  # cnn.py:29-32
        a = np.abs(d)
        per = np.clip(a, 0.0, MAE_CLIP_VALUE).mean(axis=1)
        g = np.sign(d) * (a <= MAE_CLIP_VALUE) / n_out
        return per, g.astype(d.dtype)
    raise ValueError("unknown loss")


# --------------------------------------------------------------------------------------
# Flat parameter layout (SURVEY.md section 8 a2 [PIN]): per layer W[in][out] row-major then b[out]
# --------------------------------------------------------------------------------------
def mlp_offsets(dims):
    offs, o = [], 0
    for i in range(len(dims) - 1):
        w = o
        o += dims[i] * dims[i + 1]
        b = o
        o += dims[i + 1]
        offs.append((w, b))
    return offs, o


def mlp_n_params(dims):
    return mlp_offsets(dims)[1]


def glorot_init(dims, seed: int, dtype=np.float32):
    """Keras Dense default (glorot_uniform weights, zero bias) from the counter hash."""
    offs, n = mlp_offsets(dims)
    p = np.zeros(n, dtype=np.float64)
    for li, (w, b) in enumerate(offs):
        fan_in, fan_out = dims[li], dims[li + 1]
        lim = np.sqrt(6.0 / (fan_in + fan_out))
        u = hash_uniform(seed, 100 + li, np.arange(fan_in * fan_out, dtype=np.uint64)).astype(np.float64)
        p[w:w + fan_in * fan_out] = (2.0 * u - 1.0) * lim
    return p.astype(dtype)


def mlp_forward(params, dims, X):
    """Hidden ReLU (cnn.py:111 is the only activation the reference uses), linear output (cnn.py:114).
    Returns list of activations [X, H1, ..., yhat]."""
    offs, _ = mlp_offsets(dims)
    acts = [X]
    h = X
    for li, (w, b) in enumerate(offs):
        W = params[w:w + dims[li] * dims[li + 1]].reshape(dims[li], dims[li + 1])
        z = h @ W + params[b:b + dims[li + 1]]
        h = np.maximum(z, 0.0) if li < len(offs) - 1 else z
        acts.append(h)
    return acts


def mlp_loss_grad_sums(params, dims, X, y, loss: int):
    """Sum (not mean) of per-sample losses and of per-sample gradients over the rows given.
    This is what one data-parallel shard contributes; divide by the GLOBAL batch afterwards."""
    params = np.asarray(params)
    dt = params.dtype
    X = np.asarray(X, dtype=dt)
    y = np.asarray(y, dtype=dt).reshape(X.shape[0], dims[-1])
    offs, n = mlp_offsets(dims)
    acts = mlp_forward(params, dims, X)
    per, g = loss_and_grad(acts[-1], y, loss)
    grads = np.zeros(n, dtype=dt)
    dz = g.astype(dt)
    for li in range(len(offs) - 1, -1, -1):
        w, b = offs[li]
        a_in = acts[li]
        grads[w:w + dims[li] * dims[li + 1]] = (a_in.T @ dz).reshape(-1)
        grads[b:b + dims[li + 1]] = dz.sum(axis=0)
        if li > 0:
            W = params[w:w + dims[li] * dims[li + 1]].reshape(dims[li], dims[li + 1])
            dz = (dz @ W.T) * (acts[li] > 0)
    return per.sum(), grads


# --------------------------------------------------------------------------------------
# Optimisers (Keras 1.x semantics; state dict is mutated in place)
# --------------------------------------------------------------------------------------
OPT_SGD, OPT_ADAM = 0, 1


def sgd_config(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True):  # cnn.py:117
    return dict(kind=OPT_SGD, lr=lr, momentum=momentum, decay=decay, nesterov=bool(nesterov))


def adam_config(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, decay=0.0):
    return dict(kind=OPT_ADAM, lr=lr, beta1=beta1, beta2=beta2, eps=eps, decay=decay)


def opt_state(n, dtype=np.float64):
    return dict(it=0, m=np.zeros(n, dtype=dtype), v=np.zeros(n, dtype=dtype))


def opt_step(params, grad_mean, cfg, st):
    """One optimiser update with the batch-MEAN gradient; `it` counts batches (Keras `iterations`)."""
    dt = params.dtype.type
    it = st["it"]
    lr_t = cfg["lr"] / (1.0 + cfg["decay"] * it)
    if cfg["kind"] == OPT_SGD:
        v = dt(cfg["momentum"]) * st["m"] - dt(lr_t) * grad_mean
        st["m"][...] = v
        if cfg["nesterov"]:
            params += dt(cfg["momentum"]) * v - dt(lr_t) * grad_mean
        else:
            params += v
    else:
        t = it + 1
        b1, b2 = cfg["beta1"], cfg["beta2"]
        lr_t = lr_t * np.sqrt(1.0 - b2 ** t) / (1.0 - b1 ** t)
        st["m"][...] = dt(b1) * st["m"] + dt(1.0 - b1) * grad_mean
        st["v"][...] = dt(b2) * st["v"] + dt(1.0 - b2) * grad_mean * grad_mean
        params -= dt(lr_t) * st["m"] / (np.sqrt(st["v"]) + dt(cfg["eps"]))
    st["it"] = it + 1
    return params


def mlp_fit_steps(params, dims, X, y, batch, n_steps, loss, cfg, st, first_batch=0, order=None, world=1):
    """cnn.py:127-128 mini-batch loop restated: for each step take the next contiguous block of
    `batch` rows (block order optionally permuted), mean-gradient step.  world>1 emulates the
    data-parallel path of SURVEY.md section 8e: each replica sums over its contiguous 1/world slice
    of the global batch, sums are added in rank order, then scaled by 1/B_global."""
    losses = []
    n_batches = X.shape[0] // batch
    for s in range(n_steps):
        bi = (first_batch + s) % n_batches
        if order is not None:
            bi = int(order[bi])
        Xb, yb = X[bi * batch:(bi + 1) * batch], y[bi * batch:(bi + 1) * batch]
        shard = batch // world
        lsum, gsum = 0.0, np.zeros_like(params)
        for r in range(world):
            l_r, g_r = mlp_loss_grad_sums(params, dims, Xb[r * shard:(r + 1) * shard], yb[r * shard:(r + 1) * shard], loss)
            lsum = lsum + l_r
            gsum = gsum + g_r
        losses.append(lsum / batch)
        opt_step(params, gsum / params.dtype.type(batch), cfg, st)
    return np.asarray(losses)


# --------------------------------------------------------------------------------------
# Dynamic ANN windowing (BASELINE.json config 4; SURVEY.md section 8d [PIN])
# --------------------------------------------------------------------------------------
def lag_windows(series, n_lags):
    """Sample i (i >= n_lags-1) = rows i-n_lags+1 .. i of the raw series flattened lag-major
    (oldest row first).  Returns (n - n_lags + 1, n_lags * F)."""
    n, f = series.shape
    idx = np.arange(n_lags)[None, :] + np.arange(n - n_lags + 1)[:, None]
    return series[idx].reshape(n - n_lags + 1, n_lags * f)


# --------------------------------------------------------------------------------------
# LSTM regressor (BASELINE.json config 5; SURVEY.md section 8 a6 [PIN])
# Flat layout: per layer  K[(D_l + H)][4H] row-major (rows 0..D_l-1 input kernel, then recurrent
# kernel), b[4H]; gate column blocks in Keras order i, f, c, o; then head w[H][n_out], b[n_out].
# --------------------------------------------------------------------------------------
def lstm_offsets(d_in, hidden, n_layers, n_out=1):
    offs, o = [], 0
    d = d_in
    for _ in range(n_layers):
        k = o
        o += (d + hidden) * 4 * hidden
        b = o
        o += 4 * hidden
        offs.append((k, b, d))
        d = hidden
    hw = o
    o += hidden * n_out
    hb = o
    o += n_out
    return offs, (hw, hb), o


def lstm_init(d_in, hidden, n_layers, seed, n_out=1, dtype=np.float32):
    offs, (hw, hb), n = lstm_offsets(d_in, hidden, n_layers, n_out)
    p = np.zeros(n, dtype=np.float64)
    for li, (k, b, d) in enumerate(offs):
        cnt = (d + hidden) * 4 * hidden
        lim = np.sqrt(6.0 / (d + hidden + 4 * hidden))
        u = hash_uniform(seed, 200 + li, np.arange(cnt, dtype=np.uint64)).astype(np.float64)
        p[k:k + cnt] = (2.0 * u - 1.0) * lim
        p[b + hidden:b + 2 * hidden] = 1.0  # unit forget bias
    lim = np.sqrt(6.0 / (hidden + n_out))
    u = hash_uniform(seed, 299, np.arange(hidden * n_out, dtype=np.uint64)).astype(np.float64)
    p[hw:hw + hidden * n_out] = (2.0 * u - 1.0) * lim
    return p.astype(dtype)


def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def lstm_forward(params, d_in, hidden, n_layers, X, n_out=1, keep=False):
    """X (B, T, d_in) batch-major.  Returns yhat (B, n_out) and, if keep, the BPTT stash."""
    dt = params.dtype
    B, T, _ = X.shape
    H = hidden
    offs, (hw, hb), _ = lstm_offsets(d_in, hidden, n_layers, n_out)
    seq = X.astype(dt)
    stash = []
    for (k, b, d) in offs:
        K = params[k:k + (d + H) * 4 * H].reshape(d + H, 4 * H)
        bias = params[b:b + 4 * H]
        h = np.zeros((B, H), dtype=dt)
        c = np.zeros((B, H), dtype=dt)
        hs = np.zeros((B, T, H), dtype=dt)
        rec = dict(i=[], f=[], g=[], o=[], c=[], cprev=[], hprev=[], x=seq)
        for t in range(T):
            z = seq[:, t, :] @ K[:d] + h @ K[d:] + bias
            i, f = _sigmoid(z[:, :H]), _sigmoid(z[:, H:2 * H])
            g, o = np.tanh(z[:, 2 * H:3 * H]), _sigmoid(z[:, 3 * H:])
            if keep:
                rec["cprev"].append(c)
                rec["hprev"].append(h)
            c = f * c + i * g
            h = o * np.tanh(c)
            hs[:, t, :] = h
            if keep:
                for nm, val in (("i", i), ("f", f), ("g", g), ("o", o), ("c", c)):
                    rec[nm].append(val)
        stash.append(rec)
        seq = hs
    Wh = params[hw:hw + H * n_out].reshape(H, n_out)
    yhat = seq[:, -1, :] @ Wh + params[hb:hb + n_out]
    return (yhat, stash, seq[:, -1, :]) if keep else yhat


def lstm_loss_grad_sums(params, d_in, hidden, n_layers, X, y, loss, n_out=1):
    """Sum of per-sequence losses and gradient sums via full BPTT (manual, no autograd)."""
    dt = params.dtype
    B, T, _ = X.shape
    H = hidden
    offs, (hw, hb), n = lstm_offsets(d_in, hidden, n_layers, n_out)
    yhat, stash, hlast = lstm_forward(params, d_in, hidden, n_layers, X, n_out, keep=True)
    y = np.asarray(y, dtype=dt).reshape(B, n_out)
    per, g = loss_and_grad(yhat, y, loss)
    grads = np.zeros(n, dtype=dt)
    Wh = params[hw:hw + H * n_out].reshape(H, n_out)
    grads[hw:hw + H * n_out] = (hlast.T @ g).reshape(-1)
    grads[hb:hb + n_out] = g.sum(axis=0)
    dseq = np.zeros((B, T, H), dtype=dt)
    dseq[:, -1, :] = g @ Wh.T
    for li in range(n_layers - 1, -1, -1):
        k, b, d = offs[li]
        K = params[k:k + (d + H) * 4 * H].reshape(d + H, 4 * H)
        rec = stash[li]
        dK = np.zeros_like(K)
        db = np.zeros(4 * H, dtype=dt)
        dh_next = np.zeros((B, H), dtype=dt)
        dc_next = np.zeros((B, H), dtype=dt)
        dx = np.zeros((B, T, d), dtype=dt)
        for t in range(T - 1, -1, -1):
            i, f, gg, o, c = rec["i"][t], rec["f"][t], rec["g"][t], rec["o"][t], rec["c"][t]
            dh = dseq[:, t, :] + dh_next
            tc = np.tanh(c)
            dc = dc_next + dh * o * (1.0 - tc * tc)
            dz = np.concatenate([dc * gg * i * (1.0 - i), dc * rec["cprev"][t] * f * (1.0 - f),
                                 dc * i * (1.0 - gg * gg), dh * tc * o * (1.0 - o)], axis=1)
            xin = np.concatenate([rec["x"][:, t, :], rec["hprev"][t]], axis=1)
            dK += xin.T @ dz
            db += dz.sum(axis=0)
            dxin = dz @ K.T
            dx[:, t, :] = dxin[:, :d]
            dh_next = dxin[:, d:]
            dc_next = dc * f
        grads[k:k + (d + H) * 4 * H] = dK.reshape(-1)
        grads[b:b + 4 * H] = db
        dseq = dx
    return per.sum(), grads


def lstm_fit_steps(params, d_in, hidden, n_layers, X, y, batch, n_steps, loss, cfg, st, world=1):
    losses = []
    n_batches = X.shape[0] // batch
    for s in range(n_steps):
        bi = s % n_batches
        Xb, yb = X[bi * batch:(bi + 1) * batch], y[bi * batch:(bi + 1) * batch]
        shard = batch // world
        lsum, gsum = 0.0, np.zeros_like(params)
        for r in range(world):
            l_r, g_r = lstm_loss_grad_sums(params, d_in, hidden, n_layers, Xb[r * shard:(r + 1) * shard],
                                           yb[r * shard:(r + 1) * shard], loss)
            lsum, gsum = lsum + l_r, gsum + g_r
        losses.append(lsum / batch)
        opt_step(params, gsum / params.dtype.type(batch), cfg, st)
    return np.asarray(losses)


def synth_series(seed: int, start: int, n_seq: int, T: int, d_in: int):
    """Config 5 inputs: AR(1)-smoothed hash noise x_t = 0.75 x_{t-1} + 0.25 e_t (fp32 fma), and
    y = softsign(mean over the last 8 steps of x . w_t)."""
    seqs = np.arange(start, start + n_seq, dtype=np.uint64)
    idx = (seqs[:, None, None] * np.uint64(T) + np.arange(T, dtype=np.uint64)[None, :, None]) * np.uint64(d_in) \
        + np.arange(d_in, dtype=np.uint64)[None, None, :]
    e = (np.float32(2.0) * hash_uniform(seed, STREAM_X, idx) - np.float32(1.0)).astype(np.float32)
    X = np.zeros((n_seq, T, d_in), dtype=np.float32)
    prev = np.zeros((n_seq, d_in), dtype=np.float32)
    for t in range(T):
        prev = fma32(np.float32(0.75), prev, (np.float32(0.25) * e[:, t, :]).astype(np.float32))
        X[:, t, :] = prev
    wt = teacher_weights(seed, d_in)
    s = np.zeros(n_seq, dtype=np.float32)
    for t in range(T - 8, T):
        for k in range(d_in):
            s = fma32(X[:, t, k], wt[k], s)
    y = (s / (np.float32(1.0) + np.abs(s))).astype(np.float32)
    return X, y


# --------------------------------------------------------------------------------------
# Epoch-level host logic (cnn.py:121-123): EarlyStopping(patience=10) + best-only checkpoint
# --------------------------------------------------------------------------------------
def early_stopping_trace(val_losses, patience=10):
    """Returns (stop_epoch_index or None, best_epoch_index): Keras EarlyStopping(monitor='val_loss')
    waits `patience` epochs without strict improvement; ModelCheckpoint(save_best_only) saves on
    strict improvement."""
    best, best_ep, wait = np.inf, -1, 0
    for ep, v in enumerate(val_losses):
        if v < best:
            best, best_ep, wait = v, ep, 0
        else:
            wait += 1
            if wait >= patience:
                return ep, best_ep
    return None, best_ep


# --------------------------------------------------------------------------------------
# Periodic PARAMETER averaging (BASELINE.json "gradient/parameter averaging"): what a Spark
# driver does between partitions -- every replica runs `every` local steps on its own slice of each global
# batch (mean over its LOCAL rows), then parameters AND optimiser state are replaced by their mean over replicas.
# The snapshot has no averaging code; this pins the optional mode.  With every=1 and plain
# SGD (no momentum) it coincides with the default synchronous gradient averaging.
# --------------------------------------------------------------------------------------
def mlp_fit_steps_param_avg(params, dims, X, y, batch, n_steps, loss, cfg, world, every):
    reps = [params.copy() for _ in range(world)]
    sts = [opt_state(params.size, params.dtype) for _ in range(world)]
    shard = batch // world
    losses = []
    for s in range(n_steps):
        Xb, yb = X[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch]
        lsum = 0.0
        for r in range(world):
            l_r, g_r = mlp_loss_grad_sums(reps[r], dims, Xb[r * shard:(r + 1) * shard], yb[r * shard:(r + 1) * shard], loss)
            lsum += l_r
            opt_step(reps[r], g_r / params.dtype.type(shard), cfg, sts[r])
        losses.append(lsum / batch)
        if (s + 1) % every == 0:
            for arrs in ([reps[r] for r in range(world)], [sts[r]["m"] for r in range(world)], [sts[r]["v"] for r in range(world)]):
                mean = sum(arrs) / world
                for a in arrs:
                    a[...] = mean
    params[...] = sum(reps) / world
    return np.asarray(losses)


# --------------------------------------------------------------------------------------
# The CNN that is actually in the snapshot (cnn.py:110-114):
#   Convolution1D(1, 100, 13, activation="relu") -> Dropout(0.5) -> Flatten -> Dense(3600, 12)
# on (48, 1) inputs, valid border.  Flat layout: Wc[K][F] (tap-major), bc[F], Wd[T*F][n_out], bd[n_out],
# T = L - K + 1; Flatten is time-major (index t*F + f), as Keras flattens (steps, filters).
# Pinned where Keras-0.x/Theano cannot be consulted: cross-correlation tap order (Theano's conv2d flips the
# filter -- a relabelling of learned weights); inverted dropout (x * mask / keep at train time, identity at
# test time); the mask is the counter hash of (seed, sample id, unit) so CUDA and oracle draw the same one.
# --------------------------------------------------------------------------------------
STREAM_DROPOUT = 300


def cnn_offsets(L, K, F, n_out):
    T = L - K + 1
    wc, bc = 0, K * F
    wd = bc + F
    bd = wd + T * F * n_out
    return (wc, bc, wd, bd), bd + n_out


def cnn_init(L, K, F, n_out, seed, dtype=np.float32):
    (wc, bc, wd, bd), n = cnn_offsets(L, K, F, n_out)
    T = L - K + 1
    p = np.zeros(n, dtype=np.float64)
    lim = np.sqrt(6.0 / (K + K * F))
    p[wc:wc + K * F] = (2.0 * hash_uniform(seed, 400, np.arange(K * F, dtype=np.uint64)).astype(np.float64) - 1.0) * lim
    lim = np.sqrt(6.0 / (T * F + n_out))
    p[wd:wd + T * F * n_out] = (2.0 * hash_uniform(seed, 401, np.arange(T * F * n_out, dtype=np.uint64)).astype(np.float64) - 1.0) * lim
    return p.astype(dtype)


def dropout_mask(seed, sample0, n_rows, n_units, keep=0.5):
    """mask[r][j] = 1 if the 24-bit hash of (sample0 + r) * n_units + j is below keep * 2^24."""
    idx = (np.uint64(sample0) + np.arange(n_rows, dtype=np.uint64))[:, None] * np.uint64(n_units) + np.arange(n_units, dtype=np.uint64)[None, :]
    return (hash_u24(seed, STREAM_DROPOUT, idx) < np.uint32(keep * (1 << 24))).astype(np.float64)


def cnn_forward(params, L, K, F, n_out, X, mask=None, keep=0.5):
    (wc, bc, wd, bd), _ = cnn_offsets(L, K, F, n_out)
    T = L - K + 1
    Wc = params[wc:wc + K * F].reshape(K, F)
    win = np.stack([X[:, k:k + T] for k in range(K)], axis=2)          # (B, T, K)
    conv = win @ Wc + params[bc:bc + F]                                 # (B, T, F)
    a = np.maximum(conv, 0.0).reshape(X.shape[0], T * F)
    if mask is not None:
        a = a * mask / keep
    out = a @ params[wd:wd + T * F * n_out].reshape(T * F, n_out) + params[bd:bd + n_out]
    return out, a, win


def cnn_loss_grad_sums(params, L, K, F, n_out, X, y, loss, mask=None, keep=0.5):
    dt = params.dtype
    (wc, bc, wd, bd), n = cnn_offsets(L, K, F, n_out)
    T = L - K + 1
    out, a, win = cnn_forward(params, L, K, F, n_out, X.astype(dt), mask, keep)
    per, g = loss_and_grad(out, np.asarray(y, dtype=dt).reshape(X.shape[0], n_out), loss)
    grads = np.zeros(n, dtype=dt)
    Wd = params[wd:wd + T * F * n_out].reshape(T * F, n_out)
    grads[wd:wd + T * F * n_out] = (a.T @ g).reshape(-1)
    grads[bd:bd + n_out] = g.sum(axis=0)
    da = g @ Wd.T
    if mask is not None:
        da = da * mask / keep
    dconv = (da * (a > 0 if mask is None else (cnn_forward(params, L, K, F, n_out, X.astype(dt))[1] > 0))).reshape(-1, T, F)
    grads[wc:wc + K * F] = np.einsum("btk,btf->kf", win, dconv).reshape(-1)
    grads[bc:bc + F] = dconv.sum(axis=(0, 1))
    return per.sum(), grads
