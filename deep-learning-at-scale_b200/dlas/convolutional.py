"""The script's own model (cnn.py:110-114), Keras-0.x call shape:

    model = Sequential()
    model.add(Convolution1D(1, 100, 13, activation='relu', border_mode='valid', input_length=48))
    model.add(Dropout(0.5))
    model.add(Flatten())
    model.add(Dense(3600, 12))
    model.compile(loss=mae_clip, optimizer=SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))

Sequential delegates to CnnRegressor when it holds a Convolution1D.  Flat parameters: Wc[K][F], bc[F],
Wd[T*F][n_out] (Flatten is time-major), bd[n_out]; glorot_uniform kernels, zero biases.  Dropout is inverted
and its mask is a counter hash of (seed, global sample index of the step, unit): every rank and every step
draws its own bits and a run is reproducible from the seed (the oracle evaluates the same hash)."""
from __future__ import annotations

import numpy as np

from . import hashing, ops
from .flatmodel import FlatModel


class Convolution1D:
    """Convolution1D(input_dim, nb_filter, filter_length, activation=, border_mode=) as cnn.py:111; the length
    of the input series comes from `input_length` here or from the first fit/predict call."""

    def __init__(self, input_dim, nb_filter, filter_length, activation=None, border_mode="valid", input_length=None):
        if int(input_dim) != 1:
            raise NotImplementedError("one input channel on this path (cnn.py:111 passes input_dim=1)")
        if border_mode != "valid":
            raise NotImplementedError("border_mode='valid' only (cnn.py:111)")
        if activation != "relu":
            raise NotImplementedError("Convolution1D activation is ReLU on this path (cnn.py:111)")
        self.nb_filter, self.filter_length, self.input_length = int(nb_filter), int(filter_length), input_length


class Dropout:
    def __init__(self, p):
        if not 0.0 <= p < 1.0:
            raise ValueError("dropout rate must be in [0, 1)")
        self.p = float(p)


class Flatten:
    pass


class CnnRegressor(FlatModel):
    def __init__(self, L, K, F, n_out, drop=0.5, seed=ops.BASE_SEED, device=None):
        super().__init__(seed, device)
        self.L, self.K, self.F, self.n_out = int(L), int(K), int(F), int(n_out)
        self.keep = 1.0 - float(drop)

    @property
    def shape(self):
        return (self.L, self.K, self.F, self.n_out)

    def _count_params(self):
        return ops.cnn_n_params(*self.shape)

    def _init_params(self):
        K, F, TF, O = self.K, self.F, (self.L - self.K + 1) * self.F, self.n_out
        lim = np.sqrt(6.0 / (K + K * F))
        wc = (2.0 * hashing.uniform(self.seed, 400, K * F).astype(np.float64) - 1.0) * lim
        lim = np.sqrt(6.0 / (TF + O))
        wd = (2.0 * hashing.uniform(self.seed, 401, TF * O).astype(np.float64) - 1.0) * lim
        return np.concatenate([wc, np.zeros(F), wd, np.zeros(O)]).astype(np.float32)

    def _make_workspace(self, n_rows):
        return ops.cnn_workspace(*self.shape, n_rows, self.device)

    def _shape_x(self, X):
        return X.reshape(-1, self.L)       # (n, L) or Keras' (n, L, 1)

    def _grad(self, X, y, grad_sum, ws, train, global_rows):
        keep = self.keep if train else 1.0
        sample0 = self.iterations * int(global_rows) + self.rank * X.shape[0]
        return ops.cnn_grad(*self.shape, self.loss, self.params, X, y, grad_sum, ws, keep=keep,
                            dropout_seed=self.seed, sample0=sample0)

    def _fit_steps(self, X, Y, batch, order):
        """cnn.py:127-128's inner loop over the batches `order` of one epoch, in ONE call (dlas_cnn_fit_steps)."""
        import torch
        losses = torch.empty(len(order), dtype=torch.float32, device=self.device)
        ops.cnn_fit_steps(*self.shape, self.loss, self.cfg, self.params, self.m, self.v, X, Y, batch, order, self.iterations,
                          self.keep, self.seed, losses, self._workspace(batch))
        return losses

    def _forward(self, X, ws):
        return ops.cnn_predict(*self.shape, self.params, X, ws)

    def _describe(self):
        return {"cnn": list(self.shape) + [1.0 - self.keep]}
