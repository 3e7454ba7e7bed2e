"""LSTM flow regressor (Readme.md:21; BASELINE.json config 5): stacked LSTM layers over
(n_seq, T, d_in) series, last hidden state -> Dense(hidden -> 1).  Keras call shape:

    model = Sequential()
    model.add(LSTM(128, input_shape=(48, 8), return_sequences=True))
    model.add(LSTM(128))
    model.add(Dense(1))
    model.compile(loss="mse", optimizer=SGD(...)); model.fit(X, y, batch_size=4096, nb_epoch=...)

Sequential delegates to LSTMRegressor when it holds LSTM layers.  Gate order i,f,c,o, sigmoid
recurrent activation, tanh activation, unit forget bias."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, hashing, ops
from .flatmodel import FlatModel


class LSTM:
    def __init__(self, units, input_shape=None, return_sequences=False, input_dim=None, input_length=None):
        self.units = int(units)
        if input_shape is None and input_dim is not None:
            input_shape = (input_length, input_dim)
        self.input_shape = input_shape
        self.return_sequences = return_sequences


class LSTMRegressor(FlatModel):
    def __init__(self, d_in, hidden, n_layers, T, seed=ops.BASE_SEED, device=None, mode=_lib.MODE_FP32):
        super().__init__(seed, device)
        self.d_in, self.hidden, self.n_layers, self.T = int(d_in), int(hidden), int(n_layers), int(T)
        self.mode = int(mode)     # MODE_TC: the stages with a tcgen05 kernel run on the tensor cores (include/dlas_b200.h)

    def _init_params(self):
        H, p, o, d = self.hidden, [], 0, self.d_in
        for li in range(self.n_layers):
            cnt = (d + H) * 4 * H
            lim = np.sqrt(6.0 / (d + H + 4 * H))
            u = hashing.uniform(self.seed, 200 + li, cnt).astype(np.float64)
            b = np.zeros(4 * H, dtype=np.float32)
            b[H:2 * H] = 1.0
            p += [((2.0 * u - 1.0) * lim).astype(np.float32), b]
            d = H
        lim = np.sqrt(6.0 / (H + 1))
        u = hashing.uniform(self.seed, 299, H).astype(np.float64)
        p += [((2.0 * u - 1.0) * lim).astype(np.float32), np.zeros(1, dtype=np.float32)]
        return np.concatenate(p)

    def _count_params(self):
        n = ops.lstm_n_params(self.d_in, self.hidden, self.n_layers)
        if n == 0:
            raise _lib.DlasError(-5, "LSTM shape not supported")
        if not ops.lstm_mode_supported(self.d_in, self.hidden, self.n_layers, self.mode):
            raise _lib.DlasError(-5, "LSTM d_in=%d hidden=%d is not supported in mode %d" % (self.d_in, self.hidden, self.mode))
        return n

    def _make_workspace(self, n_seq):
        return ops.lstm_workspace(self.d_in, self.hidden, self.n_layers, self.T, n_seq, self.device)

    def _shape_x(self, X):
        return X.reshape(-1, self.T, self.d_in)

    def _grad(self, X, y, grad_sum, ws, train, global_rows):
        return ops.lstm_grad(self.d_in, self.hidden, self.n_layers, self.T, self.loss, self.params, X, y, grad_sum, ws, mode=self.mode)

    def _forward(self, X, ws):
        return ops.lstm_predict(self.d_in, self.hidden, self.n_layers, self.T, self.params, X, ws, mode=self.mode)

    def _describe(self):
        return {"lstm": [self.d_in, self.hidden, self.n_layers, self.T]}
