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

import time

import numpy as np
import torch

from . import _lib, hashing, ops, parallel


class LSTM:
    def __init__(self, units, input_shape=None, return_sequences=False, input_dim=None, input_length=None):
        self.units = int(units)
        if input_shape is None and input_dim is not None:
            input_shape = (input_length, input_dim)
        self.input_shape = input_shape
        self.return_sequences = return_sequences


class LSTMRegressor:
    def __init__(self, d_in, hidden, n_layers, T, seed=ops.BASE_SEED, device=None):
        self.d_in, self.hidden, self.n_layers, self.T = int(d_in), int(hidden), int(n_layers), int(T)
        self.seed, self.device = seed, device
        self.params, self.loss, self.iterations = None, None, 0
        self._ws, self._ws_rows = None, -1

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

    def compile(self, loss, optimizer):
        from .models import _LOSSES
        self.loss = _LOSSES[loss]
        self.cfg = optimizer.cfg()
        _lib.load()
        if not torch.cuda.is_available():
            raise RuntimeError("dlas needs a B200 GPU: there is no CPU path")
        self.device = torch.device(self.device if self.device is not None else "cuda:%d" % torch.cuda.current_device())
        n = ops.lstm_n_params(self.d_in, self.hidden, self.n_layers)
        if n == 0:
            raise _lib.DlasError(-5, "LSTM shape not supported")
        self.n_params = n
        if self.params is None:
            self.params = torch.from_numpy(self._init_params()).to(self.device)
        self.m = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.v = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.grad_sum = torch.zeros(n + 1, dtype=torch.float32, device=self.device)
        self.iterations = 0
        self.rank, self.world = parallel.world_info()
        return self

    def _workspace(self, n_seq):
        if n_seq > self._ws_rows:
            self._ws = ops.lstm_workspace(self.d_in, self.hidden, self.n_layers, self.T, n_seq, self.device)
            self._ws_rows = n_seq
        return self._ws

    def _dev(self, x, y=None):
        X = torch.as_tensor(x, dtype=torch.float32).to(self.device).reshape(-1, self.T, self.d_in).contiguous()
        if y is None:
            return X
        return X, torch.as_tensor(y, dtype=torch.float32).to(self.device).reshape(-1, 1).contiguous()

    def set_weights_flat(self, flat):
        t = torch.as_tensor(np.asarray(flat, dtype=np.float32))
        if self.params is None:
            self.params = t.to(self.device if self.device is not None else "cuda")
        else:
            self.params.copy_(t)

    def get_weights_flat(self):
        return self.params.detach().cpu().numpy()

    def train_on_batch(self, X, y, global_rows=None):
        rows = global_rows if global_rows is not None else X.shape[0] * self.world
        ws = self._workspace(X.shape[0])
        loss_out = torch.empty(1, dtype=torch.float32, device=self.device)
        parallel.dp_step(
            lambda: ops.lstm_grad(self.d_in, self.hidden, self.n_layers, self.T, self.loss, self.params, X, y,
                                  self.grad_sum, ws),
            lambda scale: ops.opt_step(self.cfg, self.params, self.m, self.v, self.grad_sum, scale, self.iterations,
                                       loss_out),
            self.grad_sum, rows)
        self.iterations += 1
        return loss_out

    def fit(self, x, y, validation_data=None, verbose=2, callbacks=None, nb_epoch=None, epochs=None, batch_size=32,
            shuffle=True, log=print):
        n_epochs = nb_epoch if nb_epoch is not None else (epochs if epochs is not None else 1)
        X, Y = self._dev(x, y)
        lb = batch_size // self.world
        if lb < 1 or lb * self.world != batch_size:
            raise ValueError("batch_size %d must be a positive multiple of the world size %d" % (batch_size, self.world))
        n_local = X.shape[0]
        n_full, rem = n_local // lb, n_local % lb
        history = {"loss": [], "val_loss": []}
        for epoch in range(n_epochs):
            t0 = time.time()
            order = hashing.batch_order(self.seed, epoch, n_full) if shuffle else np.arange(n_full)
            losses, weights = [], []
            for s in range(n_full):
                b = int(order[s])
                losses.append(self.train_on_batch(X[b * lb:(b + 1) * lb], Y[b * lb:(b + 1) * lb], global_rows=batch_size))
                weights.append(lb)
            if rem:
                losses.append(self.train_on_batch(X[n_full * lb:], Y[n_full * lb:], global_rows=rem * self.world))
                weights.append(rem)
            lv = torch.cat(losses).cpu().numpy()
            logs = {"loss": float((lv * np.asarray(weights)).sum() / max(1, n_local))}
            if validation_data is not None:
                logs["val_loss"] = self.evaluate(*validation_data)
            history["loss"].append(logs["loss"])
            history["val_loss"].append(logs.get("val_loss"))
            if verbose and self.rank == 0:
                log("Epoch %d/%d - %.1fs - loss: %.6f%s" % (epoch + 1, n_epochs, time.time() - t0, logs["loss"],
                    (" - val_loss: %.6f" % logs["val_loss"]) if "val_loss" in logs else ""))
            stop = False
            for cb in callbacks or []:
                stop = bool(cb.on_epoch_end(self, epoch, logs)) or stop
            if stop:
                break
        return history

    def predict(self, x):
        X = self._dev(x)
        return ops.lstm_predict(self.d_in, self.hidden, self.n_layers, self.T, self.params, X, self._workspace(X.shape[0]))

    def evaluate(self, x, y):
        """Mean loss: the loss sum comes out of the same forward/backward kernels (grad_sum[n])."""
        X, Y = self._dev(x, y)
        scratch = torch.empty_like(self.grad_sum)
        ops.lstm_grad(self.d_in, self.hidden, self.n_layers, self.T, self.loss, self.params, X, Y, scratch,
                      self._workspace(X.shape[0]))
        pack = torch.stack([scratch[-1], torch.tensor(float(X.shape[0]), device=self.device)])
        parallel.allreduce_sum_(pack)
        return float(pack[0].item() / max(1.0, pack[1].item()))

    test_on_batch = evaluate

    def save_weights(self, path):
        import os
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        torch.save({"lstm": [self.d_in, self.hidden, self.n_layers, self.T], "params": self.params.detach().cpu()}, path)
