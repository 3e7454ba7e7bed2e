"""Shared trainer behind the non-MLP model families (LSTM stack, the script's CNN): flat fp32 parameter
vector, per-batch gradient SUMS from a dlas_*_grad call, cross-rank sum (dlas/parallel.py), one
dlas_opt_step -- the same step shape as Sequential's MLP path (cnn.py:117-132)."""
from __future__ import annotations

import time

import numpy as np
import torch

from . import _lib, hashing, ops, parallel


class FlatModel:
    """Subclasses provide: _count_params(), _init_params(), _make_workspace(n_rows), _shape_x(X), n_out,
    _grad(X, y, grad_sum, ws, train, global_rows), _forward(X, ws), _describe() (checkpoint header)."""
    n_out = 1

    def __init__(self, seed=ops.BASE_SEED, device=None):
        self.seed, self.device = seed, device
        self.params, self.loss, self.iterations = None, None, 0
        self._ws, self._ws_rows = None, -1

    def compile(self, loss, optimizer):
        from .models import _LOSSES
        self.loss = _LOSSES[loss]
        self.cfg = optimizer.cfg()
        _lib.load()
        if not torch.cuda.is_available():
            raise RuntimeError("dlas needs a B200 GPU: there is no CPU path")
        self.device = torch.device(self.device if self.device is not None else "cuda:%d" % torch.cuda.current_device())
        n = self._count_params()
        self.n_params = n
        if self.params is None:
            self.params = torch.from_numpy(self._init_params()).to(self.device)
        self.m = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.v = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.grad_sum = torch.zeros(n + 1, dtype=torch.float32, device=self.device)
        self.iterations = 0
        self.rank, self.world = parallel.world_info()
        return self

    def _workspace(self, n_rows):
        if n_rows > self._ws_rows:
            self._ws = self._make_workspace(n_rows)
            self._ws_rows = n_rows
        return self._ws

    def _dev(self, x, y=None):
        X = self._shape_x(torch.as_tensor(x, dtype=torch.float32).to(self.device)).contiguous()
        if y is None:
            return X
        return X, torch.as_tensor(y, dtype=torch.float32).to(self.device).reshape(-1, self.n_out).contiguous()

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
            lambda: self._grad(X, y, self.grad_sum, ws, True, rows),
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
            if self.world == 1 and n_full and hasattr(self, "_fit_steps"):
                # single GPU: the whole epoch's batch loop is one library call (no Python between the steps)
                losses.append(self._fit_steps(X, Y, lb, [int(b) for b in order[:n_full]]))
                weights += [lb] * n_full
                self.iterations += n_full
            else:
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
        return self._forward(X, self._workspace(X.shape[0]))

    def evaluate(self, x, y):
        """Mean loss in test mode: the loss sum comes out of the same kernels (grad_sum[n])."""
        X, Y = self._dev(x, y)
        scratch = torch.empty_like(self.grad_sum)
        self._grad(X, Y, scratch, self._workspace(X.shape[0]), False, X.shape[0])
        pack = torch.stack([scratch[-1], torch.tensor(float(X.shape[0]), device=self.device)])
        parallel.allreduce_sum_(pack)
        return float(pack[0].item() / max(1.0, pack[1].item()))

    def test_on_batch(self, x, y):
        return self.evaluate(x, y)

    def save_weights(self, path):
        import os
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        blob = dict(self._describe())
        blob["params"] = self.params.detach().cpu()
        torch.save(blob, path)
