"""Host-side mirror of the Keras call sequence the reference scripts use (cnn.py:110-132):

    model = Sequential(); model.add(Dense(in, out, activation="relu")); ...
    model.compile(loss=mae_clip, optimizer=SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    model.fit(X, y, validation_data=(Xv, yv), verbose=2, callbacks=[EarlyStopping(...),
              ModelCheckpoint(path, save_best_only=True)], nb_epoch=1000, batch_size=20)
    model.test_on_batch(Xt, yt)

Same names, argument meaning and defaults; everything numeric runs in libdlas_b200.so on the GPU.
Data-parallel training (one process per GPU, torch.distributed): each rank passes ITS shard of
every global batch; gradient sums are all-reduced once per step in a single flat bucket and the
optimiser sees the global-batch mean."""
from __future__ import annotations

import os
import time

import numpy as np
import torch

from . import _lib, hashing, ops, parallel


# ---- losses / optimisers / layers / callbacks (Keras names) -----------------------------------
def mae_clip(y_true=None, y_pred=None):
    """cnn.py:29-32: mean(clip(|y_true - y_pred|, 0, 6), axis=-1).  Passed to compile(loss=mae_clip);
    the arithmetic is the kernel's loss epilogue, this object only selects it."""
    raise RuntimeError("mae_clip is evaluated inside the CUDA kernels; pass it to compile(loss=mae_clip)")


_LOSSES = {"mse": _lib.LOSS_MSE, "mean_squared_error": _lib.LOSS_MSE, "mae_clip": _lib.LOSS_MAE_CLIP,
           mae_clip: _lib.LOSS_MAE_CLIP}


class SGD:
    """keras.optimizers.SGD as called at cnn.py:117."""

    def __init__(self, lr=0.01, momentum=0.0, decay=0.0, nesterov=False):
        self.lr, self.momentum, self.decay, self.nesterov = lr, momentum, decay, nesterov

    def cfg(self):
        return ops.make_opt_cfg(_lib.OPT_SGD, self.lr, self.momentum, self.decay, self.nesterov)


class Adam:
    """keras.optimizers.Adam (epsilon added outside the bias-corrected square root)."""

    def __init__(self, lr=0.001, beta_1=0.9, beta_2=0.999, epsilon=1e-7, decay=0.0):
        self.lr, self.beta_1, self.beta_2, self.epsilon, self.decay = lr, beta_1, beta_2, epsilon, decay

    def cfg(self):
        return ops.make_opt_cfg(_lib.OPT_ADAM, self.lr, 0.0, self.decay, False, self.beta_1, self.beta_2, self.epsilon)


class Dense:
    """Dense(input_dim, output_dim) -- the Keras-0.x positional form of cnn.py:114 -- or
    Dense(output_dim, input_dim=...) / Dense(units) after a first layer."""

    def __init__(self, *args, activation=None, input_dim=None):
        if len(args) == 2:
            self.input_dim, self.output_dim = int(args[0]), int(args[1])
        elif len(args) == 1:
            self.input_dim, self.output_dim = (None if input_dim is None else int(input_dim)), int(args[0])
        else:
            raise TypeError("Dense(input_dim, output_dim) or Dense(output_dim, input_dim=...)")
        if activation not in (None, "linear", "relu"):
            raise NotImplementedError("activation %r: the kernels implement relu (hidden) and linear (output)" % (activation,))
        self.activation = "linear" if activation is None else activation


class EarlyStopping:
    """keras.callbacks.EarlyStopping(monitor='val_loss', patience=10), cnn.py:121."""

    def __init__(self, monitor="val_loss", patience=0):
        self.monitor, self.patience = monitor, patience
        self.best, self.wait, self.stopped_epoch = np.inf, 0, None

    def on_epoch_end(self, model, epoch, logs):
        cur = logs.get(self.monitor)
        if cur is None:
            return False
        if cur < self.best:
            self.best, self.wait = cur, 0
            return False
        self.wait += 1
        if self.wait >= self.patience:
            self.stopped_epoch = epoch
            return True
        return False


class ModelCheckpoint:
    """keras.callbacks.ModelCheckpoint(path, save_best_only=True), cnn.py:122.  File format: a
    torch.save dict {dims, params} (the reference's HDF5 writer is not available)."""

    def __init__(self, filepath, monitor="val_loss", save_best_only=False):
        self.filepath, self.monitor, self.save_best_only = filepath, monitor, save_best_only
        self.best = np.inf

    def on_epoch_end(self, model, epoch, logs):
        cur = logs.get(self.monitor)
        if self.save_best_only:
            if cur is None or not cur < self.best:
                return False
            self.best = cur
        if model.rank == 0:
            model.save_weights(self.filepath)
        return False


class GraphStepper:
    """Replays a captured training step (Sequential.graph_stepper) -- or, when the whole step is the library's two
    launches with device-resident state (dlas_mlp_dp_step_at), enqueues n steps with ONE call and no graph: launches
    chained by programmatic dependent launch overlap their prologues, which replays of a one-step graph do not."""

    def __init__(self, model, graph, state, ring, enqueue=None):
        self.model, self.graph, self.state, self.ring, self.enqueue = model, graph, state, ring, enqueue

    def run(self, n_steps):
        if self.enqueue is not None:
            self.enqueue(n_steps)
        else:
            for _ in range(n_steps):
                self.graph.replay()
        self.model.iterations += n_steps

    def recent_losses(self, n):
        """The batch-mean losses of the last n steps (device ring buffer -> host)."""
        it = int(self.state[1].item()) if self.state is not None else self.model.iterations
        idx = torch.arange(it - n, it, device=self.ring.device) % self.ring.numel()
        return self.ring[idx].cpu().numpy()


# ---- the model -------------------------------------------------------------------------------
class Sequential:
    def __init__(self, layers=None, mode="fp32", seed=ops.BASE_SEED, device=None):
        self.layers = []
        self.mode = {"fp32": _lib.MODE_FP32, "tc": _lib.MODE_TC}[mode] if isinstance(mode, str) else int(mode)
        self.seed = seed
        self.device = device
        self.params = None
        self.loss = None
        self.iterations = 0
        self._impl = None   # LSTMRegressor when the stack holds LSTM layers
        for l in layers or []:
            self.add(l)

    # -- build (cnn.py:110-114)
    def add(self, layer):
        from .recurrent import LSTM
        from .convolutional import Convolution1D, Dropout, Flatten
        if isinstance(layer, LSTM) or any(isinstance(l, LSTM) for l in self.layers):
            if isinstance(layer, Dense) and layer.input_dim is None:
                layer.input_dim = self.layers[-1].units
            self.layers.append(layer)
            return
        if isinstance(layer, (Convolution1D, Dropout, Flatten)) or any(isinstance(l, Convolution1D) for l in self.layers):
            self.layers.append(layer)       # the stack of cnn.py:110-114, checked in compile
            return
        if not isinstance(layer, Dense):
            raise NotImplementedError("Dense, LSTM and the Convolution1D/Dropout/Flatten stack of cnn.py are on the B200 hot path")
        if layer.input_dim is None:
            if not self.layers:
                raise ValueError("first layer needs input_dim")
            layer.input_dim = self.layers[-1].output_dim
        elif self.layers and layer.input_dim != self.layers[-1].output_dim:
            raise ValueError("layer input_dim %d != previous output_dim %d" % (layer.input_dim, self.layers[-1].output_dim))
        self.layers.append(layer)

    @property
    def dims(self):
        return [self.layers[0].input_dim] + [l.output_dim for l in self.layers]

    # -- compile (cnn.py:118)
    def _compile_lstm(self, loss, optimizer):
        from .recurrent import LSTM, LSTMRegressor
        rec = [l for l in self.layers if isinstance(l, LSTM)]
        head = self.layers[len(rec):]
        if self.layers[:len(rec)] != rec or len(head) != 1 or not isinstance(head[0], Dense) or head[0].output_dim != 1:
            raise NotImplementedError("supported stack: LSTM layers followed by one Dense(1)")
        if rec[0].input_shape is None or len({l.units for l in rec}) != 1:
            raise NotImplementedError("first LSTM needs input_shape=(T, d_in); all LSTM layers share one width")
        T, d_in = rec[0].input_shape
        return self._delegate(LSTMRegressor(d_in, rec[0].units, len(rec), T, seed=self.seed, device=self.device, mode=self.mode),
                              loss, optimizer)

    def _delegate(self, impl, loss, optimizer):
        self._impl = impl
        if self.params is not None:
            impl.params = self.params
        impl.compile(loss, optimizer if optimizer is not None else SGD())
        for name in ("fit", "predict", "evaluate", "test_on_batch", "train_on_batch", "set_weights_flat",
                     "get_weights_flat", "save_weights"):
            setattr(self, name, getattr(impl, name))
        return self

    def _compile_cnn(self, loss, optimizer):
        from .convolutional import CnnRegressor, Convolution1D, Dropout, Flatten
        kinds = [type(l) for l in self.layers]
        if kinds not in ([Convolution1D, Dropout, Flatten, Dense], [Convolution1D, Flatten, Dense]):
            raise NotImplementedError("supported stack: Convolution1D -> [Dropout] -> Flatten -> Dense (cnn.py:110-114)")
        conv, dense = self.layers[0], self.layers[-1]
        if dense.activation != "linear":
            raise NotImplementedError("the output layer is linear on this path (cnn.py:114)")
        L = conv.input_length
        if dense.input_dim is not None:      # Dense(3600, 12): T*F fixes the series length
            if dense.input_dim % conv.nb_filter:
                raise ValueError("Dense input_dim %d is not a multiple of nb_filter %d" % (dense.input_dim, conv.nb_filter))
            L_dense = dense.input_dim // conv.nb_filter + conv.filter_length - 1
            if L is not None and L != L_dense:
                raise ValueError("input_length %d disagrees with Dense input_dim %d" % (L, dense.input_dim))
            L = L_dense
        if L is None:
            raise ValueError("give Convolution1D(input_length=...) or Dense(input_dim, output_dim)")
        drop = self.layers[1].p if len(self.layers) == 4 else 0.0
        return self._delegate(CnnRegressor(L, conv.filter_length, conv.nb_filter, dense.output_dim, drop, seed=self.seed,
                                           device=self.device), loss, optimizer)

    def compile(self, loss="mse", optimizer=None, averaging="gradient", average_every=1, exchange="auto"):
        """averaging="gradient" (default): synchronous mean-gradient averaging every step, the parity semantics of
        SURVEY.md 8e.  averaging="parameter": Spark-driver-style periodic parameter averaging -- `average_every`
        local steps per rank between averages of parameters and optimiser state.
        exchange (world > 1, gradient averaging): "auto" = the library's two-launch step with the NVLink push/poll
        exchange inside its finishing kernel (dlas_mlp_dp_step_at) where the mode supports it and the ranks' memory
        can be mapped, else NCCL; "peer" = require it; "nccl" = all-reduce of the flat bucket between our kernels."""
        if averaging not in ("gradient", "parameter"):
            raise ValueError("averaging must be 'gradient' or 'parameter'")
        if exchange not in ("auto", "peer", "nccl"):
            raise ValueError("exchange must be 'auto', 'peer' or 'nccl'")
        self.exchange = exchange
        self._recv = None
        self._recv_tried = False
        self.dp_path = None
        self.average_every = int(average_every) if averaging == "parameter" else 0
        if not self.layers:
            raise ValueError("add layers before compile")
        if loss not in _LOSSES:
            raise ValueError("unknown loss %r" % (loss,))
        from .recurrent import LSTM
        if any(isinstance(l, LSTM) for l in self.layers):
            return self._compile_lstm(loss, optimizer)
        from .convolutional import Convolution1D
        if any(isinstance(l, Convolution1D) for l in self.layers):
            return self._compile_cnn(loss, optimizer)
        for l in self.layers[:-1]:
            if l.activation != "relu":
                raise NotImplementedError("hidden layers are ReLU on this path (cnn.py:111)")
        if self.layers[-1].activation != "linear":
            raise NotImplementedError("the output layer is linear on this path (cnn.py:114)")
        if loss not in _LOSSES:
            raise ValueError("unknown loss %r" % (loss,))
        self.loss = _LOSSES[loss]
        self.optimizer = optimizer if optimizer is not None else SGD()
        self.cfg = self.optimizer.cfg()
        _lib.load()  # fail here, loudly, if the CUDA extension is missing
        if not torch.cuda.is_available():
            raise RuntimeError("dlas needs a B200 GPU: there is no CPU path")
        self.device = torch.device(self.device if self.device is not None else "cuda:%d" % torch.cuda.current_device())
        dims = self.dims
        if not ops.mlp_mode_supported(dims, self.mode):
            raise _lib.DlasError(-5, "layer widths %s are not supported in mode %d" % (dims, self.mode))
        n = ops.mlp_n_params(dims)
        self.n_params = n
        if self.params is None:
            self.params = torch.from_numpy(self._glorot(dims)).to(self.device)
        self.m = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.v = torch.zeros(n, dtype=torch.float32, device=self.device)
        self.grad_sum = torch.zeros(n + 1, dtype=torch.float32, device=self.device)  # [n] = loss sum
        self.workspace = ops.mlp_workspace(dims, self.mode, self.device)
        self.iterations = 0
        self.rank, self.world = parallel.world_info()
        return self

    def _glorot(self, dims):
        """Keras Dense default: glorot_uniform kernel, zero bias (drawn from the counter hash)."""
        p = np.zeros(sum(a * b + b for a, b in zip(dims[:-1], dims[1:])), dtype=np.float32)
        o = 0
        for li, (a, b) in enumerate(zip(dims[:-1], dims[1:])):
            lim = np.sqrt(6.0 / (a + b))
            u = hashing.uniform(self.seed, 100 + li, a * b).astype(np.float64)
            p[o:o + a * b] = ((2.0 * u - 1.0) * lim).astype(np.float32)
            o += a * b + b
        return p

    def set_weights_flat(self, flat):
        t = torch.as_tensor(np.asarray(flat, dtype=np.float32))
        if self.params is None:
            self.params = t.to(self.device if self.device is not None else "cuda")
        else:
            self.params.copy_(t)

    def get_weights_flat(self):
        return self.params.detach().cpu().numpy()

    def save_weights(self, path):
        d = os.path.dirname(path)
        if d:
            os.makedirs(d, exist_ok=True)
        torch.save({"dims": self.dims, "params": self.params.detach().cpu()}, path)

    def load_weights(self, path):
        ck = torch.load(path)
        if list(ck["dims"]) != self.dims:
            raise ValueError("checkpoint dims %s != model dims %s" % (ck["dims"], self.dims))
        self.set_weights_flat(ck["params"].numpy())

    # -- helpers
    def _to_dev(self, a, cols):
        if isinstance(a, ops.LagWindows):       # lag windows over a resident series: read in place with a row stride
            if a.shape[1] != cols:
                raise ValueError("LagWindows of width %d given to a model with %d inputs" % (a.shape[1], cols))
            return a
        if isinstance(a, torch.Tensor):
            t = a.to(device=self.device, dtype=torch.float32)
        else:
            t = torch.as_tensor(np.asarray(a, dtype=np.float32), device=self.device)
        return t.reshape(-1, cols).contiguous()

    def _dp_recv(self):
        """The receive buffer of the push/poll exchange (world > 1, tensor-core mode), created and mapped on first use;
        None when this model trains through NCCL instead (fp32 mode, exchange="nccl", parameter averaging, or the
        ranks' memory cannot be mapped).  Every rank takes the same branch: the mapping is a collective."""
        if self._recv_tried:
            return self._recv
        self._recv_tried = True
        self.dp_path = "single" if self.world == 1 else "nccl"
        if self.world == 1 or self.exchange == "nccl" or getattr(self, "average_every", 0) > 0:
            return None
        n_recv = ops.mlp_dp_recv_bytes(self.dims, self.mode, self.world)
        if not n_recv:
            if self.exchange == "peer":
                raise _lib.DlasError(-5, "exchange='peer': dlas_mlp_dp_step_at does not support mode %d / widths %s" % (self.mode, self.dims))
            return None
        try:
            self._recv = parallel.PeerRecvBuffer(n_recv, self.device)
        except Exception as e:   # no symmetric-memory support between these GPUs: NCCL path (still no CPU path)
            if self.exchange == "peer":
                raise
            print("dlas: peer-memory receive buffer unavailable (%s); using the NCCL all-reduce" % (e,))
            return None
        self._ring = torch.zeros(4096, dtype=torch.float32, device=self.device)
        self.dp_path = "peer-push"
        return self._recv

    def _push_steps(self, X, Y, local_rows, global_rows, order=None, n_steps=1):
        """n_steps data-parallel steps through dlas_mlp_dp_step_at (host-driven flavour) on batches `order` of
        (X, Y); returns the ring indices holding their batch-mean losses."""
        recv = self._recv
        if isinstance(X, ops.LagWindows):
            raise NotImplementedError("LagWindows inputs train through exchange='nccl' at world > 1")
        n_batches = max(1, X.shape[0] // local_rows)
        if order is not None:
            n_steps = len(order)
        if n_steps > self._ring.numel():
            self._ring = torch.zeros(n_steps, dtype=torch.float32, device=self.device)
        it0 = self.iterations
        ops.mlp_dp_step_at(self.dims, self.loss, self.mode, self.cfg, self.params, self.m, self.v, X, Y, local_rows, None,
                           recv.bases, self.rank, self.world, 1.0 / float(global_rows), n_batches, self._ring, self.workspace,
                           n_steps=n_steps, cursor0=0, iteration0=it0, batch_order=order)
        self.iterations += n_steps
        return (torch.arange(it0, it0 + n_steps, device=self.device) % self._ring.numel())

    def train_on_batch(self, X, y, global_rows=None):
        """One optimiser step on this rank's shard (X, y) of one global batch; returns the device
        scalar holding the global batch-mean loss (the LOCAL shard's mean loss in parameter-averaging mode)."""
        dims = self.dims
        if self.world > 1 and self._dp_recv() is not None:
            rows = global_rows if global_rows is not None else X.shape[0] * self.world
            return self._ring[self._push_steps(X, y, X.shape[0], rows)]
        if getattr(self, "average_every", 0) > 0 and self.world > 1:
            # periodic PARAMETER averaging (optional mode): local step on the local mean gradient, no per-step collective;
            # every `average_every` steps parameters and optimiser state are replaced by their mean over ranks
            loss_out = torch.empty(1, dtype=torch.float32, device=self.device)
            ops.mlp_grad(dims, self.loss, self.mode, self.params, X, y, self.grad_sum, self.workspace)
            ops.opt_step(self.cfg, self.params, self.m, self.v, self.grad_sum, 1.0 / X.shape[0], self.iterations, loss_out)
            self.iterations += 1
            if self.iterations % self.average_every == 0:
                parallel.average_([self.params, self.m, self.v])
            return loss_out
        rows = global_rows if global_rows is not None else X.shape[0] * self.world
        loss_out = torch.empty(1, dtype=torch.float32, device=self.device)
        parallel.dp_step(
            lambda: ops.mlp_grad(dims, self.loss, self.mode, self.params, X, y, self.grad_sum, self.workspace),
            lambda scale: ops.opt_step(self.cfg, self.params, self.m, self.v, self.grad_sum, scale,
                                       self.iterations, loss_out),
            self.grad_sum, rows)
        self.iterations += 1
        return loss_out

    # -- fit (cnn.py:127-128)
    def fit(self, x=None, y=None, validation_data=None, verbose=2, callbacks=None, nb_epoch=None, epochs=None,
            batch_size=32, shuffle=True, log=print):
        if self.loss is None:
            raise RuntimeError("compile() the model before fit()")
        n_epochs = nb_epoch if nb_epoch is not None else (epochs if epochs is not None else 1)
        dims = self.dims
        X = self._to_dev(x, dims[0])
        Y = self._to_dev(y, dims[-1])
        if validation_data is not None:
            Xv, Yv = self._to_dev(validation_data[0], dims[0]), self._to_dev(validation_data[1], dims[-1])
        local_batch = batch_size // self.world
        if local_batch < 1 or local_batch * self.world != batch_size:
            raise ValueError("batch_size %d must be a positive multiple of the world size %d" % (batch_size, self.world))
        n_local = X.shape[0]
        n_full = n_local // local_batch
        rem = n_local - n_full * local_batch
        if self.world > 1:
            # every rank must run the same number of steps of the same sizes: a rank with one batch more would wait for
            # sums that never come (unequal shards are a caller error, not something to paper over)
            mine = torch.tensor([n_full, -n_full, rem, -rem], dtype=torch.int64, device=self.device)
            torch.distributed.all_reduce(mine, op=torch.distributed.ReduceOp.MAX)
            if int(mine[0]) != -int(mine[1]) or int(mine[2]) != -int(mine[3]):
                raise ValueError("fit: ranks hold different numbers of rows (this rank: %d full batches + %d rows); shard the "
                                 "rows with dlas.parallel.local_rows so that every rank owns the same share of every batch" % (n_full, rem))
            self._dp_recv()
        history = {"loss": [], "val_loss": []}
        callbacks = callbacks or []
        for epoch in range(n_epochs):
            t0 = time.time()
            order = hashing.batch_order(self.seed, epoch, n_full) if shuffle else np.arange(n_full, dtype=np.int64)
            losses = torch.empty(n_full + (1 if rem else 0), dtype=torch.float32, device=self.device)
            if self.world == 1:
                if n_full:
                    ops.mlp_fit_steps(dims, self.loss, self.mode, self.cfg, self.params, self.m, self.v, X, Y,
                                      local_batch, n_full, self.iterations, losses, self.grad_sum, self.workspace,
                                      batch_order=order)
                    self.iterations += n_full
                if rem:  # Keras also trains on the ragged last batch
                    Xr, Yr = X[n_full * local_batch:], Y[n_full * local_batch:]
                    ops.mlp_fit_steps(dims, self.loss, self.mode, self.cfg, self.params, self.m, self.v, Xr, Yr,
                                      rem, 1, self.iterations, losses[n_full:], self.grad_sum, self.workspace)
                    self.iterations += 1
            elif self._recv is not None:
                # the library's own data-parallel step (two launches, exchange inside the finishing kernel), one call per epoch
                if n_full:
                    losses[:n_full] = self._ring[self._push_steps(X, Y, local_batch, batch_size, order=order)]
                if rem:
                    losses[n_full:] = self._ring[self._push_steps(X[n_full * local_batch:], Y[n_full * local_batch:], rem, rem * self.world)]
            else:
                for s in range(n_full):
                    b = int(order[s])
                    lo = self.train_on_batch(X[b * local_batch:(b + 1) * local_batch], Y[b * local_batch:(b + 1) * local_batch],
                                             global_rows=batch_size)
                    losses[s:s + 1].copy_(lo)
                if rem:
                    lo = self.train_on_batch(X[n_full * local_batch:], Y[n_full * local_batch:], global_rows=rem * self.world)
                    losses[n_full:].copy_(lo)
            # Keras reports the sample-weighted mean of the batch losses
            w = torch.full_like(losses, float(local_batch))
            if rem:
                w[-1] = float(rem)
            logs = {"loss": float((losses * w).sum().item() / max(1, n_local))}
            ops.mlp_status(dims, self.mode, self.workspace)     # raises if a step failed on the device (timeout, fp16 range)
            if validation_data is not None:
                logs["val_loss"] = self.evaluate(Xv, Yv)
            history["loss"].append(logs["loss"])
            history["val_loss"].append(logs.get("val_loss"))
            if verbose and self.rank == 0:
                msg = "Epoch %d/%d - %.1fs - loss: %.6f" % (epoch + 1, n_epochs, time.time() - t0, logs["loss"])
                if "val_loss" in logs:
                    msg += " - val_loss: %.6f" % logs["val_loss"]
                log(msg)
            stop = False
            for cb in callbacks:
                stop = bool(cb.on_epoch_end(self, epoch, logs)) or stop
            if stop:
                break
        return history

    def graph_stepper(self, X, Y, local_batch, global_rows=None, ring_len=4096, warmup_steps=3, peer=None):
        """Capture ONE whole training step -- gradient kernels on this rank's next `local_batch` rows of the
        resident (X, Y), the all-reduce of the flat bucket, the optimiser update, the cursor advance -- in a
        CUDA graph and return a GraphStepper whose .run(n) replays it n times with no per-step host work.
        The batch cursor and the iteration counter live in device memory (dlas_step_state), so every replay
        trains on the next batch with the right learning-rate schedule.  `warmup_steps` real steps run first."""
        dims = self.dims
        n_batches = X.shape[0] // local_batch
        if n_batches < 1:
            raise ValueError("need at least one whole batch of %d rows" % local_batch)
        rows = global_rows if global_rows is not None else local_batch * self.world
        state = torch.tensor([0, self.iterations], dtype=torch.int64, device=self.device)
        ring = torch.zeros(ring_len, dtype=torch.float32, device=self.device)
        Y = Y.reshape(-1, dims[-1])

        # peer=True: the all-reduce and the update are ONE kernel of ours over NVLink peer memory
        # (dlas_dp_allreduce_update); peer=False: NCCL all-reduce between our kernels; None: peer when world > 1
        # and the ranks' memory can be mapped, else NCCL
        bucket, recv = None, None
        if self.world == 1 and peer and ops.mlp_dp_recv_bytes(dims, self.mode, 1):
            recv = parallel.LocalRecvBuffer(ops.mlp_dp_recv_bytes(dims, self.mode, 1), self.device)   # same two launches, no peers
        if self.world > 1 and peer is not False:
            try:
                n_recv = ops.mlp_dp_recv_bytes(dims, self.mode, self.world)
                if n_recv:      # tensor-core mode: forward/backward + ONE kernel for reduce, exchange, update and advance
                    recv = self._recv if self._recv is not None else parallel.PeerRecvBuffer(n_recv, self.device)
                    self._recv, self._recv_tried = recv, True
                    if not hasattr(self, "_ring"):
                        self._ring = torch.zeros(4096, dtype=torch.float32, device=self.device)
                else:
                    bucket = parallel.PeerBucket(self.n_params, self.device)
            except Exception as e:   # no symmetric-memory support between these GPUs: NCCL path (still no CPU path)
                if peer:
                    raise
                print("dlas: peer-memory bucket unavailable (%s); using the NCCL all-reduce" % (e,))
        self.dp_path = "peer-push" if recv is not None else "peer-kernel" if bucket is not None else ("nccl" if self.world > 1 else "single")

        if recv is not None:
            # host-driven: nothing to capture, the step is two launches; the host only counts batches
            it0 = self.iterations

            def enqueue(n):
                done = self.iterations - it0
                ops.mlp_dp_step_at(dims, self.loss, self.mode, self.cfg, self.params, self.m, self.v, X, Y, local_batch, None,
                                   recv.bases, self.rank, self.world, 1.0 / rows, n_batches, ring, self.workspace, n_steps=n,
                                   cursor0=done % n_batches, iteration0=self.iterations)
            enqueue(warmup_steps)
            self.iterations += warmup_steps
            st = GraphStepper(self, None, None, ring, enqueue)
            st.bucket = recv        # keeps the mapped memory alive as long as the stepper
            return st

        def one_step():
            if bucket is not None:
                ops.mlp_grad_at(dims, self.loss, self.mode, self.params, X, Y, local_batch, state, bucket.buf, self.workspace,
                                parity_stride=bucket.parity_stride)
                ops.dp_allreduce_update(self.cfg, self.params, self.m, self.v, bucket.bases, self.rank, self.world,
                                        bucket.parity_stride, bucket.flags_offset, 1.0 / rows, state, ring)
            else:
                ops.mlp_grad_at(dims, self.loss, self.mode, self.params, X, Y, local_batch, state, self.grad_sum, self.workspace)
                parallel.allreduce_sum_(self.grad_sum)
                ops.opt_step_at(self.cfg, self.params, self.m, self.v, self.grad_sum, 1.0 / rows, state, ring)
            ops.step_advance(state, n_batches)

        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup_steps):
                one_step()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            one_step()
        self.iterations += warmup_steps               # capture records the step, it does not run it
        st = GraphStepper(self, graph, state, ring)
        st.bucket = bucket      # keeps the mapped memory alive as long as the graph
        return st

    def fit_host(self, x_host, y_host, batch_size, n_steps, losses_host=None, first_batch=0):
        """Train `n_steps` steps streaming batches from (pinned) HOST tensors: the host->device copy
        of step s+1 runs on a copy stream under the kernels of step s (two device buffers); each
        step's batch-mean loss is read back into `losses_host` (pinned, n_steps floats).  With
        torch.distributed, x_host/y_host hold this rank's shard and batch_size is the GLOBAL batch.
        Returns losses_host; the caller synchronises."""
        dims = self.dims
        lb = batch_size // self.world
        n_batches = x_host.shape[0] // lb
        if losses_host is None:
            losses_host = torch.empty(n_steps, dtype=torch.float32).pin_memory()
        if not hasattr(self, "_hb") or self._hb[0][0].shape[0] != lb:
            self._hb = [(torch.empty((lb, dims[0]), dtype=torch.float32, device=self.device),
                         torch.empty((lb, dims[-1]), dtype=torch.float32, device=self.device)) for _ in range(2)]
            self._copy_stream = torch.cuda.Stream(device=self.device)
            self._ready = [torch.cuda.Event() for _ in range(2)]
            self._free = [torch.cuda.Event() for _ in range(2)]
            self._host_loss = torch.zeros(64, dtype=torch.float32, device=self.device)
        main = torch.cuda.current_stream()
        for ev in self._free:
            ev.record(main)
        yv = y_host.reshape(-1, dims[-1])
        for s in range(n_steps):
            b = (first_batch + s) % n_batches
            k = s & 1
            with torch.cuda.stream(self._copy_stream):
                self._copy_stream.wait_event(self._free[k])
                self._hb[k][0].copy_(x_host[b * lb:(b + 1) * lb], non_blocking=True)
                self._hb[k][1].copy_(yv[b * lb:(b + 1) * lb], non_blocking=True)
                self._ready[k].record(self._copy_stream)
            main.wait_event(self._ready[k])
            if self.world == 1 and getattr(self, "average_every", 0) == 0:
                # single GPU: the library's step loop on the staging buffer (gradient kernel + reduction fused with the update)
                lo = self._host_loss[s % 64:s % 64 + 1]
                ops.mlp_fit_steps(dims, self.loss, self.mode, self.cfg, self.params, self.m, self.v, self._hb[k][0], self._hb[k][1],
                                  lb, 1, self.iterations, lo, self.grad_sum, self.workspace)
                self.iterations += 1
            else:   # world > 1: dlas_mlp_dp_step_at on the staging buffer (push/poll exchange), or the NCCL step
                lo = self.train_on_batch(self._hb[k][0], self._hb[k][1], global_rows=batch_size)
            losses_host[s:s + 1].copy_(lo, non_blocking=True)
            self._free[k].record(main)
        return losses_host

    # -- inference (cnn.py:132)
    def predict(self, x):
        X = self._to_dev(x, self.dims[0])
        return ops.mlp_predict(self.dims, self.mode, self.params, X, workspace=self.workspace)

    def evaluate(self, x, y):
        """Mean loss over (x, y); with torch.distributed the mean is over all ranks' rows."""
        X = self._to_dev(x, self.dims[0])
        Y = self._to_dev(y, self.dims[-1])
        ls = ops.mlp_eval_loss(self.dims, self.loss, self.mode, self.params, X, Y, self.workspace)
        cnt = torch.tensor([float(X.shape[0])], dtype=torch.float32, device=self.device)
        if self.world > 1:
            pack = torch.cat([ls, cnt])
            torch.distributed.all_reduce(pack)
            ls, cnt = pack[:1], pack[1:]
        return float(ls.item() / max(1.0, cnt.item()))

    test_on_batch = evaluate
