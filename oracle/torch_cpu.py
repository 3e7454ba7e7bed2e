"""Multi-threaded fp32 CPU port of the training step (torch CPU tensors, explicit forward /
backward / Keras-SGD -- no autograd).  TEST INFRASTRUCTURE: used only as bench.py's
`cpu_baseline` / `--impl reference` leg and cross-checked against dlas_oracle in tests.  It stands
in for the reference's CPU path: Spark local[*] + Keras/Theano are unavailable on this box
(BASELINE.md section 2), and the reference's own script cannot run.
Follows cnn.py:127-128 (batch loop), cnn.py:117 (SGD rule), cnn.py:29-32 (mae_clip)."""
import numpy as np
import torch


class TorchCpuMlp:
    def __init__(self, dims, params_flat, lr=0.001, momentum=0.99, decay=1e-6, nesterov=True, loss="mse"):
        self.dims = list(dims)
        self.W, self.b, self.vW, self.vb = [], [], [], []
        o = 0
        p = np.asarray(params_flat, dtype=np.float32)
        for a, c in zip(dims[:-1], dims[1:]):
            self.W.append(torch.from_numpy(p[o:o + a * c].reshape(a, c).copy())); o += a * c
            self.b.append(torch.from_numpy(p[o:o + c].copy())); o += c
            self.vW.append(torch.zeros(a, c)); self.vb.append(torch.zeros(c))
        self.lr, self.momentum, self.decay, self.nesterov, self.loss = lr, momentum, decay, nesterov, loss
        self.it = 0

    def flat(self):
        return np.concatenate([np.concatenate([W.numpy().ravel(), b.numpy()]) for W, b in zip(self.W, self.b)])

    def step(self, X, y):
        """One mini-batch step; returns the batch-mean loss."""
        B = X.shape[0]
        acts = [X]
        h = X
        L = len(self.W)
        for l in range(L):
            h = torch.addmm(self.b[l], h, self.W[l])
            if l < L - 1:
                h = torch.relu_(h)
            acts.append(h)
        d = acts[-1] - y
        n_out = self.dims[-1]
        if self.loss == "mse":
            loss = float((d * d).sum() / (B * n_out))
            dz = d * (2.0 / (B * n_out))
        else:
            a = d.abs()
            loss = float(a.clamp(max=6.0).sum() / (B * n_out))
            dz = torch.sign(d) * (a <= 6.0) / (B * n_out)
        lr_t = self.lr / (1.0 + self.decay * self.it)
        for l in range(L - 1, -1, -1):
            gW = acts[l].t() @ dz
            gb = dz.sum(dim=0)
            if l > 0:
                dz = (dz @ self.W[l].t()) * (acts[l] > 0)
            for p, v, g in ((self.W[l], self.vW[l], gW), (self.b[l], self.vb[l], gb)):
                v.mul_(self.momentum).add_(g, alpha=-lr_t)
                if self.nesterov:
                    p.add_(v, alpha=self.momentum).add_(g, alpha=-lr_t)
                else:
                    p.add_(v)
        self.it += 1
        return loss


def gilbert_numpy_f32(P, S, R, threads=1):
    """The physical baseline as the vectorised NumPy fp32 expression, optionally chunked over threads."""
    if threads <= 1:
        return P * np.power(S, np.float32(1.89)) / (np.float32(435.0) * np.power(R, np.float32(0.546)))
    from concurrent.futures import ThreadPoolExecutor
    out = np.empty_like(P)
    n = P.shape[0]
    cuts = np.linspace(0, n, threads + 1).astype(np.int64)

    def work(i):
        s = slice(cuts[i], cuts[i + 1])
        out[s] = P[s] * np.power(S[s], np.float32(1.89)) / (np.float32(435.0) * np.power(R[s], np.float32(0.546)))

    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, range(threads)))
    return out
