"""A few training steps of the dynamic ANN 192-256-256-1 (24-lag windows over an 8-feature series, 65,536 windows per step) in
tensor-core mode, for the profiler:  ncu ... python tools/dyn_step.py [steps]"""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import __graft_entry__ as G  # noqa: E402

G.build()
import numpy as np  # noqa: E402
import torch  # noqa: E402

import dlas  # noqa: E402
from dlas import ops  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
mode = sys.argv[2] if len(sys.argv) > 2 else "tc"
dd = [192, 256, 256, 1]
series, ys = ops.synth_rows(65536 + 23, 8)
Wd, yd = dlas.LagWindows(series, 24), ys[23:].contiguous()
mdl = dlas.Sequential([dlas.Dense(192, 256, activation="relu"), dlas.Dense(256, 256, activation="relu"), dlas.Dense(256, 1)], mode=mode)
mdl.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
ls = torch.zeros(steps, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for rep in range(3):
    e0.record()
    ops.mlp_fit_steps(dd, mdl.loss, mdl.mode, mdl.cfg, mdl.params, mdl.m, mdl.v, Wd, yd, 65536, steps, 0, ls, mdl.grad_sum, mdl.workspace,
                      batch_order=np.zeros(steps, dtype=np.int64))
    e1.record()
    torch.cuda.synchronize()
    ops.mlp_status(dd, mdl.mode, mdl.workspace)
    print("%s: %.1f us / step, %.3f G samples/s, losses %s" % (mode, e0.elapsed_time(e1) / steps * 1e3, 65536 * steps / e0.elapsed_time(e1) / 1e6,
                                                              ls[:2].tolist()))
