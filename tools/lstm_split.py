"""Debug: one LSTM 2x128, T=48 training step at batch 4096 (configs[4]) -- run under
`ncu --metrics gpu__time_duration.sum --clock-control none --csv` to see how the step splits over its kernels."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "deep-learning-at-scale_b200"))
import torch  # noqa: E402
import dlas  # noqa: E402
from dlas import ops  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "fp32"
Xs, ys = ops.synth_series(4096, 48, 8)
kw = {"mode": mode} if mode != "fp32" else {}
ml = dlas.Sequential([dlas.LSTM(128, input_shape=(48, 8), return_sequences=True), dlas.LSTM(128), dlas.Dense(1)], **kw)
ml.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
for _ in range(2):
    lo = ml.train_on_batch(Xs, ys)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    lo = ml.train_on_batch(Xs, ys)
e1.record()
torch.cuda.synchronize()
print("mode %s: %.3f ms per step, loss %.6f" % (mode, e0.elapsed_time(e1) / 3, float(lo.item())))
