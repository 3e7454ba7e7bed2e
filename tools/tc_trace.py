"""Debug: one 65,536-row gradient call of the tcgen05 kernel with DLAS_TC_TRACE set; prints the SM-clock
phase breakdown of CTA 0 (epilogue thread 0 and the MMA issuer)."""
import os
import sys

PKG = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "deep-learning-at-scale_b200")
sys.path.insert(0, PKG)
os.environ["DLAS_B200_LIB"] = os.path.join(PKG, "lib", "libdlas_b200_trace.so")   # python deep-learning-at-scale_b200/build.py --trace
import torch  # noqa: E402
import dlas  # noqa: E402
from dlas import ops  # noqa: E402

DIMS = [32, 256, 256, 1]
X, y = ops.synth_rows(65536, 32)
p = torch.randn(ops.mlp_n_params(DIMS), device="cuda") * 0.05
ws = ops.mlp_workspace(DIMS, dlas.MODE_TC)
gs = torch.zeros(p.numel() + 1, device="cuda")
for _ in range(3):
    ops.mlp_grad(DIMS, 0, dlas.MODE_TC, p, X, y, gs, ws)
torch.cuda.synchronize()
os.environ["DLAS_TC_TRACE"] = "/tmp/tc_trace.txt"
ops.mlp_grad(DIMS, 0, dlas.MODE_TC, p, X, y, gs, ws)
torch.cuda.synchronize()
ev = [l.split() for l in open("/tmp/tc_trace.txt")]
t0 = min(int(t) for w, k, t in ev)      # common origin: both roles stamp the same SM clock
for who in ("E", "M", "F"):
    seq = [(int(k), int(t)) for w, k, t in ev if w == who]
    out = " ".join("%d:%d" % (k, t - t0) for k, t in seq)
    open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gpurun_out", "tc_trace_%s.txt" % who), "w").write(out + "\n")
    print(who, out[:1500])
