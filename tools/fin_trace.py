"""Debug (trace library: python deep-learning-at-scale_b200/build.py --trace): globaltimer stamps of the finishing kernel of
the data-parallel step (block 0 and the last block): reduction + push, poll for the peers' words, update.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/fin_trace.py"""
import ctypes as C
import os
import sys

PKG = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "deep-learning-at-scale_b200")
sys.path.insert(0, PKG)
os.environ["DLAS_B200_LIB"] = os.path.join(PKG, "lib", "libdlas_b200_trace.so")
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import dlas  # noqa: E402
from dlas import ops, _lib  # noqa: E402

rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
B = 65536
X, y = ops.synth_rows(B * 8, 32, start=rank * B * 8)
model = dlas.Sequential([dlas.Dense(32, 256, activation="relu"), dlas.Dense(256, 256, activation="relu"), dlas.Dense(256, 1)], mode="tc")
model.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
st = model.graph_stepper(X, y, B, global_rows=B * world, warmup_steps=5, peer=True)
st.run(59)
torch.cuda.synchronize()
buf = (C.c_ulonglong * (64 * 2 * 4))()
lib = _lib.load()
lib.dlas_debug_fin_trace.argtypes = [C.c_void_p]
assert lib.dlas_debug_fin_trace(buf) == 0
t = np.array(buf[:], dtype=np.int64).reshape(64, 2, 4)
its = [i for i in range(64) if t[i, 0, 0] > 0]
its.sort(key=lambda i: t[i, 0, 0])
its = its[-30:]
per = np.diff([t[i, 0, 0] for i in its]) / 1e3
out = ["rank %d: step period us: median %.1f min %.1f max %.1f" % (rank, np.median(per), per.min(), per.max())]
for b, name in ((0, "block 0"), (1, "last block")):
    d = np.array([[t[i, b, 1] - t[i, b, 0], t[i, b, 2] - t[i, b, 1], t[i, b, 3] - t[i, b, 2]] for i in its]) / 1e3
    out.append("rank %d %s: reduce+push %.1f us, poll %.1f us (max %.1f), update %.1f us; kernel start of last block after block 0: %.1f us"
               % (rank, name, np.median(d[:, 0]), np.median(d[:, 1]), d[:, 1].max(), np.median(d[:, 2]),
                  np.median([t[i, 1, 0] - t[i, 0, 0] for i in its]) / 1e3))
print("\n".join(out), flush=True)
if world > 1:
    dist.barrier()
    torch.cuda.synchronize()
    os._exit(0)
