"""Data-parallel parity on real GPUs (run under torchrun with WORLD_SIZE ranks, NCCL):
every rank trains on ITS shard of each global batch through the public facade; rank 0 compares the
per-step losses and final weights with the single-process fp64 oracle on the whole batches.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_check.py [fp32|tc]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "deep-learning-at-scale_b200"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import dlas  # noqa: E402
from dlas import parallel  # noqa: E402
import dlas_oracle as O  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "fp32"
# path: eager = train_on_batch per step (tc: the library's push/poll step; fp32: NCCL); nccl = train_on_batch with the NCCL
# all-reduce forced; fit = model.fit with the shuffled epoch order (the reference's call, cnn.py:127-128); graph = replayed
# step + NCCL; peer = stepper + our NVLink kernels; pavgK = parameter averaging every K steps
path = "eager"
if "-" in mode:
    mode, path = mode.split("-")
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
if mode == "lstm":
    # data-parallel LSTM (configs[4] shape, smaller batch): every rank trains on its slice of each global batch, the flat
    # bucket is all-reduced by NCCL between our kernels; compared with the single-process fp64 oracle
    lmode = path if path in ("fp32", "tc") else "fp32"
    d_in, H, L, T, batch, steps = 8, 128, 2, 12, 256, 4
    Xs, ys = O.synth_series(O.BASE_SEED, 0, batch * steps, T, d_in)
    p0 = O.lstm_init(d_in, H, L, O.BASE_SEED)
    mine = parallel.local_rows(Xs.shape[0], batch, rank, world).numpy()
    ml = dlas.Sequential([dlas.LSTM(H, input_shape=(T, d_in), return_sequences=True), dlas.LSTM(H), dlas.Dense(1)], mode=lmode)
    ml.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    ml.set_weights_flat(p0)
    Xl, yl = torch.from_numpy(Xs[mine]).cuda(), torch.from_numpy(ys[mine]).cuda().reshape(-1, 1)
    lb = batch // world
    got = torch.cat([ml.train_on_batch(Xl[s * lb:(s + 1) * lb], yl[s * lb:(s + 1) * lb], global_rows=batch) for s in range(steps)]).cpu().numpy()
    w = ml.get_weights_flat()
    chk = torch.tensor([float(np.abs(w).sum())], device="cuda", dtype=torch.float64)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    ok = True
    if rank == 0:
        pr = p0.astype(np.float64)
        ref = O.lstm_fit_steps(pr, d_in, H, L, Xs.astype(np.float64), ys.astype(np.float64), batch, steps, O.LOSS_MSE, O.sgd_config(),
                               O.opt_state(pr.size))
        lerr = float(np.abs(got / ref - 1).max())
        werr = float(np.linalg.norm(w - pr) / np.linalg.norm(pr))
        same = float(hi.item() - lo.item())
        tol_l, tol_w = (1e-4, 1e-4) if lmode == "fp32" else (1e-3, 1e-4)
        ok = lerr < tol_l and werr < tol_w and same == 0.0
        print("dp_check world=%d mode=lstm path=%s: max loss rel err %.3g, weights rel err %.3g, rank spread %.3g -> %s"
              % (world, lmode, lerr, werr, same, "OK" if ok else "FAIL"), flush=True)
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.broadcast(flag, 0)
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(int(flag.item()))
dims = [32, 256, 256, 1]
batch, steps = 8192, 12
X, y = O.synth_rows(O.BASE_SEED, 0, batch * steps, 32)
p0 = O.glorot_init(dims, O.BASE_SEED)
mine = parallel.local_rows(X.shape[0], batch, rank, world).numpy()
model = dlas.Sequential([dlas.Dense(32, 256, activation="relu"), dlas.Dense(256, 256, activation="relu"), dlas.Dense(256, 1)], mode=mode)
every = int(path[4:]) if path.startswith("pavg") else 0      # e.g. fp32-pavg3: parameter averaging every 3 steps
model.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True),
              averaging="parameter" if every else "gradient", average_every=max(1, every),
              exchange="nccl" if path in ("nccl", "graph") else "auto")
model.set_weights_flat(p0)
Xl, yl = torch.from_numpy(X[mine]).cuda(), torch.from_numpy(y[mine]).cuda().reshape(-1, 1)
lb = batch // world
order = None
if path == "fit":
    from dlas import hashing
    order = hashing.batch_order(model.seed, 0, steps)
    hist = model.fit(Xl, yl, nb_epoch=1, batch_size=batch, shuffle=True, verbose=0)
    got = model._ring[torch.arange(steps, device="cuda") % model._ring.numel()].cpu().numpy() if model._recv is not None else None
    assert model.dp_path == ("peer-push" if mode == "tc" else "nccl"), model.dp_path
    epoch_loss = hist["loss"][0]
elif path in ("eager", "nccl") or every:
    losses = []
    for s in range(steps):
        losses.append(model.train_on_batch(Xl[s * lb:(s + 1) * lb], yl[s * lb:(s + 1) * lb], global_rows=batch))
    got = torch.cat(losses).cpu().numpy()
    if not every:
        assert model.dp_path == ("peer-push" if (mode == "tc" and path == "eager") else "nccl"), model.dp_path
else:
    st = model.graph_stepper(Xl, yl, lb, global_rows=batch, warmup_steps=3, peer=(path == "peer"))
    st.run(steps - 3)
    torch.cuda.synchronize()
    got = st.recent_losses(steps)
    assert model.dp_path == (("peer-push" if mode == "tc" else "peer-kernel") if path == "peer" else "nccl"), model.dp_path
if every:
    parallel.average_([model.params])          # the oracle reports the replica mean after the last step
w = model.get_weights_flat()
# every rank must hold identical weights after the all-reduced steps
chk = torch.tensor([float(np.abs(w).sum())], device="cuda", dtype=torch.float64)
lo, hi = chk.clone(), chk.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN)
dist.all_reduce(hi, op=dist.ReduceOp.MAX)
ok = True
if rank == 0:
    pr = p0.astype(np.float64)
    if every:   # per-step loss here is the rank-0 shard's local mean; compare weights (after a final average) only
        ref = O.mlp_fit_steps_param_avg(pr, dims, X.astype(np.float64), y.astype(np.float64), batch, steps, O.LOSS_MSE,
                                        O.sgd_config(), world, every)
        got = ref
    else:
        ref = O.mlp_fit_steps(pr, dims, X.astype(np.float64), y.astype(np.float64), batch, steps, O.LOSS_MSE, O.sgd_config(), O.opt_state(pr.size),
                              order=order)
        if got is None:      # fit through NCCL keeps no per-step ring: compare the epoch mean Keras reports
            got, ref = np.asarray([epoch_loss]), np.asarray([ref.mean()])
    lerr = float(np.abs(got / ref - 1).max())
    werr = float(np.linalg.norm(w - pr) / np.linalg.norm(pr))
    same = float(hi.item() - lo.item())
    tol_l = 1e-4 if mode == "fp32" else 3e-4
    ok = lerr < tol_l and werr < 1e-4 and same == 0.0
    print("dp_check world=%d mode=%s path=%s: max loss rel err %.3g, weights rel err %.3g, rank spread %.3g -> %s"
          % (world, mode, path, lerr, werr, same, "OK" if ok else "FAIL"), flush=True)
flag = torch.tensor([0 if ok else 1], device="cuda")
dist.broadcast(flag, 0)
torch.cuda.synchronize()
sys.stdout.flush()
os._exit(int(flag.item()))     # captured collectives keep communicator resources alive: skip the collective teardown
