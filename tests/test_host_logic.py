"""CPU tests of the host side: Keras-shaped façade argument handling, per-epoch batch order,
callbacks, sharding, and the data-parallel step over gloo with world_size 2 (the compute inside the
step comes from the oracle here -- on the GPU box it is libdlas_b200.so)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import dlas_oracle as O
from dlas import hashing, parallel
from dlas.models import Dense, EarlyStopping, ModelCheckpoint, Sequential, SGD, Adam, mae_clip


def test_batch_order_matches_oracle():
    for ep in (0, 1, 7):
        np.testing.assert_array_equal(hashing.batch_order(O.BASE_SEED, ep, 333), O.batch_order(O.BASE_SEED, ep, 333))


def test_glorot_matches_oracle():
    dims = [8, 64, 64, 1]
    m = Sequential([Dense(8, 64, activation="relu"), Dense(64, 64, activation="relu"), Dense(64, 1)], seed=O.BASE_SEED)
    np.testing.assert_allclose(m._glorot(dims), O.glorot_init(dims, O.BASE_SEED), rtol=1e-7)


def test_keras_call_shapes():
    m = Sequential()
    m.add(Dense(3600, 12))                       # cnn.py:114 positional (input_dim, output_dim)
    assert m.dims == [3600, 12]
    m2 = Sequential([Dense(64, input_dim=8, activation="relu"), Dense(1)])
    assert m2.dims == [8, 64, 1]
    with pytest.raises(ValueError):
        Sequential([Dense(8, 64), Dense(32, 1)])
    with pytest.raises(NotImplementedError):
        Dense(8, 8, activation="tanh")
    sgd = SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True)  # cnn.py:117
    c = sgd.cfg()
    assert (c.kind, c.nesterov) == (0, 1) and c.lr == pytest.approx(0.001) and c.momentum == pytest.approx(0.99)
    a = Adam().cfg()
    assert a.kind == 1 and a.eps == pytest.approx(1e-7)
    with pytest.raises(RuntimeError):
        mae_clip(None, None)


def test_early_stopping_and_checkpoint_match_oracle(tmp_path):
    v = [5, 4, 3] + [3.5] * 12
    es = EarlyStopping(monitor="val_loss", patience=10)            # cnn.py:121

    class M:
        rank = 0
        saved = []

        def save_weights(self, path):
            self.saved.append(path)

    mk = ModelCheckpoint(str(tmp_path / "models/cnn.mdl"), save_best_only=True)  # cnn.py:122
    model, stop_ep = M(), None
    for ep, val in enumerate(v):
        mk.on_epoch_end(model, ep, {"val_loss": val})
        if es.on_epoch_end(model, ep, {"val_loss": val}):
            stop_ep = ep
            break
    assert (stop_ep, len(model.saved)) == (O.early_stopping_trace(v, 10)[0], 3)


def test_shards_cover_the_global_batch():
    for world in (1, 2, 4, 8):
        rows = torch.cat([parallel.local_rows(64 * 5, 64, r, world) for r in range(world)])
        assert sorted(rows.tolist()) == list(range(320))
        lo, hi = parallel.shard_bounds(64, world - 1, world)
        assert hi == 64 and hi - lo == 64 // world
    with pytest.raises(ValueError):
        parallel.shard_bounds(10, 0, 4)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _dp_worker(rank, world, port, golden, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = np.load(golden)
    dims, batch, n_steps = g["dims"].tolist(), int(g["batch"]), int(g["n_steps"])
    X, y = g["X"].astype(np.float64), g["y"].astype(np.float64)
    mine = parallel.local_rows(X.shape[0], batch, rank, world).numpy()
    Xl, yl = X[mine], y[mine]
    lb = batch // world
    p = g["params0"].astype(np.float64)
    st, cfg = O.opt_state(p.size), O.sgd_config()
    bucket = torch.zeros(p.size + 1, dtype=torch.float64)
    losses = []
    for s in range(n_steps):
        def grad_fn():
            ls, gr = O.mlp_loss_grad_sums(p, dims, Xl[s * lb:(s + 1) * lb], yl[s * lb:(s + 1) * lb], O.LOSS_MSE)
            bucket[:-1] = torch.from_numpy(gr)
            bucket[-1] = ls

        def step_fn(scale):
            O.opt_step(p, bucket[:-1].numpy() * scale, cfg, st)
            losses.append(float(bucket[-1]) * scale)

        parallel.dp_step(grad_fn, step_fn, bucket, batch)
    if rank == 0:
        np.savez(out, losses=np.array(losses), final=p)
    dist.barrier()
    dist.destroy_process_group()


def test_dp_step_gloo_world2_equals_single_process(golden_dir, tmp_path):
    golden = os.path.join(golden_dir, "dp_equivalence.npz")
    out = str(tmp_path / "dp2.npz")
    mp.spawn(_dp_worker, args=(2, _free_port(), golden, out), nprocs=2, join=True)
    r, g = np.load(out), np.load(golden)
    np.testing.assert_allclose(r["losses"], g["losses_w1"], rtol=1e-12)
    np.testing.assert_allclose(r["final"], g["final_w1"], rtol=1e-9, atol=1e-13)


def _avg_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    a = torch.full((5,), float(rank + 1), dtype=torch.float64)
    b = torch.arange(3, dtype=torch.float64) * (rank + 1)
    parallel.average_([a, b])
    if rank == 0:
        np.savez(out, a=a.numpy(), b=b.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_parameter_average_gloo_world2(tmp_path):
    out = str(tmp_path / "avg.npz")
    mp.spawn(_avg_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = np.load(out)
    np.testing.assert_allclose(r["a"], 1.5)
    np.testing.assert_allclose(r["b"], np.arange(3) * 1.5)


def test_bind_host_to_gpu_is_a_no_op_without_nvml_or_on_one_node():
    """dlas.parallel.bind_host_to_gpu narrows the CPU affinity to the GPU's NUMA node before pinned buffers are allocated; with
    no NVML device (this container) or a single-node host it must change nothing and say so."""
    import os
    from dlas import parallel
    before = os.sched_getaffinity(0)
    assert parallel.bind_host_to_gpu(0) == 0
    assert os.sched_getaffinity(0) == before
