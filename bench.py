#!/usr/bin/env python
"""Benchmark of the hot path (BASELINE.json metric): static-ANN 32-256-256-1 mini-batch training
samples/s on synthetic rows, data-parallel over N B200s, plus Gilbert rows/s against HBM.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

A "step" is one optimiser step (forward, loss, backward, gradient all-reduce when N>1, SGD update)
over one batch of 65,536 rows per GPU.  Rank 0 prints ONE JSON line.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "deep-learning-at-scale_b200"))

DIMS = [32, 256, 256, 1]
TRAIN_FLOP_PER_SAMPLE = 427_520       # 2*sum(in*out) forward + 2x that backward minus the first layer's dX
BYTES_PER_SAMPLE = 132                # 32 features + 1 target, fp32
GILBERT_BYTES_PER_ROW = 16            # three fp32 reads + one fp32 write
WORKLOAD = "static ANN 32-256-256-1, MSE, SGD(lr=1e-3, m=0.99, decay=1e-6, nesterov), 100M synthetic rows (configs[2])"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback (B200_PROFILING.md)")


class ClockSampler:
    """Samples SM clock and throttle reasons while the timed region runs (NVML).  ONE process samples, and only its own
    GPU: NVML calls from eight processes at once -- or from one process to eight GPUs -- were measured to stall the
    ranks' kernels (see main)."""

    def __init__(self, indices, period=0.004):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._th = None
        self.period = period
        self.nv, self.hs = None, []
        if not indices:
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.hs = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in indices]
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.hs[0], pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    NAMES = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
             "hw_power_brake": 0x80, "sync_boost": 0x10, "applications_clocks_setting": 0x2}

    def _run(self):
        while not self._stop.is_set():
            self.sample_once()
            self._stop.wait(self.period)

    def sample_once(self):
        nv = self.nv
        if not nv:
            return
        for h in self.hs:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in self.NAMES.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass

    def __enter__(self):
        if self.nv:
            self._th = threading.Thread(target=self._run, daemon=True)
            self._th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._th:
            self._th.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "gpus_sampled": len(self.hs)}


def n_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def _oracle_path():
    p = os.path.join(ROOT, "oracle")
    if p not in sys.path:
        sys.path.insert(0, p)


def cpu_port_rate(rows_per_step, n_steps, warmup=1, dims=None, n_blocks=4):
    """Times oracle/torch_cpu.py (the CPU port of the same training step) -- bench.py executes oracle/ only here, in
    gilbert_cpu_baseline (both: the cpu_baseline legs) and in dp_parity_gate (as the checker, outside every timed
    region).  Returns (samples/s, seconds, threads)."""
    _oracle_path()
    import numpy as np
    import torch
    import dlas_oracle as O
    import torch_cpu as T
    dims = dims or DIMS
    torch.set_num_threads(n_cores())
    n = rows_per_step * n_blocks
    rng = np.random.default_rng(0)   # values do not matter for timing; shapes and dtypes do
    X = torch.from_numpy((rng.random((n, dims[0]), dtype=np.float32) * 2 - 1))
    y = torch.from_numpy((rng.random((n, 1), dtype=np.float32) * 2 - 1))
    m = T.TorchCpuMlp(dims, O.glorot_init(dims, O.BASE_SEED))
    for s in range(warmup):
        m.step(X[:rows_per_step], y[:rows_per_step])
    t0 = time.perf_counter()
    for s in range(n_steps):
        b = s % n_blocks
        m.step(X[b * rows_per_step:(b + 1) * rows_per_step], y[b * rows_per_step:(b + 1) * rows_per_step])
    dt = time.perf_counter() - t0
    return rows_per_step * n_steps / dt, dt, torch.get_num_threads()


def gilbert_cpu_baseline(n_rows=20_000_000):
    """the physical model as the vectorised NumPy fp32 expression over a >= 20 M-row slice of the same
    synthetic inputs, one thread and all host threads (chunked)."""
    _oracle_path()
    import dlas_oracle as O
    import torch_cpu as T
    P, S, R = O.synth_gilbert(O.BASE_SEED, 0, n_rows)
    out = {}
    for name, thr in (("numpy_f32_1_thread", 1), ("numpy_f32_all_threads", n_cores())):
        T.gilbert_numpy_f32(P[:1 << 20], S[:1 << 20], R[:1 << 20], threads=thr)
        t0 = time.perf_counter()
        T.gilbert_numpy_f32(P, S, R, threads=thr)
        dt = time.perf_counter() - t0
        out[name] = {"value": n_rows / dt, "unit": "rows/s", "cores": thr, "kind": "port",
                     "sample": "%d rows of the config-2 inputs, q = P*S**1.89/(435*R**0.546) in NumPy fp32 (%.2f s)" % (n_rows, dt)}
    return out


def traffic_from_profiles(kernel, rows):
    """roofline.traffic: DRAM bytes per launch of `kernel` from the committed ncu capture (profiles/ncu_traffic.json, written
    by tools/ncu_summary.py), scaled by rows when the launch size differs.  (None, None) without a capture."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[kernel]
        b = (d["dram_bytes_read"] + d["dram_bytes_write"]) * rows / d["rows_per_launch"]
        return b, d["source"] + ("" if rows == d["rows_per_launch"] else " (scaled from %d rows per launch)" % d["rows_per_launch"])
    except Exception:
        return None, None


def dp_parity_gate(dlas, torch, dist, mode, rank, world, local_rows_fn):
    """BASELINE.md 5.5 "parity gates run before timing": 12 steps at global batch 8,192 of the benchmarked network through
    the SAME stepper the timed region uses (Sequential.graph_stepper), compared with the single-process fp64 oracle on the
    whole batches, plus the spread of a weight checksum over the ranks.  Returns the dp_parity object."""
    _oracle_path()
    import numpy as np
    import dlas_oracle as O
    batch, steps = 8192, 12
    X, y = O.synth_rows(O.BASE_SEED, 0, batch * steps, DIMS[0])
    p0 = O.glorot_init(DIMS, O.BASE_SEED)
    mine = local_rows_fn(X.shape[0], batch, rank, world).numpy()
    m = dlas.Sequential([dlas.Dense(32, 256, activation="relu"), dlas.Dense(256, 256, activation="relu"), dlas.Dense(256, 1)], mode=mode)
    m.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
    m.set_weights_flat(p0)
    Xl, yl = torch.from_numpy(X[mine]).cuda(), torch.from_numpy(y[mine]).cuda().reshape(-1, 1)
    st = m.graph_stepper(Xl, yl, batch // world, global_rows=batch, warmup_steps=3)
    st.run(steps - 3)
    torch.cuda.synchronize()
    got = st.recent_losses(steps)
    w = m.get_weights_flat()
    chk = torch.tensor([float(np.abs(w.astype(np.float64)).sum())], device="cuda", dtype=torch.float64)
    lo, hi = chk.clone(), chk.clone()
    if world > 1:
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    pr = p0.astype(np.float64)
    ref = O.mlp_fit_steps(pr, DIMS, X.astype(np.float64), y.astype(np.float64), batch, steps, O.LOSS_MSE, O.sgd_config(), O.opt_state(pr.size))
    lerr = float(np.abs(got / ref - 1).max())
    werr = float(np.linalg.norm(w - pr) / np.linalg.norm(pr))
    tol_l = 1e-4 if mode == "fp32" else 3e-4
    res = {"path": m.dp_path, "steps": steps, "global_batch": batch, "loss_rel_err": lerr, "weights_rel_err": werr,
           "rank_spread": float(hi.item() - lo.item()), "tolerance": {"loss": tol_l, "weights": 1e-4},
           "checked_against": "fp64 oracle on the whole batches (oracle/dlas_oracle.py), before the timed region"}
    res["ok"] = bool(lerr < tol_l and werr < 1e-4 and res["rank_spread"] == 0.0 and np.isfinite(got).all())
    del st
    return res


def reference_arm(args, rank):
    """--impl reference: the reference's CPU path.  Spark local[*] + Keras/Theano cannot run on this
    box and the reference script is not executable, so this times the CPU
    port of the same step (oracle/torch_cpu.py) with all host threads, on rank 0 only."""
    if rank != 0:
        return
    rate1, dt1, _ = cpu_port_rate(8192, 2, warmup=1)
    budget = 150.0 / max(1, args.steps + args.warmup)          # seconds per step so the run stays within minutes
    rows = int(min(65536, max(4096, (rate1 * min(budget, 1.0)) // 4096 * 4096)))
    rate, dt, thr = cpu_port_rate(rows, args.steps, warmup=max(1, args.warmup))
    line = {"impl": "reference", "metric": "static_ann_train_samples_per_s", "value": rate, "unit": "samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "rows_per_step": rows, "host": "cpu"},
            "cpu_baseline": {"value": rate, "unit": "samples/s", "cores": thr, "kind": "port",
                             "sample": "%d steps of %d rows, torch-CPU fp32 port of the step (Spark/Keras unavailable)" % (args.steps, rows)},
            "e2e": {"value": rate, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


_REAL_STDOUT = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: library banners written to fd 1 (NCCL prints its version there) are
    sent to stderr for the rest of the run; emit() writes the line to the real stdout."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--batch", type=int, default=65536, help="rows per GPU per step")
    ap.add_argument("--mode", default="auto", choices=["auto", "fp32", "tc"])
    ap.add_argument("--gilbert-rows", type=int, default=200_000_000)
    ap.add_argument("--gilbert-variant", type=int, default=-1)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the dynamic-ANN / 8-64-64-1 / LSTM side measurements")
    ap.add_argument("--no-sustained", action="store_true", help="skip the >= 2 s pass")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return reference_arm(args, rank)
    if args.warmup < 3:
        args.warmup = 3

    import numpy as np
    import torch
    import torch.distributed as dist
    import dlas
    from dlas import ops

    torch.cuda.set_device(local_rank)
    host_cpus = 0
    if world > 1:
        from dlas import parallel as _par
        host_cpus = _par.bind_host_to_gpu(local_rank)     # pinned batch buffers on the GPU's own NUMA node (e2e)
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    N, K, W, B = world, args.steps, args.warmup, args.batch
    pk = peaks()

    mode = args.mode
    if mode == "auto":
        mode = "tc" if ops.mlp_mode_supported(DIMS, dlas.MODE_TC) else "fp32"
    model = dlas.Sequential([dlas.Dense(32, 256, activation="relu"), dlas.Dense(256, 256, activation="relu"),
                             dlas.Dense(256, 1)], mode=mode)
    model.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))

    # ---- this rank's shard of the rows, generated in HBM
    rows_rank = (args.rows // N) // B * B
    X, y = ops.synth_rows(rows_rank, DIMS[0], start=rank * rows_rank)
    n_batches = rows_rank // B
    losses = torch.zeros(max(K, W), dtype=torch.float32, device="cuda")

    stepper = None
    use_graph = N > 1 or os.environ.get("DLAS_BENCH_GRAPH") == "1"     # N == 1 default: eager launches from the host loop
    if use_graph and N == 1:
        stepper = model.graph_stepper(X, y, B, global_rows=B, warmup_steps=W, peer=True)
    if N > 1:
        # one captured step (gradient kernels, all-reduce, update, cursor advance) replayed K times: no per-step host work
        stepper = model.graph_stepper(X, y, B, global_rows=B * N, warmup_steps=W)

    def run_steps(n, first, keep=True):
        if stepper is None:
            done = 0
            while done < n:     # sequential batches, wrapping at the end of the shard
                fb = (first + done) % n_batches
                cnt = min(n - done, n_batches - fb)
                ops.mlp_fit_steps(DIMS, model.loss, model.mode, model.cfg, model.params, model.m, model.v, X, y, B, cnt,
                                  model.iterations, losses[done:] if keep else None, model.grad_sum, model.workspace, first_batch=fb)
                model.iterations += cnt
                done += cnt
        else:
            stepper.run(n)

    def barrier():
        torch.cuda.synchronize()
        if N > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity gate of the N > 1 path, before any timing; a failed gate fails the run
    dp_parity = None
    if N > 1 or os.environ.get("DLAS_BENCH_GATE") == "1":
        from dlas import parallel
        dp_parity = dp_parity_gate(dlas, torch, dist, mode, rank, N, parallel.local_rows)
        if not dp_parity["ok"]:
            if rank == 0:
                sys.stderr.write("bench.py: data-parallel parity gate FAILED: %s\n" % json.dumps(dp_parity))
                emit({"metric": "static_ann_train_samples_per_s", "value": None, "n_gpus": N, "invalid": "dp_parity gate failed",
                      "dp_parity": dp_parity})
            torch.cuda.synchronize()
            os._exit(3)

    run_steps(W, 0)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # Clocks.  N = 1: NVML sampled every 4 ms while the timed region runs (cost measured: 0.2 us per step).  N > 1: NVML in
    # rank 0's process slows the ranks' lock-step whenever it is used around the region (measured at N=8, us per step:
    # NVML never initialised 71; one sample before + one after the region 115; rank 0 sampling its own GPU inside the
    # region 93; rank 0 sampling all eight 94; a sampler in every rank 99).  So at N > 1 the timed pass runs with NVML
    # untouched and reports the SM clock the forward/backward kernel measures itself INSIDE the region (clock64 /
    # globaltimer, dlas_prof_sm_clock); NVML clock and throttle reasons are then sampled every 4 ms during a second,
    # untimed pass of the same K steps right after it.
    no_nvml = os.environ.get("DLAS_BENCH_NO_SAMPLER") == "1"
    in_region = (N == 1) and not no_nvml
    clk = ClockSampler([local_rank] if (in_region and rank == 0) else [], period=0.004)
    if mode == "tc":
        ops.prof_sm_clock(reset=True)
    with clk:
        e0.record()
        run_steps(K, W)
        e1.record()
        barrier()
    device_mhz = ops.prof_sm_clock(reset=True) if mode == "tc" else None
    if N > 1 and not no_nvml:
        clk = ClockSampler([local_rank] if rank == 0 else [], period=0.004)
        with clk:
            run_steps(K, W + K)
            barrier()
    ms = e0.elapsed_time(e1)
    final_loss = float(losses[K - 1].item()) if stepper is None else float(stepper.recent_losses(1)[0])
    steps_done = W + K + (K if (N > 1 and not no_nvml) else 0)
    # ---- sustained pass: the same steps for >= 2 s, so that the fraction can also be quoted against the SUSTAINED peak
    # (the K-step region above is a burst of about a millisecond); clocks as above
    sustained = None
    if not args.no_sustained:
        ks = int(min(200_000, max(K, 2.2 / max(1e-6, ms / K * 1e-3))))
        if N > 1:
            tks = torch.tensor([ks], device="cuda")
            dist.all_reduce(tks, op=dist.ReduceOp.MAX)
            ks = int(tks.item())
        clk_s = ClockSampler([local_rank] if (in_region and rank == 0) else [], period=0.02)
        if mode == "tc":
            ops.prof_sm_clock(reset=True)
        barrier()
        with clk_s:
            e0.record()
            run_steps(ks, steps_done, keep=False)
            e1.record()
            barrier()
        steps_done += ks
        ts = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if N > 1:
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
        s_ms = float(ts.item())
        sustained = {"steps": ks, "seconds": s_ms * 1e-3, "value": N * B * ks / (s_ms * 1e-3), "unit": "samples/s",
                     "us_per_step": s_ms / ks * 1e3, "sm_mhz_in_kernel": ops.prof_sm_clock(reset=True) if mode == "tc" else None,
                     "clocks": clk_s.summary() if in_region else "NVML not sampled at N > 1 (see clocks.when)"}
    # Kernel time for the roofline: a SECOND pass over the same kind of steps with CUDA events around every launch of
    # the forward/backward kernel on its own stream (dlas_prof_*).  Kept out of the pass above because an event between
    # two kernels breaks their programmatic-dependent-launch overlap and would tax `value`; kernels inside a replayed
    # graph (N>1) cannot carry events at all, so there the same kernel is launched eagerly on the same shapes.
    ops.prof_read(ops.PROF_MLP_GRAD)
    ops.prof_enable(True)
    if stepper is None:
        run_steps(K, steps_done)
    else:
        for s_ in range(10):
            b_ = s_ % n_batches
            ops.mlp_grad(DIMS, model.loss, model.mode, model.params, X[b_ * B:(b_ + 1) * B], y[b_ * B:(b_ + 1) * B],
                         model.grad_sum, model.workspace)
    barrier()
    ops.prof_enable(False)
    kern_ms, kern_n = ops.prof_read(ops.PROF_MLP_GRAD)
    t = torch.tensor([ms], device="cuda")
    if N > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = N * B * K / (ms * 1e-3)

    # ---- end to end through the public API with HOST buffers (H2D per step, loss D2H per step)
    ring = min(32, n_batches)
    xh = torch.empty((ring * B, DIMS[0]), dtype=torch.float32).pin_memory()
    yh = torch.empty((ring * B, 1), dtype=torch.float32).pin_memory()
    xh.copy_(X[:ring * B]); yh.copy_(y[:ring * B])
    lh = torch.empty(max(K, W), dtype=torch.float32).pin_memory()
    model.fit_host(xh, yh, B * N, W, lh)
    barrier()
    e0.record()
    model.fit_host(xh, yh, B * N, K, lh)
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if N > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = N * B * K / (float(t.item()) * 1e-3)
    dp_path_e2e = "single" if N == 1 else getattr(model, "dp_path", None)
    del xh, yh

    # ---- inference (a8: model.predict / test_on_batch, cnn.py:132) over resident rows: rows/s and its HBM fraction
    # (algorithmic bytes = 32 fp32 features read + 1 fp32 prediction written per row)
    n_pred = int(min(rows_rank, 8_388_608))
    yhat = torch.empty((n_pred, 1), dtype=torch.float32, device="cuda")
    for _ in range(2):
        ops.mlp_predict(DIMS, model.mode, model.params, X[:n_pred], out=yhat)
    barrier()
    e0.record()
    for r_ in range(5):
        ops.mlp_predict(DIMS, model.mode, model.params, X[:n_pred], out=yhat)     # 1.07 GB per pass >> L2
    e1.record()
    barrier()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if N > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    pred_s = float(t.item()) * 1e-3 / 5
    predict = {"rows_per_s": N * n_pred / pred_s, "rows_per_gpu_per_call": n_pred, "ms_per_call": pred_s * 1e3,
               "kernel": "forward-only kind of the tcgen05 kernel (mlp_tc_kernel<K_PREDICT>)" if mode == "tc" else "fp32 CUDA-core forward",
               "hbm_frac": n_pred * (DIMS[0] * 4 + 4) / pred_s / 1e9 / pk["hbm"],
               "tensor_frac_of_bf16_burst": n_pred * 147_968 / pred_s / 1e12 / pk["tf_burst"]}
    del yhat

    # ---- Gilbert rows/s vs HBM (config 2), sharded by rows, no collective
    del X, y
    torch.cuda.empty_cache()
    g_rows = args.gilbert_rows // N
    P, S, R = ops.synth_gilbert(g_rows, start=rank * g_rows)
    q = torch.empty_like(P)
    variants = [args.gilbert_variant] if args.gilbert_variant >= 0 else [0, 7]
    gil = {}
    for var in variants:
        for _ in range(3):
            ops.gilbert(P, S, R, out=q, variant=var)
        barrier()
        ops.prof_read(ops.PROF_GILBERT)
        ops.prof_enable(True)
        reps = 10
        e0.record()
        for _ in range(reps):
            ops.gilbert(P, S, R, out=q, variant=var)   # 3.2 GB per pass >> 126 MB L2
        e1.record()
        barrier()
        ops.prof_enable(False)
        gk_ms, gk_n = ops.prof_read(ops.PROF_GILBERT)
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if N > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        rows_s = N * g_rows * reps / (float(t.item()) * 1e-3)
        gbs = g_rows * GILBERT_BYTES_PER_ROW / (gk_ms / gk_n * 1e-3) / 1e9
        g_traffic, g_src = traffic_from_profiles("gilbert_vec_kernel", g_rows)
        gil[str(var)] = {"rows_per_s": rows_s, "kernel_ms": gk_ms / gk_n,
                         "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                                      "frac": gbs / pk["hbm"], "traffic": g_traffic, "traffic_source": g_src}}
    best = max(gil, key=lambda k: gil[k]["rows_per_s"])
    del P, S, R, q

    # ---- the other model families of SURVEY.md 8d, N=1 only, a few steps each (reported beside the headline, not the metric)
    extras = {}
    if N == 1 and not args.no_extras:
        def timed(fn, n):
            for _ in range(2):
                fn()
            torch.cuda.synchronize()
            e0.record()
            for _ in range(n):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) * 1e-3 / n
        # the headline network at other batch sizes: what each launch's fixed cost (weights -> fp16 tiles, accumulator drain,
        # partial reduction + update) weighs against the per-tile work
        if mode == "tc":
            sweep = {}
            ns = 20
            Xw, yw = ops.synth_rows(1048576 * ns, DIMS[0])        # 2.8 GB: every step of a sweep point reads new rows
            ls = torch.zeros(ns, device="cuda")
            for rows_b in (16384, 262144, 1048576):
                t = timed(lambda: ops.mlp_fit_steps(DIMS, model.loss, model.mode, model.cfg, model.params, model.m, model.v, Xw, yw,
                                                    rows_b, ns, 0, ls, model.grad_sum, model.workspace), 2)
                sweep[str(rows_b)] = {"samples_per_s": rows_b * ns / t, "us_per_step": t / ns * 1e6}
            extras["static_ann_32_256_256_1_rows_per_step_sweep"] = sweep
            del Xw, yw
        # dynamic ANN (config 4): 24 lags x 8 features -> 192-256-256-1, 65,536 windows per step; tensor-core mode = layer 1 in
        # mlp_tc_l1.cu around the CTA-pair kernel (five launches per step), fp32 mode = the CUDA-core kernels
        dd = [192, 256, 256, 1]
        series, ys = ops.synth_rows(65536 + 23, 8)
        Wd = dlas.LagWindows(series, 24)       # windows read in place from the raw series (row stride 8): 36 B per sample
        yd = ys[23:].contiguous()
        for dm in ("fp32", "tc"):
            if dm == "tc" and not (mode == "tc" and ops.mlp_mode_supported(dd, dlas.MODE_TC)):
                continue
            mdl = dlas.Sequential([dlas.Dense(192, 256, activation="relu"), dlas.Dense(256, 256, activation="relu"), dlas.Dense(256, 1)], mode=dm)
            mdl.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
            t = timed(lambda: mdl.train_on_batch(Wd, yd), 5)
            ls4 = torch.zeros(4, device="cuda")
            tf = timed(lambda: ops.mlp_fit_steps(dd, mdl.loss, mdl.mode, mdl.cfg, mdl.params, mdl.m, mdl.v, Wd, yd, 65536, 4, 0, ls4,
                                                 mdl.grad_sum, mdl.workspace, batch_order=np.zeros(4, dtype=np.int64)), 3) / 4
            ops.mlp_status(dd, mdl.mode, mdl.workspace)
            tp = timed(lambda: mdl.predict(Wd), 5)
            extras["dynamic_ann_192_256_256_1" + ("" if dm == "tc" or mode != "tc" else "_fp32")] = {
                "samples_per_s": 65536 / tf, "us_per_step": tf * 1e6, "train_on_batch_samples_per_s": 65536 / t, "mode": dm,
                "predict_rows_per_s": 65536 / tp,
                "rows_per_step": 65536,
                "input": "24-lag windows gathered from the raw (n, 8) series inside the kernels "
                         "(dlas_mlp_fit_steps_strided): 36 algorithmic bytes per sample, no materialised copy",
                "frac_of_bf16_sustained_peak": 591360 * 65536 / tf / 1e12 / pk["tf_sust"]}
            del mdl
        # static ANN 8-64-64-1 (configs[0]): 1 M rows, one whole epoch through model.fit at batch 4096 (and one 65,536-row
        # step for comparison with round 1); the CPU port runs the same epoch
        X8, y8 = ops.synth_rows(1_000_000, 8)
        m8 = dlas.Sequential([dlas.Dense(8, 64, activation="relu"), dlas.Dense(64, 64, activation="relu"), dlas.Dense(64, 1)],
                             mode=("tc" if ops.mlp_mode_supported([8, 64, 64, 1], dlas.MODE_TC) and mode == "tc" else "fp32"))
        m8.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
        t = timed(lambda: m8.train_on_batch(X8[:65536], y8[:65536]), 10)
        m8.fit(X8, y8, nb_epoch=1, batch_size=4096, verbose=0)
        torch.cuda.synchronize()
        t0_ = time.perf_counter()
        m8.fit(X8, y8, nb_epoch=1, batch_size=4096, verbose=0)
        torch.cuda.synchronize()
        t_epoch = time.perf_counter() - t0_
        e8 = {"samples_per_s": 65536 / t, "mode": "fp32" if m8.mode == dlas.MODE_FP32 else "tc", "rows_per_step": 65536,
              "epoch_1M_rows_batch_4096": {"samples_per_s": 1_000_000 / t_epoch, "seconds": t_epoch, "through": "model.fit (wall clock)"}}
        if not args.no_cpu_baseline:
            rate8, dt8, thr8 = cpu_port_rate(4096, 1_000_000 // 4096, warmup=2, dims=[8, 64, 64, 1], n_blocks=1_000_000 // 4096)
            e8["cpu_baseline"] = {"value": rate8, "unit": "samples/s", "cores": thr8, "kind": "port",
                                  "sample": "one epoch over 1 M rows at batch 4096 (%.2f s), torch-CPU fp32 port" % dt8}
        extras["static_ann_8_64_64_1"] = e8
        # the headline network at the reference's own batch size of 20 (cnn.py:128), for the record: launch-bound
        Xr, yr = ops.synth_rows(20 * 2000, DIMS[0])
        l20 = torch.zeros(2000, device="cuda")
        for md_, nm_ in ((dlas.MODE_FP32, "fp32"),) + (((dlas.MODE_TC, "tc"),) if mode == "tc" else ()):
            ws_ = ops.mlp_workspace(DIMS, md_)
            mm_ = torch.zeros_like(model.params)
            pp_ = model.params.clone()
            t = timed(lambda: ops.mlp_fit_steps(DIMS, model.loss, md_, model.cfg, pp_, mm_, mm_.clone(), Xr, yr, 20, 2000, 0, l20,
                                                model.grad_sum, ws_), 2)
            extras.setdefault("static_ann_32_256_256_1_batch_20", {})[nm_] = {"samples_per_s": 20 * 2000 / t, "us_per_step": t / 2000 * 1e6}
        del Xr, yr
        # LSTM 2x128, T=48, batch 4096 (config 5)
        Xs, ysq = ops.synth_series(4096, 48, 8)
        for lm in ("fp32", "tc"):
            if lm == "tc" and not ops.lstm_mode_supported(8, 128, 2, dlas.MODE_TC):
                continue
            ml = dlas.Sequential([dlas.LSTM(128, input_shape=(48, 8), return_sequences=True), dlas.LSTM(128), dlas.Dense(1)], mode=lm)
            ml.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
            t = timed(lambda: ml.train_on_batch(Xs, ysq), 5)
            tp = timed(lambda: ml.predict(Xs), 5)
            extras["lstm_2x128_T48_batch4096" + ("" if lm == "tc" else "_fp32")] = {
                "sequences_per_s": 4096 / t, "ms_per_step": t * 1e3, "mode": lm,
                "frac_of_bf16_sustained_peak": 57.41e6 * 4096 / t / 1e12 / pk["tf_sust"],
                "predict_sequences_per_s": 4096 / tp}
            del ml
        # the script's own CNN (cnn.py:110-128) at its own batch of 20 and at 4096 rows per step, dropout on, through the
        # epoch loop of the library (dlas_cnn_fit_steps: 6 launches per step, no Python between steps)
        for cb in (20, 4096):
            nb_ = 32
            Wc, _ = ops.synth_series(cb * nb_, 60, 1)
            Wc = Wc.reshape(cb * nb_, 60)
            Xc, yc = Wc[:, :48].contiguous(), Wc[:, 48:].contiguous()
            mc = dlas.Sequential([dlas.Convolution1D(1, 100, 13, activation="relu"), dlas.Dropout(0.5), dlas.Flatten(), dlas.Dense(3600, 12)])
            mc.compile(loss=dlas.mae_clip, optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
            impl = mc._impl
            t = timed(lambda: impl._fit_steps(Xc, yc, cb, list(range(nb_))), 3) / nb_
            extras["cnn_48_conv13x100_dense12_batch%d" % cb] = {"samples_per_s": cb / t, "us_per_step": t * 1e6, "mode": "fp32",
                                                                 "launches_per_step": 6}
        del series, ys, Wd, yd, X8, y8, Xs, ysq, Wc, Xc, yc

    if N > 1 and not args.no_extras:
        def timed_dp(fn, n):
            for _ in range(2):
                fn()
            barrier()
            e0.record()
            for _ in range(n):
                fn()
            e1.record()
            barrier()
            tt = torch.tensor([e0.elapsed_time(e1)], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt.item()) * 1e-3 / n
        # configs[4]: LSTM 2x128, T = 48, GLOBAL batch 4096 split over the ranks (strong scaling, as BASELINE.json states it)
        lb = 4096 // N
        Xs, ysq = ops.synth_series(lb, 48, 8, start=rank * lb)
        lm = "tc" if ops.lstm_mode_supported(8, 128, 2, dlas.MODE_TC) else "fp32"
        ml = dlas.Sequential([dlas.LSTM(128, input_shape=(48, 8), return_sequences=True), dlas.LSTM(128), dlas.Dense(1)], mode=lm)
        ml.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
        t = timed_dp(lambda: ml.train_on_batch(Xs, ysq, global_rows=4096), 5)
        extras["lstm_2x128_T48_batch4096"] = {"sequences_per_s": 4096 / t, "ms_per_step": t * 1e3, "mode": lm, "scaling": "strong",
                                              "sequences_per_gpu_per_step": lb, "exchange": "nccl all-reduce of the 807 KB bucket"}
        del ml, Xs, ysq
        # configs[3]: dynamic ANN 192-256-256-1, 65,536 windows per GPU per step read in place from the raw series (weak scaling)
        series, ys = ops.synth_rows(65536 + 23, 8, start=rank * 65536)
        Wd = dlas.LagWindows(series, 24)
        dm = "tc" if (mode == "tc" and ops.mlp_mode_supported([192, 256, 256, 1], dlas.MODE_TC)) else "fp32"
        mdl = dlas.Sequential([dlas.Dense(192, 256, activation="relu"), dlas.Dense(256, 256, activation="relu"), dlas.Dense(256, 1)], mode=dm)
        mdl.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True), exchange="nccl")
        yd = ys[23:].contiguous()
        t = timed_dp(lambda: mdl.train_on_batch(Wd, yd, global_rows=65536 * N), 5)
        extras["dynamic_ann_192_256_256_1"] = {"samples_per_s": 65536 * N / t, "mode": dm, "rows_per_gpu_per_step": 65536,
                                               "scaling": "weak", "exchange": "nccl all-reduce of the 462 KB bucket"}
        del mdl, series, ys, Wd, yd

    if rank == 0:
        tc_traffic, tc_src = traffic_from_profiles("mlp_tc_kernel", B) if mode == "tc" else (None, None)
        kern_s = kern_ms / max(1, kern_n) * 1e-3
        tfs = TRAIN_FLOP_PER_SAMPLE * B / kern_s / 1e12
        # my kernels per step: fp32 mode transpose + forward/backward + reduce + update; tc mode forward/backward +
        # (reduce fused with the update) at N=1, forward/backward + reduce + update at N>1 (NCCL's kernel not counted)
        # tc mode: forward/backward + one finishing kernel (reduce [+ NVLink push/poll exchange] + update [+ cursor advance]) at any N
        launches_per_step = (4 if mode == "fp32" else 2) if N == 1 else (5 if mode == "fp32" else 2)
        line = {
            "metric": "static_ann_train_samples_per_s", "value": value, "unit": "samples/s", "n_gpus": N, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if mode == "fp32" else
                     "f16-operands/f32-accumulate (reduced-range mode: |value| <= 65504, 11-bit significand; stated tolerance vs the "
                     "fp64 oracle: per-step loss 3e-4, weights 1e-4 -- tests/test_mlp_tc_gpu.py)",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "rows_per_gpu_per_step": B, "global_batch": B * N, "mode": mode,
                       "rows_resident_per_gpu": rows_rank, "l2": "inputs (%.1f GB per GPU) larger than L2; every step reads new rows"
                       % (rows_rank * BYTES_PER_SAMPLE / 1e9), "parallelism": "dp%d" % N, "final_loss": final_loss,
                       "dp_path": getattr(model, "dp_path", None) or ("single" if N == 1 else None),
                       "dp_path_e2e": dp_path_e2e, "host_cpus_bound_to_gpu_numa_node": host_cpus},
            # the K timed steps are a burst of about a millisecond: quoted against the BURST peak; `sustained` below carries
            # the >= 2 s figure against the sustained peak
            "roofline": {"bound": "tensor", "achieved": tfs, "peak": pk["tf_burst"], "unit": "TFLOP/s", "frac": tfs / pk["tf_burst"],
                         "frac_of_sustained_peak": tfs / pk["tf_sust"],
                         "traffic": tc_traffic, "traffic_source": tc_src,
                         "kernel": "mlp forward+loss+backward", "kernel_ms": kern_s * 1e3,
                         "kernel_share_of_step": (kern_ms / max(1, kern_n)) / (ms / K),
                         "kernel_timing": ("second pass of %d steps with CUDA events around every launch" % K) if N == 1
                         else "10 eager launches on the same shapes (graph replays cannot carry events)",
                         "peak_source": pk["src"] + ", bf16 burst",
                         "hbm_frac": BYTES_PER_SAMPLE * B / kern_s / 1e9 / pk["hbm"]},
            "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": B * BYTES_PER_SAMPLE, "d2h_bytes_per_step": 4},
            "gpu_launches": launches_per_step * K,
            "clocks": dict(clk.summary(), sm_mhz_in_kernel=device_mhz,
                           when=("NVML every 4 ms inside the timed region" if in_region else
                                 "sm_mhz_in_kernel: measured by the forward/backward kernel inside the timed region; NVML "
                                 "clock and reasons: every 4 ms during an identical untimed pass right after it (NVML "
                                 "use around the timed region slows the ranks' lock-step, see bench.py)")),
            "sustained": sustained,
            "predict": predict,
            "other_models": extras,
            "gilbert": {"metric": "gilbert_rows_per_s", "rows": g_rows * N, "variant": int(best), **gil[best], "all_variants": gil},
        }
        if dp_parity is not None:
            line["dp_parity"] = dp_parity
        if sustained is not None:
            sustained["frac_of_sustained_peak"] = sustained["value"] / N * TRAIN_FLOP_PER_SAMPLE / 1e12 / pk["tf_sust"]
            sustained["note"] = "whole step (both kernels, launch hand-overs) over >= 2 s, against bf16_tflops_sustained"
        if N == 1 and not args.no_cpu_baseline:
            line["gilbert"]["cpu_baseline"] = gilbert_cpu_baseline()
            rate1, _, _ = cpu_port_rate(8192, 2, warmup=1)
            rows = int(min(65536, max(4096, (rate1 * 1.0) // 4096 * 4096)))
            n_steps = int(max(3, min(2000, 15.0 * rate1 / rows)))      # ~15 s of CPU work
            rate, dt, thr = cpu_port_rate(rows, n_steps, warmup=1)
            line["cpu_baseline"] = {"value": rate, "unit": "samples/s", "cores": thr, "kind": "port",
                                    "sample": "%d steps of %d rows (%.1f s) of the same 32-256-256-1 step, torch-CPU fp32 port; "
                                              "Spark local[*]/Keras unavailable on this box" % (n_steps, rows, dt)}
        emit(line)
    if N > 1:
        # captured NCCL work keeps communicator resources alive: leave without the collective teardown
        del stepper
        barrier()
        sys.stdout.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
