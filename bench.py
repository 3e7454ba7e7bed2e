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


def cpu_port_rate(rows_per_step, n_steps, warmup=1):
    """Times oracle/torch_cpu.py (the CPU port of the same training step) -- the ONLY place bench.py
    executes anything under oracle/.  Returns (samples/s, seconds, threads)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import torch
    import dlas_oracle as O
    import torch_cpu as T
    torch.set_num_threads(n_cores())
    n = rows_per_step * 4
    rng = np.random.default_rng(0)   # values do not matter for timing; shapes and dtypes do
    X = torch.from_numpy((rng.random((n, DIMS[0]), dtype=np.float32) * 2 - 1))
    y = torch.from_numpy((rng.random((n, 1), dtype=np.float32) * 2 - 1))
    m = T.TorchCpuMlp(DIMS, O.glorot_init(DIMS, O.BASE_SEED))
    for s in range(warmup):
        m.step(X[:rows_per_step], y[:rows_per_step])
    t0 = time.perf_counter()
    for s in range(n_steps):
        b = s % 4
        m.step(X[b * rows_per_step:(b + 1) * rows_per_step], y[b * rows_per_step:(b + 1) * rows_per_step])
    dt = time.perf_counter() - t0
    return rows_per_step * n_steps / dt, dt, torch.get_num_threads()


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
    if world > 1:
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

    def run_steps(n, first):
        if stepper is None:
            done = 0
            while done < n:     # sequential batches, wrapping at the end of the shard
                fb = (first + done) % n_batches
                cnt = min(n - done, n_batches - fb)
                ops.mlp_fit_steps(DIMS, model.loss, model.mode, model.cfg, model.params, model.m, model.v, X, y, B, cnt,
                                  model.iterations, losses[done:], model.grad_sum, model.workspace, first_batch=fb)
                model.iterations += cnt
                done += cnt
        else:
            stepper.run(n)

    def barrier():
        torch.cuda.synchronize()
        if N > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
    # Kernel time for the roofline: a SECOND pass over the same kind of steps with CUDA events around every launch of
    # the forward/backward kernel on its own stream (dlas_prof_*).  Kept out of the pass above because an event between
    # two kernels breaks their programmatic-dependent-launch overlap and would tax `value`; kernels inside a replayed
    # graph (N>1) cannot carry events at all, so there the same kernel is launched eagerly on the same shapes.
    ops.prof_read(ops.PROF_MLP_GRAD)
    ops.prof_enable(True)
    if stepper is None:
        run_steps(K, W + K)
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
    del xh, yh

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
        gil[str(var)] = {"rows_per_s": rows_s, "kernel_ms": gk_ms / gk_n,
                         "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                                      "frac": gbs / pk["hbm"],
                                      "traffic": (3.179e9 * g_rows / 200_000_000)}}   # ncu: 2.400 GB read + 0.779 GB written per 200M rows (profiles/)
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
        # dynamic ANN (config 4): 24 lags x 8 features -> 192-256-256-1, fp32 kernel, 65,536 windows per step
        dd = [192, 256, 256, 1]
        series, ys = ops.synth_rows(65536 + 23, 8)
        Wd = ops.lag_windows(series, 24)
        mdl = dlas.Sequential([dlas.Dense(192, 256, activation="relu"), dlas.Dense(256, 256, activation="relu"), dlas.Dense(256, 1)])
        mdl.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
        yd = ys[23:].contiguous()
        t = timed(lambda: mdl.train_on_batch(Wd, yd), 5)
        extras["dynamic_ann_192_256_256_1"] = {"samples_per_s": 65536 / t, "mode": "fp32", "rows_per_step": 65536,
                                               "frac_of_bf16_sustained_peak": 591360 * 65536 / t / 1e12 / pk["tf_sust"]}
        # static ANN 8-64-64-1 (config 1 shape), fp32 kernel
        X8, y8 = ops.synth_rows(65536, 8)
        m8 = dlas.Sequential([dlas.Dense(8, 64, activation="relu"), dlas.Dense(64, 64, activation="relu"), dlas.Dense(64, 1)])
        m8.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
        t = timed(lambda: m8.train_on_batch(X8, y8), 10)
        extras["static_ann_8_64_64_1"] = {"samples_per_s": 65536 / t, "mode": "fp32", "rows_per_step": 65536}
        # LSTM 2x128, T=48, batch 4096 (config 5)
        Xs, ysq = ops.synth_series(4096, 48, 8)
        ml = dlas.Sequential([dlas.LSTM(128, input_shape=(48, 8), return_sequences=True), dlas.LSTM(128), dlas.Dense(1)])
        ml.compile(loss="mse", optimizer=dlas.SGD(lr=0.001, momentum=0.99, decay=1e-6, nesterov=True))
        t = timed(lambda: ml.train_on_batch(Xs, ysq), 3)
        extras["lstm_2x128_T48_batch4096"] = {"sequences_per_s": 4096 / t, "ms_per_step": t * 1e3, "mode": "fp32",
                                              "frac_of_bf16_sustained_peak": 57.41e6 * 4096 / t / 1e12 / pk["tf_sust"]}
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

    if rank == 0:
        kern_s = kern_ms / max(1, kern_n) * 1e-3
        tfs = TRAIN_FLOP_PER_SAMPLE * B / kern_s / 1e12
        # my kernels per step: fp32 mode transpose + forward/backward + reduce + update; tc mode forward/backward +
        # (reduce fused with the update) at N=1, forward/backward + reduce + update at N>1 (NCCL's kernel not counted)
        # tc mode: forward/backward + one finishing kernel (reduce [+ NVLink push/poll exchange] + update [+ cursor advance]) at any N
        launches_per_step = (4 if mode == "fp32" else 2) if N == 1 else (5 if mode == "fp32" else 2)
        line = {
            "metric": "static_ann_train_samples_per_s", "value": value, "unit": "samples/s", "n_gpus": N, "steps": K,
            "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if mode == "fp32" else "f16-operands/f32-accumulate (10-bit mantissa, as TF32)",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "rows_per_gpu_per_step": B, "global_batch": B * N, "mode": mode,
                       "rows_resident_per_gpu": rows_rank, "l2": "inputs (%.1f GB per GPU) larger than L2; every step reads new rows"
                       % (rows_rank * BYTES_PER_SAMPLE / 1e9), "parallelism": "dp%d" % N, "final_loss": final_loss},
            "roofline": {"bound": "tensor", "achieved": tfs, "peak": pk["tf_sust"], "unit": "TFLOP/s", "frac": tfs / pk["tf_sust"],
                         "traffic": (10.56e6 if (mode == "tc" and B == 65536) else None),   # ncu dram read+write per launch, profiles/r01_ncu_full_tc_final.txt
                         "kernel": "mlp forward+loss+backward", "kernel_ms": kern_s * 1e3,
                         "kernel_share_of_step": (kern_ms / max(1, kern_n)) / (ms / K),
                         "kernel_timing": ("second pass of %d steps with CUDA events around every launch" % K) if N == 1
                         else "10 eager launches on the same shapes (graph replays cannot carry events)",
                         "peak_source": pk["src"] + ", bf16 sustained",
                         "hbm_frac": BYTES_PER_SAMPLE * B / kern_s / 1e9 / pk["hbm"]},
            "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": B * BYTES_PER_SAMPLE, "d2h_bytes_per_step": 4},
            "gpu_launches": launches_per_step * K,
            "clocks": dict(clk.summary(), sm_mhz_in_kernel=device_mhz,
                           when=("NVML every 4 ms inside the timed region" if in_region else
                                 "sm_mhz_in_kernel: measured by the forward/backward kernel inside the timed region; NVML "
                                 "clock and reasons: every 4 ms during an identical untimed pass right after it (NVML "
                                 "use around the timed region slows the ranks' lock-step, see bench.py)")),
            "other_models": extras,
            "gilbert": {"metric": "gilbert_rows_per_s", "rows": g_rows * N, "variant": int(best), **gil[best], "all_variants": gil},
        }
        if N == 1 and not args.no_cpu_baseline:
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
