"""Data-parallel plumbing (one process per GPU, torch.distributed; NCCL over NVLink on B200, gloo
in the CPU tests).  Replaces the Spark-driver averaging that BASELINE.json asks for (the snapshot
has none): rows of every global batch are split into `world` contiguous
slices, each rank computes gradient SUMS over its slice, one all-reduce of ONE flat bucket
(n_params gradients + the loss sum) per step, then every rank applies the same update with
grad_scale = 1/global_batch.  The path has no other collective."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def world_info():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_bounds(global_batch, rank, world):
    """Rows [lo, hi) of a global batch that rank `rank` owns; the union over ranks is the batch."""
    if global_batch % world:
        raise ValueError("global batch %d is not a multiple of the world size %d" % (global_batch, world))
    per = global_batch // world
    return rank * per, (rank + 1) * per


def local_rows(n_rows, global_batch, rank, world):
    """Index tensor of the rows of a dataset (n_rows, whole batches only) that belong to `rank`
    when every consecutive block of `global_batch` rows is one step's batch."""
    lo, hi = shard_bounds(global_batch, rank, world)
    n_batches = n_rows // global_batch
    base = torch.arange(n_batches, dtype=torch.int64)[:, None] * global_batch
    return (base + torch.arange(lo, hi, dtype=torch.int64)[None, :]).reshape(-1)


def allreduce_sum_(bucket):
    """In-place SUM all-reduce of the flat bucket (no-op for a single process)."""
    _, world = world_info()
    if world > 1:
        dist.all_reduce(bucket, op=dist.ReduceOp.SUM)
    return bucket


def dp_step(grad_fn, step_fn, bucket, global_rows):
    """One synchronous data-parallel step: grad_fn() fills `bucket` with this rank's sums,
    the bucket is all-reduced, step_fn(1/global_rows) applies the update.  Returns the bucket."""
    grad_fn()
    allreduce_sum_(bucket)
    step_fn(1.0 / float(global_rows))
    return bucket


class PeerBucket:
    """This rank's gradient bucket in memory that every peer GPU of the box can load from (torch symmetric memory:
    CUDA VMM allocations exchanged between the ranks' processes), laid out as dlas_dp_allreduce_update expects:
    two parity copies of (n_params gradient sums + loss sum), then `world` arrival flags.  PyTorch only allocates and
    maps the memory; the reduction over NVLink and the update are dlas_dp_allreduce_update's one kernel."""

    def __init__(self, n_params, device):
        import torch.distributed._symmetric_memory as symm_mem
        self.rank, self.world = world_info()
        self.parity_stride = (n_params + 1 + 63) // 64 * 64
        self.flags_offset = 2 * self.parity_stride
        total = self.flags_offset + 64
        self.buf = symm_mem.empty(total, dtype=torch.float32, device=device)
        self.handle = symm_mem.rendezvous(self.buf, dist.group.WORLD)
        self.buf.zero_()                      # flags start at 0; iteration it announces it+1 >= 1
        torch.cuda.synchronize(device)
        dist.barrier()                        # nobody announces before everybody has cleared
        self.bases = torch.tensor([int(p) for p in self.handle.buffer_ptrs], dtype=torch.int64, device=device)


class PeerRecvBuffer:
    """This rank's RECEIVE buffer for dlas_mlp_dp_step_at, in memory every peer GPU of the box can store to (torch
    symmetric memory): [2 parities][world sources][words] of 4-byte floats, zeroed once (zero bits = not arrived; the
    kernel clears what it has consumed).  PyTorch only
    allocates and maps it; the pushes over NVLink, the polling, the sum and the update are the library's kernel."""

    def __init__(self, n_bytes, device):
        import torch.distributed._symmetric_memory as symm_mem
        self.rank, self.world = world_info()
        self.buf = symm_mem.empty(n_bytes // 4, dtype=torch.float32, device=device)
        self.handle = symm_mem.rendezvous(self.buf, dist.group.WORLD)
        self.buf.zero_()                      # all-zero bits: a word that has not arrived
        torch.cuda.synchronize(device)
        dist.barrier()                        # nobody pushes before everybody has cleared
        self.bases = torch.tensor([int(p) for p in self.handle.buffer_ptrs], dtype=torch.int64, device=device)


class LocalRecvBuffer:
    """world == 1: dlas_mlp_dp_step_at has nobody to exchange with; the buffer is ordinary device memory."""

    def __init__(self, n_bytes, device):
        self.rank, self.world = 0, 1
        self.buf = torch.zeros(n_bytes // 4, dtype=torch.float32, device=device)
        self.bases = torch.tensor([self.buf.data_ptr()], dtype=torch.int64, device=device)


def average_(tensors):
    """In-place MEAN over ranks of each tensor (periodic parameter averaging).  NCCL averages inside the collective;
    gloo (CPU tests) has no AVG, so there the sum is divided on the host."""
    _, world = world_info()
    if world <= 1:
        return
    for t in tensors:
        if dist.get_backend() == "nccl":
            dist.all_reduce(t, op=dist.ReduceOp.AVG)
        else:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
            t /= world


def bind_host_to_gpu(local_rank):
    """One process per GPU on a multi-socket host: restrict this process to the CPUs the driver reports as local to its GPU
    (NVML ideal CPU affinity), BEFORE any pinned host buffer is allocated, so that the buffers it streams batches from are
    first-touched on the GPU's own NUMA node and the host->device copies do not cross the socket link.  Returns the number
    of CPUs kept, or 0 when nothing was changed (no NVML, single node, or an affinity mask set by the launcher already)."""
    try:
        import pynvml
        all_cpus = os.sched_getaffinity(0)
        if len(all_cpus) < (os.cpu_count() or 0):
            return 0                                  # the launcher has bound us already: keep its choice
        pynvml.nvmlInit()
        try:
            h = pynvml.nvmlDeviceGetHandleByIndex(int(local_rank))
            n_words = ((os.cpu_count() or 64) + 63) // 64
            words = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
        finally:
            pynvml.nvmlShutdown()
        cpus = {w * 64 + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1} & all_cpus
        if not cpus or cpus == all_cpus:
            return 0
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return 0
