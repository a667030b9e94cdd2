"""Multi-GPU plumbing for the scoring path: events are sharded across ranks, models are replicated, and NOTHING is
exchanged while scoring.  torch.distributed is used only to line the ranks up (barrier) and to reduce
the per-rank timings / counts for reporting."""
from __future__ import annotations

import os
from typing import Sequence, Tuple


def env_rank() -> Tuple[int, int, int]:
    """(rank, world_size, local_rank) from the torchrun environment; single process when absent."""
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def shard(n_total: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) slice of a stream of n_total events for `rank` (sizes differ by at most one)."""
    base, rem = divmod(n_total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def init(backend: str, device=None):
    """Initialises the default process group when WORLD_SIZE > 1; returns torch.distributed or None."""
    rank, world, local = env_rank()
    if world <= 1:
        return None
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    if not dist.is_initialized():
        kwargs = {"device_id": device} if (device is not None and backend == "nccl") else {}
        dist.init_process_group(backend, rank=rank, world_size=world, **kwargs)
    return dist


def max_over_ranks(dist, values: Sequence[float], device=None):
    """Element-wise MAX of per-rank timings (a multi-GPU number is the slowest rank's)."""
    import torch
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t]


def sum_over_ranks(dist, values: Sequence[float], device=None):
    import torch
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t]


def whole_job_rate(units_per_rank_per_step: int, world: int, steps: int, max_ms: float) -> float:
    """events/s of the whole job: everything all ranks processed divided by the slowest rank's time."""
    return units_per_rank_per_step * world * steps / (max_ms / 1e3)


def bind_to_gpu_numa_node(local_rank: int):
    """Pins this process to the CPUs of the NUMA node its GPU hangs off, before any pinned buffer is allocated: pinned
    pages are then local to the GPU's root complex, so N ranks feeding N GPUs do not all pull through one socket
    (the reference's deployment analogue: one subtask per GPU, S/package.scala:63-65).  Returns the node or None when
    the topology cannot be read; never raises."""
    import os
    try:
        import torch
        p = torch.cuda.get_device_properties(local_rank)
        bdf = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & os.sched_getaffinity(0)
        if allowed:
            os.sched_setaffinity(0, allowed)
            return node
    except Exception:
        pass
    return None
