"""Multi-rank plumbing for the measurement harness: one process per GPU, events sharded by rank
exactly as the MPI producer shards them (ref: psana_ray/producer.py:138-140 + psana MPIDataSource),
NO collective on the data path.  torch.distributed is used for the barrier around the timed region
and for the max-over-ranks of the timing only (backend nccl on GPUs, gloo on CPU)."""
from __future__ import annotations

import os


def env_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def init(backend: str, device=None):
    import torch.distributed as dist
    rank, world, _ = env_world()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if device is not None:
            dist.init_process_group(backend, device_id=device)
        else:
            dist.init_process_group(backend)
    return rank, world


def barrier():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def max_over_ranks(seconds: float, device="cpu") -> float:
    """Slowest rank's time: the job is done when the last rank is."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return float(seconds)
    t = torch.tensor([seconds], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_floats(value: float, device="cpu"):
    """Every rank's value, on every rank (list of world_size floats)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return [float(value)]
    t = torch.tensor([value], dtype=torch.float64, device=device)
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [float(o.item()) for o in out]


def shard(nevents: int, rank: int, size: int):
    """Event numbers of one rank: nevent % size == rank."""
    return range(rank, nevents, size)


def whole_job_throughput(units_per_rank: int, world: int, seconds_max: float) -> float:
    """Weak scaling: every rank processed units_per_rank units; value = all units / slowest rank's time."""
    return world * units_per_rank / seconds_max


def finalize():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


def bind_to_gpu_numa(device_index: int):
    """Pin this process to the CPUs next to its GPU (NVML's CPU affinity of the device) BEFORE it allocates pinned
    host memory, so the staging buffers are first-touched on the GPU's NUMA node and H2D copies do not cross the
    socket interconnect (8 ranks x ~55 GB/s is a host-memory question).  Returns the CPU list, or None
    when NVML / affinity control is unavailable (then nothing changes).  PSANA_RAY_NO_NUMA_BIND=1 disables it."""
    if os.environ.get("PSANA_RAY_NO_NUMA_BIND"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = device_index
        if vis:                                              # NVML numbers physical devices
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if device_index < len(ids) and ids[device_index].isdigit():
                phys = int(ids[device_index])
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None

