"""Clip sharding across the GPUs of one box (one process per GPU).

The hot path is embarrassingly parallel over clips -- every reduction (normaliser statistics,
LayerNorm rows, attention over one clip's positions) is per clip -- so ranks take contiguous clip
ranges, weights are replicated, and there is NO collective on the data path (the reference shards
the same way with ``common_utils.shard``, transcription_pretrain_no_voiceprints.py:231-233).
torch.distributed is used only for the rendezvous, the timing barrier and the max-over-ranks of the
elapsed time; ``gather_states`` is the optional all-gather of encoder states for a single consumer.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def env_rank_world() -> tuple[int, int, int]:
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def shard_range(n_items: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [start, stop) of ``n_items`` for ``rank``; sizes differ by at most one."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"rank {rank} of {world}")
    base, extra = divmod(n_items, world)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def init_process_group(backend: str | None = None) -> tuple[int, int, int]:
    rank, local_rank, world = env_rank_world()
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kwargs = {}
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            kwargs["device_id"] = torch.device("cuda", local_rank)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    return rank, local_rank, world


def bind_host_to_gpu(device_index: int) -> list[int]:
    """Pin this process (and so its pinned staging buffers, first-touch) to the CPUs NVML reports as local to the GPU.
    With one process per GPU every rank moves ~15 GB/s through host memory on the host-facing path; without the binding
    half of that crosses the socket interconnect.  Returns the CPU list, or [] when NVML / affinity is unavailable."""
    try:
        import pynvml
        pynvml.nvmlInit()
        props = torch.cuda.get_device_properties(device_index)
        bus = f"{props.pci_domain_id:08x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
        handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        n_cpu = os.cpu_count() or 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, (n_cpu + 63) // 64)
        cpus = [i for i in range(n_cpu) if (mask[i // 64] >> (i % 64)) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return []


def barrier():
    if dist.is_initialized():
        dist.barrier()


def max_over_ranks(value: float, device=None) -> float:
    """Max of a scalar over all ranks (identity for a single process)."""
    if not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value: float, device=None) -> float:
    if not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device or ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def gather_states(states: torch.Tensor) -> torch.Tensor:
    """Optional: all-gather ``(B_local, T, D)`` encoder states into ``(B_total, T, D)`` on every rank
    (NCCL over NVLink/NVSwitch on the GPU box; off the timed hot path)."""
    if not dist.is_initialized():
        return states
    world = dist.get_world_size()
    out = torch.empty((world * states.shape[0],) + tuple(states.shape[1:]), dtype=states.dtype, device=states.device)
    dist.all_gather_into_tensor(out, states.contiguous())
    return out
