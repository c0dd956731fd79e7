"""Process-group runtime: setup()/cleanup() with the gloo fallback on CPU.

Same contract and log lines as the reference helpers (reference utils.py:5-19):
backend = nccl when CUDA is available else gloo, env:// rendezvous.  Differences, each fixing a
defect recorded in SURVEY.md §0:
  * MASTER_ADDR / MASTER_PORT default to 127.0.0.1:29500 instead of failing when unset;
  * the CUDA device is chosen from LOCAL_RANK (or rank % device_count), not the global rank, and
    is set BEFORE the NCCL communicator is created;
  * an optional timeout is forwarded so a dead peer surfaces as an error.
"""
from __future__ import annotations

import datetime
import os

import torch
import torch.distributed as dist


def pick_backend(backend: str | None = None) -> str:
    return backend or ("nccl" if torch.cuda.is_available() else "gloo")


def local_device_index(rank: int) -> int:
    if "LOCAL_RANK" in os.environ:
        return int(os.environ["LOCAL_RANK"])
    n = torch.cuda.device_count()
    return rank % n if n else 0


def setup(rank: int, world_size: int, backend: str | None = None, timeout_s: float | None = None):
    """Join the process group; returns the torch.device this rank computes on."""
    backend = pick_backend(backend)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    kwargs = {}
    if timeout_s is not None:
        kwargs["timeout"] = datetime.timedelta(seconds=timeout_s)
    device = torch.device("cpu")
    if backend == "nccl":
        idx = local_device_index(rank)
        torch.cuda.set_device(idx)
        device = torch.device("cuda", idx)
        kwargs["device_id"] = device
    dist.init_process_group(backend=backend, world_size=world_size, rank=rank, **kwargs)
    print(f"Rank: {rank} has initialized its process group with world size {world_size}", flush=True)
    return device


def cleanup() -> None:
    if not dist.is_initialized():
        return
    rank = dist.get_rank()
    dist.destroy_process_group()
    print(f"Rank {rank} cleaned up.", flush=True)


def control_device() -> torch.device:
    """Device control-plane tensors must live on for the active backend (NCCL cannot move CPU
    tensors — the reference's CPU flag tensors break there)."""
    if dist.is_initialized() and dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")
