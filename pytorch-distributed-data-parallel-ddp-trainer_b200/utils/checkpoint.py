"""Checkpoint save / discovery / resume, file-compatible with the reference in both directions.

On-disk contract (reference train_ddp.py:204-209): ``./checkpoints/epoch_{N}.pt``
= ``{"epoch": int, "model": OrderedDict (no "module." prefix, fp32 CPU tensors, `_metadata`),
"optimizer": torch.optim.SGD.state_dict()-shaped dict}``, written by rank 0 after every epoch.

Resume intent (reference train_ddp.py:49-185): rank 0 alone finds and reads the newest file,
every rank ends with identical weights and ``start_epoch = saved_epoch + 1``.  The reference
does this with 32 blocking broadcasts and crashes doing so (SURVEY.md §5.4 bugs B1-B3); here it
is one flat parameter broadcast plus one small header broadcast.
"""
from __future__ import annotations

import os
import re
from pathlib import Path

import torch
import torch.distributed as dist

CHECKPOINT_DIR = "./checkpoints"
_EPOCH_RE = re.compile(r"^epoch_(\d+)\.pt$")


def sgd_state_dict(lr: float, num_params: int) -> dict:
    """Exactly what ``torch.optim.SGD(params, lr=lr).state_dict()`` emits for plain SGD
    (reference train_ddp.py:41,208): no per-parameter state, one param group."""
    return {
        "state": {},
        "param_groups": [{
            "lr": lr, "momentum": 0, "dampening": 0, "weight_decay": 0, "nesterov": False,
            "maximize": False, "foreach": None, "differentiable": False, "fused": None,
            "params": list(range(num_params)),
        }],
    }


def find_latest(directory: str = CHECKPOINT_DIR) -> Path | None:
    """Newest checkpoint by ctime (the reference's rule, train_ddp.py:55-57) restricted to files
    named ``epoch_N.pt`` (a stray file no longer crashes torch.load); ctime ties — e.g. a fresh
    clone — resolve to the highest epoch number."""
    d = Path(directory)
    if not d.exists():
        return None
    cands = []
    for f in d.iterdir():
        m = _EPOCH_RE.match(f.name)
        if f.is_file() and m:
            cands.append((f.stat().st_ctime, int(m.group(1)), f))
    if not cands:
        return None
    return max(cands, key=lambda c: (c[0], c[1]))[2]


def save_checkpoint(epoch: int, model_state, optimizer_state: dict,
                    directory: str = CHECKPOINT_DIR) -> str:
    """Write ``epoch_{epoch}.pt`` (write-to-temp then rename, so a crash never leaves a
    truncated file for the next resume to pick up)."""
    os.makedirs(directory, exist_ok=True)
    path = os.path.join(directory, f"epoch_{epoch}.pt")
    tmp = path + ".tmp"
    torch.save({"epoch": int(epoch), "model": model_state, "optimizer": optimizer_state}, tmp)
    os.replace(tmp, path)
    return path


def load_checkpoint(path) -> dict:
    ckpt = torch.load(str(path), map_location="cpu", weights_only=True)
    for key in ("epoch", "model"):
        if key not in ckpt:
            raise KeyError(f"{path}: not a trainer checkpoint (missing '{key}')")
    return ckpt


def discover(rank: int, device: torch.device, directory: str = CHECKPOINT_DIR):
    """Rank 0 looks for a checkpoint (creating the directory if needed) and tells everyone
    (reference train_ddp.py:49-63: flag broadcast + barrier).  Returns (found, path-on-rank-0)."""
    flag = torch.zeros(1, dtype=torch.int64, device=device)
    latest = None
    if rank == 0:
        latest = find_latest(directory)
        if latest is not None:
            flag[0] = 1
        else:
            os.makedirs(directory, exist_ok=True)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(flag, src=0)
        dist.barrier()
    return bool(flag.item()), latest


def resume(rank: int, model, latest, device: torch.device, sync_params, lr_holder: dict | None = None) -> int:
    """Rank 0 reads ``latest`` into the model; ``sync_params()`` then broadcasts the flat
    parameter arena from rank 0 (one message) and a two-number header (start epoch, lr) follows.
    Returns start_epoch = saved epoch + 1."""
    header = torch.zeros(2, dtype=torch.float64, device=device)
    if rank == 0:
        ckpt = load_checkpoint(latest)
        model.load_state_dict(ckpt["model"])
        header[0] = ckpt["epoch"] + 1
        groups = (ckpt.get("optimizer") or {}).get("param_groups") or [{}]
        header[1] = float(groups[0].get("lr", float("nan")))
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(header, src=0)
    sync_params()
    h = header.cpu()
    if lr_holder is not None and h[1] == h[1]:
        lr_holder["lr"] = float(h[1])
    return int(h[0].item())
