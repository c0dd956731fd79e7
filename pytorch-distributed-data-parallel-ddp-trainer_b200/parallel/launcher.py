"""Launcher: one training process per rank, started either by torchrun or by ourselves.

Covers the reference entry point (reference train_ddp.py:215-228) and fixes its three launch
defects: `--world_size` exists, rendezvous variables are filled in
automatically for the self-spawn path, and under torchrun (RANK/WORLD_SIZE/LOCAL_RANK present)
the process does NOT spawn again.
"""
from __future__ import annotations

import os
import socket
from typing import Callable, Sequence

import torch


def under_torchrun() -> bool:
    return "RANK" in os.environ and "WORLD_SIZE" in os.environ


def default_world_size() -> int:
    """WORLD_SIZE under torchrun, else the visible GPU count, else 2 CPU processes (the
    reference's hard-coded value, train_ddp.py:221)."""
    if "WORLD_SIZE" in os.environ:
        return int(os.environ["WORLD_SIZE"])
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    return n if n > 0 else 2


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(worker: Callable, world_size: int | None, args: Sequence = (), max_restarts: int = 0) -> None:
    """Run ``worker(rank, world_size, *args)`` on every rank.

    torchrun: this process IS one rank (its own ``--max-restarts`` applies).  Otherwise spawn
    ``world_size`` local processes (spawn start method, like the reference) with MASTER_ADDR/PORT set
    to a free local port.  ``max_restarts`` > 0 turns the reference's recovery story — "rerun the same
    command" (README.md:74), which resumes from the newest checkpoint — into something the launcher
    does itself: when any rank dies, the whole group is torn down (spawn does that) and started again
    on a fresh port, up to ``max_restarts`` times.
    """
    if under_torchrun():
        rank = int(os.environ["RANK"])
        ws = int(os.environ["WORLD_SIZE"])
        if world_size is not None and world_size != ws:
            print(f"--world_size {world_size} ignored: torchrun set WORLD_SIZE={ws}", flush=True)
        worker(rank, ws, *args)
        return
    ws = world_size if world_size is not None else default_world_size()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    user_port = "MASTER_PORT" in os.environ
    attempt = 0
    while True:
        if not user_port:
            os.environ["MASTER_PORT"] = str(free_port())
        try:
            if ws == 1:
                worker(0, 1, *args)
            else:
                from torch.multiprocessing.spawn import spawn
                spawn(worker, args=(ws, *args), nprocs=ws)
            return
        except KeyboardInterrupt:
            raise
        except Exception as e:
            if attempt >= max_restarts:
                raise
            attempt += 1
            first = str(e).strip().splitlines()[-1] if str(e).strip() else type(e).__name__
            print(f"launcher: worker group failed ({first}); restart {attempt}/{max_restarts} "
                  f"(training resumes from the newest checkpoint)", flush=True)
