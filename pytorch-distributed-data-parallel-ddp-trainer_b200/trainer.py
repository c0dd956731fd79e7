"""Per-rank trainer: setup -> (resume) -> train loop -> save -> cleanup.

Same phases, log lines, checkpoint cadence and defaults as the reference worker
(reference train_ddp.py:17-212: SGD lr=0.01, CrossEntropy, rank-0 loss print every 100 batches,
``checkpoints/epoch_{N}.pt`` after every epoch, auto-resume from the newest one), with the
defects listed in SURVEY.md §0 fixed: data and model live on the rank's device, the resume path
works on every rank, and cleanup runs even when training raises.
"""
from __future__ import annotations

import os
import sys
import time

import torch
import torch.distributed as dist

from .parallel.engine import make_engine
from .parallel.process_group import cleanup, control_device, setup
from .utils import checkpoint as ckpt
from .utils.data import get_dataloader
from .utils.trace import StepTimer, TraceWriter, nvtx_range

LEARNING_RATE = 0.01   # reference train_ddp.py:41
LOG_INTERVAL = 100     # reference train_ddp.py:201


def _maybe_inject_failure(rank: int, epoch: int) -> None:
    """Fault injection for the recovery tests: ``DDP_B200_FAIL_ONCE="<rank>:<epoch>:<marker file>"``
    makes that rank raise at the start of that epoch, once (the marker file records that it happened)."""
    spec = os.environ.get("DDP_B200_FAIL_ONCE")
    if not spec:
        return
    r, e, marker = spec.split(":", 2)
    if int(r) == rank and int(e) == epoch and not os.path.exists(marker):
        with open(marker, "w") as f:
            f.write(f"rank {rank} epoch {epoch}\n")
        raise RuntimeError(f"injected failure on rank {rank} at epoch {epoch}")


def ddp_train(rank: int, world_size: int, epochs: int, batch_size: int, *, data_root: str = "./data",
              ckpt_dir: str = ckpt.CHECKPOINT_DIR, lr: float = LEARNING_RATE,
              log_interval: int = LOG_INTERVAL, dataset=None, backend: str | None = None,
              max_batches: int | None = None, stats: dict | None = None) -> None:
    """rank: this process' rank; world_size: processes in the job; epochs: passes over the data;
    batch_size: per-rank batch (reference train_ddp.py:17-25)."""
    device = setup(rank=rank, world_size=world_size, backend=backend)
    engine = None
    trace = TraceWriter(rank=rank)
    try:
        print(f"Rank {rank} initialized", flush=True)
        engine = make_engine(rank, world_size, device, batch_cap=batch_size, lr=lr)
        print(f"Rank {rank} model wrapped in DDP", flush=True)
        target = engine.acquire_inputs if device.type == "cuda" else None
        loader, sampler = get_dataloader(batch_size, world_size, rank, device=device, root=data_root,
                                         dataset=dataset, target=target)
        print(f"Rank {rank}: Dataloader ready", flush=True)
        print(f"Rank {rank}: Loss and Optimizer ready", flush=True)

        cdev = control_device()
        found, latest = ckpt.discover(rank, cdev, ckpt_dir)
        if not found:
            start_epoch = 0
            print(f"Rank {rank}: No checkpoint found, starting from scratch.", flush=True)
        else:
            holder = {"lr": lr}
            start_epoch = ckpt.resume(rank, engine.model, latest, cdev, engine.sync_params, holder)
            engine.set_lr(holder["lr"])
            print(f"Rank {rank}: Resumed from checkpoint, next epoch {start_epoch}", flush=True)

        timer = StepTimer(device)
        step_stream = getattr(engine, "main", None)
        for epoch in range(start_epoch, epochs):
            _maybe_inject_failure(rank, epoch)
            sampler.set_epoch(epoch)
            print(f"Rank {rank}: Starting epoch {epoch}", flush=True)
            t0 = time.perf_counter()
            seen = 0
            with trace.span(f"epoch {epoch}"):
                for batch_idx, (images, labels) in enumerate(loader):
                    if max_batches is not None and batch_idx >= max_batches:
                        break
                    timer.begin(step_stream)
                    with nvtx_range("step"):
                        loss = engine.step(images, labels)
                    timer.end(step_stream)
                    seen += int(images.shape[0])
                    if rank == 0 and batch_idx % log_interval == 0:
                        # the only host<-device read in the loop, at the reference's cadence
                        value = engine.read_loss(loss)
                        print(f"Epoch {epoch} | Batch {batch_idx} | Loss: {value:.4f}", flush=True)
                        trace.counter("loss", loss=value)
            if device.type == "cuda":
                engine.synchronize()
            step_stats = timer.summary(world_size, cdev)
            dt = torch.tensor([time.perf_counter() - t0, float(seen)], dtype=torch.float64, device=cdev)
            if world_size > 1:
                tmax = dt[:1].clone()
                dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
                total = dt[1:].clone()
                dist.all_reduce(total, op=dist.ReduceOp.SUM)
                dt = torch.cat([tmax, total])
            if rank == 0:
                secs, imgs = float(dt[0]), float(dt[1])
                print(f"Epoch {epoch} done | {imgs / max(secs, 1e-9):.0f} img/s | "
                      f"{secs * 1e3 / max(1, len(loader)):.3f} ms/step wall, "
                      f"{step_stats['ms_mean']:.3f} ms/step on device (max over ranks)", flush=True)
                trace.counter("throughput", img_per_s=imgs / max(secs, 1e-9))
                if stats is not None:
                    stats.setdefault("img_per_s", []).append(imgs / max(secs, 1e-9))
                    stats.setdefault("device_ms_per_step", []).append(step_stats["ms_mean"])
                ckpt.save_checkpoint(epoch, engine.state_dict(), engine.optimizer_state_dict(), ckpt_dir)
    except Exception as e:
        # The parent reports whichever rank's exception it sees first, often a peer's "connection closed":
        # every rank names its own failure before the group is torn down, so the root cause is in the log.
        print(f"Rank {rank}: failed with {type(e).__name__}: {e}", file=sys.stderr, flush=True)
        raise
    finally:
        if engine is not None:
            try:
                # on an error path the peers may never reach their close(): do not wait for them
                engine.close(sync_peers=sys.exc_info()[0] is None)
            except Exception as e:  # pragma: no cover - best effort on an already failing path
                print(f"Rank {rank}: engine close failed: {e}", flush=True)
        trace.close()
        cleanup()
