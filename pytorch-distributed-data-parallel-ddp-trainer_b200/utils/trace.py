"""Tracing and step timing (worklist row W15).

The reference has no profiling hooks at all (train_ddp.py:1-14 imports; only ``print``).  Here:

* ``nvtx_range(name)`` — an NVTX range when CUDA is available (shows up in Nsight Systems /
  ``ncu --nvtx``), a no-op otherwise.  The engine brackets every phase and bucket update of an
  eagerly launched step with these; graph replays carry one range per step.
* ``StepTimer`` — device-side step latency from CUDA events recorded on the launching stream, with
  NO per-step host synchronisation: events are kept in a ring and resolved only when a summary is
  requested (the reference's logging cadence, every 100th batch).  On CPU it falls back to
  ``time.perf_counter``.  ``summary()`` reduces with MAX over ranks, the only honest multi-GPU
  number.
* ``TraceWriter`` — optional Chrome-trace (``chrome://tracing`` / Perfetto) JSON of the step
  timeline, enabled with ``DDPB_TRACE=<path>``; one file per rank (``<path>.rank<r>.json``).
"""
from __future__ import annotations

import contextlib
import json
import os
import time

import torch
import torch.distributed as dist


def _nvtx_ok() -> bool:
    return torch.cuda.is_available()


@contextlib.contextmanager
def nvtx_range(name: str):
    if _nvtx_ok():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


def nvtx_mark(name: str) -> None:
    if _nvtx_ok():
        torch.cuda.nvtx.mark(name)


class StepTimer:
    """Ring of (start, end) timestamps per step.  ``begin()``/``end()`` cost two event records on
    the given CUDA stream (or two perf_counter reads on CPU) and never block the host."""

    def __init__(self, device, capacity: int = 256):
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.capacity = int(capacity)
        self._i = 0
        self._pending: list = []          # indices whose end has been recorded but not resolved
        self._ms: list[float] = []        # resolved step times since the last summary
        if self.cuda:
            self._ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                        for _ in range(self.capacity)]
        else:
            self._t = [0.0] * self.capacity

    def begin(self, stream=None) -> None:
        k = self._i % self.capacity
        if self.cuda:
            if len(self._pending) >= self.capacity:   # ring full: resolve the oldest (it finished long ago)
                self._resolve(1)
            self._ev[k][0].record(stream or torch.cuda.current_stream(self.device))
        else:
            self._t[k] = time.perf_counter()

    def end(self, stream=None) -> None:
        k = self._i % self.capacity
        if self.cuda:
            self._ev[k][1].record(stream or torch.cuda.current_stream(self.device))
            self._pending.append(k)
        else:
            self._ms.append((time.perf_counter() - self._t[k]) * 1e3)
        self._i += 1

    def _resolve(self, count: int | None = None) -> None:
        take = self._pending if count is None else self._pending[:count]
        for k in take:
            a, b = self._ev[k]
            b.synchronize()
            self._ms.append(a.elapsed_time(b))
        del self._pending[:len(take)]

    def summary(self, world_size: int = 1, control_device=None, reset: bool = True) -> dict:
        """{"steps", "ms_mean", "ms_max"} since the last summary; with world_size > 1 the mean and
        max are reduced with MAX over ranks (all ranks must call this together)."""
        if self.cuda:
            self._resolve()
        n = len(self._ms)
        mean = sum(self._ms) / n if n else 0.0
        mx = max(self._ms) if n else 0.0
        if world_size > 1 and dist.is_available() and dist.is_initialized():
            t = torch.tensor([mean, mx], dtype=torch.float64, device=control_device or "cpu")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            mean, mx = float(t[0]), float(t[1])
        if reset:
            self._ms = []
        return {"steps": n, "ms_mean": mean, "ms_max": mx}


class TraceWriter:
    """Minimal Chrome-trace writer: ``span(name)`` context managers on the host timeline plus
    ``counter`` samples.  Inert unless constructed with a path (or DDPB_TRACE is set)."""

    def __init__(self, path: str | None = None, rank: int = 0):
        base = path if path is not None else os.environ.get("DDPB_TRACE")
        self.path = f"{base}.rank{rank}.json" if base else None
        self.rank = rank
        self.events: list[dict] = []
        self._t0 = time.perf_counter()

    @property
    def enabled(self) -> bool:
        return self.path is not None

    def _us(self) -> float:
        return (time.perf_counter() - self._t0) * 1e6

    @contextlib.contextmanager
    def span(self, name: str, **args):
        if not self.enabled:
            with nvtx_range(name):
                yield
            return
        t = self._us()
        with nvtx_range(name):
            try:
                yield
            finally:
                self.events.append({"name": name, "ph": "X", "ts": t, "dur": self._us() - t,
                                    "pid": self.rank, "tid": 0, "args": args})

    def counter(self, name: str, **values) -> None:
        if self.enabled:
            self.events.append({"name": name, "ph": "C", "ts": self._us(), "pid": self.rank,
                                "args": values})

    def close(self) -> None:
        if not self.enabled:
            return
        tmp = self.path + ".tmp"
        with open(tmp, "w") as f:
            json.dump({"traceEvents": self.events, "displayTimeUnit": "ms"}, f)
        os.replace(tmp, self.path)
