"""Flat parameter / gradient arena with reference-named views.

All parameters of a model live in ONE contiguous fp32 buffer (and all gradients in a second
one), in state_dict order.  Consequences:

* ``state_dict()`` is a set of views under the reference's key names, so the checkpoint
  contract (reference train_ddp.py:205-209) falls out of ``torch.save``;
* gradient buckets are static slices of the gradient arena — no Reducer, no bucket rebuild,
  no copy-in / copy-out (what DDP does at reference train_ddp.py:34,199; SURVEY.md X15,X16);
* the fused all-reduce + SGD kernel updates a slice of the parameter arena in place.

Buckets are listed in gradient-ready order (last layer first), mirroring the order DDP's
rebuilt buckets take for this model.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Sequence

import torch


@dataclass(frozen=True)
class ParamSpec:
    name: str
    shape: tuple
    offset: int  # in floats

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


@dataclass(frozen=True)
class Bucket:
    """A contiguous slice [start, start+numel) of the arenas, all-reduced as one message."""
    start: int
    numel: int
    names: tuple


class ParamArena:
    def __init__(self, specs: Sequence[tuple], device="cpu", params: torch.Tensor | None = None,
                 grads: torch.Tensor | None = None, align: int = 4):
        off = 0
        self.specs: list[ParamSpec] = []
        for name, shape in specs:
            off = (off + align - 1) // align * align
            self.specs.append(ParamSpec(name, tuple(shape), off))
            off += self.specs[-1].numel
        self.numel = off
        self.padded_numel = (off + 3) // 4 * 4
        if params is None:
            params = torch.zeros(self.padded_numel, dtype=torch.float32, device=device)
        if grads is None:
            grads = torch.zeros(self.padded_numel, dtype=torch.float32, device=device)
        assert params.numel() >= self.numel and grads.numel() >= self.numel
        self.params = params
        self.grads = grads
        self.by_name = {s.name: s for s in self.specs}

    # ------------------------------------------------------------------ views
    def view(self, name: str, flat: torch.Tensor | None = None) -> torch.Tensor:
        s = self.by_name[name]
        base = self.params if flat is None else flat
        return base[s.offset:s.offset + s.numel].view(s.shape)

    def grad_view(self, name: str) -> torch.Tensor:
        return self.view(name, self.grads)

    def slice_of(self, names: Sequence[str]) -> tuple:
        """(start, numel) of the smallest contiguous range covering `names` (must be adjacent)."""
        ss = sorted((self.by_name[n] for n in names), key=lambda s: s.offset)
        start = ss[0].offset
        end = ss[-1].offset + ss[-1].numel
        covered = sum(s.numel for s in ss)
        assert end - start - covered < 4 * len(ss), "bucket parameters must be adjacent"
        return start, end - start

    def make_buckets(self, groups: Sequence[Sequence[str]]) -> list[Bucket]:
        out = []
        for names in groups:
            start, n = self.slice_of(names)
            out.append(Bucket(start, n, tuple(names)))
        return out

    # ------------------------------------------------------------------ state dict
    def state_dict(self, metadata_keys: Sequence[str]) -> OrderedDict:
        """CPU fp32 clones under the reference key names, with torch's `_metadata` attribute."""
        sd = OrderedDict()
        for s in self.specs:
            sd[s.name] = self.view(s.name).detach().to("cpu", torch.float32).clone()
        md = OrderedDict()
        for k in metadata_keys:
            md[k] = {"version": 1}
        sd._metadata = md
        return sd

    def load_state_dict(self, sd, strict: bool = True) -> None:
        keys = {k[len("module."):] if k.startswith("module.") else k: k for k in sd.keys()}
        missing = [s.name for s in self.specs if s.name not in keys]
        unexpected = [k for k in keys if k not in self.by_name]
        if strict and (missing or unexpected):
            raise RuntimeError(f"Error(s) in loading state_dict: Missing key(s): {missing}; "
                               f"Unexpected key(s): {unexpected}")
        with torch.no_grad():
            for s in self.specs:
                if s.name in keys:
                    src = sd[keys[s.name]]
                    if tuple(src.shape) != s.shape:
                        raise RuntimeError(f"size mismatch for {s.name}: {tuple(src.shape)} vs {s.shape}")
                    self.view(s.name).copy_(src.to(torch.float32))
