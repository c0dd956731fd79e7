"""Generic bucketed data-parallel wrapper for any CUDA nn.Module, built on the symmetric-memory
communicator: what DDP's Reducer + NCCL + torch.optim.SGD do (reference train_ddp.py:34,199,200),
with per-bucket fused all-reduce + 1/W + SGD kernels launched from gradient-ready hooks on a side
stream, so communication AND the optimizer overlap the rest of backward.

Parameters are re-homed into the peer-mapped arena (so the two-shot kernel can push updated
slices into every rank), gradients accumulate straight into one of two arena copies (step
parity, see csrc/comm.cu), buckets are contiguous slices in reverse registration order
(gradient-ready order), capped by ``bucket_mb``.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .comm import ALGO_AUTO, SymmetricComm


def _align(n: int, a: int) -> int:
    return (n + a - 1) // a * a


def plan_buckets(spans, bucket_mb: float):
    """[(start, end)] float ranges covering the arena front to back, each a multiple of 4 floats and at
    most ``bucket_mb`` MiB; ``spans`` = (offset, numel) of the parameters in arena order.  A cut may fall
    inside a parameter."""
    cap = max(4, int(bucket_mb * (1 << 20) / 4) // 4 * 4)
    total = _align(spans[-1][0] + spans[-1][1], 4)
    cuts, start = [], 0
    for i, (o, n) in enumerate(spans):
        end = _align(o + n, 4) if i < len(spans) - 1 else total
        while end - start >= cap:          # close full buckets, cutting through this parameter if needed
            cuts.append((start, start + cap))
            start += cap
        if i == len(spans) - 1 and end > start:
            cuts.append((start, end))
    return cuts


class BucketedDataParallel:
    def __init__(self, module: torch.nn.Module, rank: int, world_size: int, lr: float = 0.01,
                 bucket_mb: float = 25.0, algo: int = ALGO_AUTO, use_graph: bool = False):
        """``use_graph``: after two eager steps with a given input shape, the whole step (forward, backward,
        bucket exchanges on the side streams) is captured once per (shape, step parity) in a CUDA graph and
        replayed — the module must then do the same work every step (no data-dependent control flow)."""
        self.module, self.rank, self.world_size, self.lr, self.algo = module, rank, world_size, lr, algo
        self.use_graph = use_graph
        self._graphs: dict = {}
        self._static: dict = {}
        self._seen: dict = {}
        self._pool = None
        params = [p for p in module.parameters() if p.requires_grad]
        self.device = params[0].device
        # arena order = reverse registration order, so buckets fill front to back during backward
        order = list(reversed(params))
        offs, off = [], 0
        for p in order:
            off = _align(off, 4)
            offs.append(off)
            off += p.numel()
        n = _align(off, 4)
        self.numel = n
        region = _align(n * 4, 256)
        self.param_off, self.grad_off = 0, [region, 2 * region]
        self.comm = SymmetricComm(rank, world_size, 3 * region)
        self.flat_params = self.comm.tensor(0, n)
        self.flat_grads = [self.comm.tensor(o, n) for o in self.grad_off]
        self.flat_params.zero_()
        for g in self.flat_grads:
            g.zero_()
        self._views = []
        with torch.no_grad():
            for p, o in zip(order, offs):
                v = self.flat_params[o:o + p.numel()].view(p.shape)
                v.copy_(p.data)
                p.data = v
                self._views.append((p, o))
        # buckets: contiguous [start, end) float ranges, multiples of 4 floats, at most `bucket_mb` each.
        # Like DDP's Reducer (25 MiB cap) — but a bucket boundary may also fall INSIDE a parameter, so one
        # very large gradient (the widened CNN's 94 MB fc1 weight) leaves as several exchanges that run
        # side by side on different streams instead of one long kernel.
        cuts = plan_buckets([(o, p.numel()) for p, o in self._views], bucket_mb)
        self.buckets = []
        self._buckets_of = {i: [] for i in range(len(self._views))}
        for (b0, b1) in cuts:
            members = [i for i, (p, o) in enumerate(self._views) if o < b1 and o + p.numel() > b0]
            for m in members:
                self._buckets_of[m].append(len(self.buckets))
            self.buckets.append({"start": b0, "n": b1 - b0, "members": members, "pending": 0})
        self.lr_dev = torch.full((1,), lr, device=self.device)
        # bucket exchanges are spread over a few side streams so several can be in flight at once
        self.sides = [torch.cuda.Stream(self.device) for _ in range(min(3, max(1, len(self.buckets))))]
        self.side = self.sides[0]
        self._ready = [torch.cuda.Event() for _ in self.buckets]
        self._done = [torch.cuda.Event() for _ in self.sides]
        self.step_count = 0
        self._parity = 0
        for i, (p, _) in enumerate(self._views):
            p.register_post_accumulate_grad_hook(self._make_hook(i))
        # Layers that can write a weight gradient straight into a caller-given buffer (models/linear.py:
        # ``weight_grad_out``) get their slice of the arena: no zero-fill and no accumulate pass over it.
        self._direct = {}
        for mod in module.modules():
            if hasattr(mod, "weight_grad_out") and isinstance(getattr(mod, "weight", None), torch.nn.Parameter):
                for i, (p, _) in enumerate(self._views):
                    if p is mod.weight:
                        self._direct[i] = mod
        self.sync_params()

    def sync_params(self, src: int = 0) -> None:
        for mod in getattr(self, "_direct", {}).values():  # packed weight copies are stale after this
            if hasattr(mod, "_w_packed_valid"):
                mod._w_packed_valid = False
        torch.cuda.synchronize(self.device)
        if self.world_size > 1:
            self.comm.broadcast(self.param_off, self.numel * 4, root=src, slot=7)
            torch.cuda.synchronize(self.device)

    def _make_hook(self, idx: int):
        def hook(param):
            if idx in self._direct:
                p, o = self._views[idx]
                dst = self.flat_grads[self._parity][o:o + p.numel()]
                if param.grad.data_ptr() != dst.data_ptr():  # autograd cloned instead of adopting the buffer
                    dst.view(p.shape).copy_(param.grad)
            for b_i in self._buckets_of[idx]:
                b = self.buckets[b_i]
                b["pending"] -= 1
                if b["pending"] == 0:
                    self._launch_bucket(b_i)
                elif b["pending"] < 0:
                    raise RuntimeError("BucketedDataParallel: a parameter received a second gradient in one "
                                       "step (gradient accumulation / two backward passes are not supported: "
                                       "one backward per step())")
        return hook

    def _launch_bucket(self, b_i: int) -> None:
        b = self.buckets[b_i]
        cur = torch.cuda.current_stream(self.device)
        side = self.sides[b_i % len(self.sides)]
        self._ready[b_i].record(cur)
        side.wait_event(self._ready[b_i])
        goff = self.grad_off[self._parity]
        b["launched"] = True
        self.comm.allreduce_sgd(goff + b["start"] * 4, self.param_off + b["start"] * 4, b["n"],
                                self.lr_dev, self.algo, slot=b_i % 6, stream=side)
        if self.use_graph:  # an eagerly launched step is host-bound: the extra stream bookkeeping costs more there
            self._prepack_updated_layers()

    def _prepack_updated_layers(self) -> None:
        """A layer that keeps a packed copy of its weight (models/linear.py) gets it rebuilt as soon as the
        last bucket holding a piece of that weight has been launched: on a side stream, under the rest of
        backward, instead of at the start of the next forward."""
        for idx, mod in self._direct.items():
            if idx in self._prepacked or not hasattr(mod, "prepack"):
                continue
            bs = self._buckets_of[idx]
            if not all(self.buckets[b].get("launched") for b in bs):
                continue
            self._prepacked.add(idx)
            home = self.sides[bs[-1] % len(self.sides)]
            for s_i in sorted({b % len(self.sides) for b in bs}):
                side = self.sides[s_i]
                if side is not home:
                    ev = torch.cuda.Event()
                    ev.record(side)
                    home.wait_event(ev)
            with torch.cuda.stream(home):
                mod.prepack(stream=home)

    def step(self, x: torch.Tensor, y: torch.Tensor, loss_fn) -> torch.Tensor:
        """forward + backward (+ per-bucket fused all-reduce/SGD from the grad hooks)."""
        self._parity = self.step_count & 1
        if not self.use_graph:
            return self._step_eager(x, y, loss_fn)
        key = (tuple(x.shape), x.dtype, tuple(y.shape), y.dtype, self._parity, loss_fn)
        seen = self._seen.get(key, 0)
        self._seen[key] = seen + 1
        if seen < 2:          # warm-up: lazy initialisation, cuDNN plans, workspace growth
            return self._step_eager(x, y, loss_fn)
        skey = key[:4]
        if skey not in self._static:
            self._static[skey] = (torch.empty_like(x), torch.empty_like(y))
        xs, ys = self._static[skey]
        xs.copy_(x, non_blocking=True)
        ys.copy_(y, non_blocking=True)
        entry = self._graphs.get(key)
        if entry is None:
            graph = torch.cuda.CUDAGraph()
            torch.cuda.synchronize(self.device)
            with torch.cuda.graph(graph, pool=self._pool):
                loss = self._step_eager(xs, ys, loss_fn, count=False)
            if self._pool is None:
                self._pool = graph.pool()
            # the graph holds raw pointers into the layers' workspaces: keep those alive with it
            keep = [dict(m._ws) for m in self.module.modules() if isinstance(getattr(m, "_ws", None), dict)]
            entry = self._graphs[key] = (graph, loss, keep)
        graph, loss = entry[0], entry[1]
        graph.replay()
        self.step_count += 1
        return loss.clone()

    def _step_eager(self, x: torch.Tensor, y: torch.Tensor, loss_fn, count: bool = True) -> torch.Tensor:
        g = self.flat_grads[self._parity]
        for i, (p, o) in enumerate(self._views):
            v = g[o:o + p.numel()].view(p.shape)
            if i in self._direct:      # the layer's backward writes the whole gradient into the arena itself
                self._direct[i].weight_grad_out = v
                p.grad = None
            else:                      # gradients accumulate in place into this step's arena copy
                v.zero_()
                p.grad = v
        for b in self.buckets:
            b["pending"] = len(b["members"])
            b["launched"] = False
        self._prepacked = set()
        loss = loss_fn(self.module(x), y)
        loss.backward()
        # A parameter that got no gradient this step (a branch not taken, a frozen path) never fires its
        # hook: its bucket is exchanged now with the zero-filled gradient, so every rank issues the same
        # sequence of collectives whatever its local graph looked like.
        for b_i, b in enumerate(self.buckets):
            if not b["launched"]:
                self._launch_bucket(b_i)
        cur = torch.cuda.current_stream(self.device)
        for side, ev in zip(self.sides, self._done):
            ev.record(side)
            cur.wait_event(ev)  # parameters are final before the next forward
        if count:
            self.step_count += 1
        return loss.detach()

    def close(self, sync_peers: bool = True) -> None:
        """See B200DataParallel.close: peers must be done reading this rank's buffer before it goes."""
        torch.cuda.synchronize(self.device)
        self._graphs.clear()
        self._static.clear()
        if sync_peers and self.world_size > 1 and dist.is_available() and dist.is_initialized():
            dist.barrier()
        self.comm.close()
