"""SimpleCNN: conv(1->32,3x3,p1)+ReLU -> conv(32->64,3x3,p1)+ReLU -> flatten -> linear(50176->10).

Same function, parameter names, shapes and initialisation as the reference classifier
(reference model.py:4-20; state_dict keys net.0.*, net.2.*, fl.*), but built on a flat
parameter/gradient arena (arena.py) with two execution paths:

* ``SimpleCNN``      — CPU path (the reference's own gloo/CPU fallback contract,
                       reference utils.py:6): torch ops + autograd writing straight into the
                       gradient arena.  Also the numerical oracle for the GPU path.
* ``SimpleCNNB200``  — B200 path: an explicit kernel sequence over the native sm_100a library
                       (no autograd engine, no cuDNN/cuBLAS): conv1 SIMT, conv2 fwd/dgrad/wgrad
                       on tcgen05 tensor cores, fused softmax-CE, split linear layer.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import torch
import torch.nn.functional as F

from .arena import ParamArena

IMG_H = IMG_W = 28
C1, C2, NCLS = 32, 64, 10
FEATURES = C2 * IMG_H * IMG_W  # 50176

SPECS = (
    ("net.0.weight", (C1, 1, 3, 3)),
    ("net.0.bias", (C1,)),
    ("net.2.weight", (C2, C1, 3, 3)),
    ("net.2.bias", (C2,)),
    ("fl.weight", (NCLS, FEATURES)),
    ("fl.bias", (NCLS,)),
)
# torch's Module.state_dict() records one `_metadata` entry per sub-module
METADATA_KEYS = ("", "net", "net.0", "net.1", "net.2", "net.3", "net.4", "fl")
# gradient-ready order: the linear layer's gradients exist right after its backward, before the
# expensive conv backward — 96% of the bytes
BUCKET_GROUPS = (("fl.weight", "fl.bias"),
                 ("net.0.weight", "net.0.bias", "net.2.weight", "net.2.bias"))


def _default_init_(arena: ParamArena, generator: torch.Generator | None = None) -> None:
    """torch's default Conv2d/Linear init (kaiming-uniform a=sqrt(5) == U(-1/sqrt(fan_in), ..)
    for weight and bias), drawn in module-construction order so that, under the same seed, the
    values equal what ``SimpleCNN()`` of the reference produces (reference model.py:8-16)."""
    with torch.no_grad():
        for wname, bname in (("net.0.weight", "net.0.bias"), ("net.2.weight", "net.2.bias"),
                             ("fl.weight", "fl.bias")):
            w = arena.view(wname)
            fan_in = w[0].numel()
            bound = 1.0 / math.sqrt(fan_in)
            wcpu = torch.empty(w.shape, dtype=torch.float32)
            # kaiming_uniform_(a=sqrt(5)), spelled exactly as torch.nn.init computes it so the
            # bounds agree to the last bit: gain = sqrt(2/(1+a^2)), bound = sqrt(3)*gain/sqrt(fan_in)
            wbound = math.sqrt(3.0) * (math.sqrt(2.0 / (1 + math.sqrt(5) ** 2)) / math.sqrt(fan_in))
            wcpu.uniform_(-wbound, wbound, generator=generator)
            bcpu = torch.empty(arena.view(bname).shape, dtype=torch.float32)
            bcpu.uniform_(-bound, bound, generator=generator)
            w.copy_(wcpu)
            arena.view(bname).copy_(bcpu)


class _ModelBase:
    """Shared state_dict / parameter plumbing (what nn.Module gives the reference)."""

    arena: ParamArena

    def state_dict(self) -> OrderedDict:
        return self.arena.state_dict(METADATA_KEYS)

    def load_state_dict(self, sd, strict: bool = True) -> None:
        self.arena.load_state_dict(sd, strict=strict)

    def named_parameters(self):
        for s in self.arena.specs:
            yield s.name, self.arena.view(s.name)

    def parameters(self):
        for _, p in self.named_parameters():
            yield p

    def buckets(self):
        return self.arena.make_buckets(BUCKET_GROUPS)

    def train(self, mode: bool = True):  # no dropout / batch-norm: nothing to switch
        return self

    def eval(self):
        return self


class SimpleCNN(_ModelBase):
    """CPU execution path (also usable on CUDA as an fp32 oracle)."""

    def __init__(self, device="cpu", params=None, grads=None, init: bool = True,
                 generator: torch.Generator | None = None):
        self.arena = ParamArena(SPECS, device=device, params=params, grads=grads)
        if init:
            _default_init_(self.arena, generator)
        self.arena.params.requires_grad_(True)
        # autograd accumulates in place into an existing .grad => gradients land in the arena
        self.arena.params.grad = self.arena.grads

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        a = self.arena
        h = F.relu(F.conv2d(x, a.view("net.0.weight"), a.view("net.0.bias"), padding=1))
        h = F.relu(F.conv2d(h, a.view("net.2.weight"), a.view("net.2.bias"), padding=1))
        return F.linear(h.flatten(1), a.view("fl.weight"), a.view("fl.bias"))

    __call__ = forward

    def forward_backward(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """loss = CE(model(x), y); gradients of the mean loss are written to the gradient arena."""
        self.arena.grads.zero_()
        loss = F.cross_entropy(self.forward(x), y)
        loss.backward()
        if self.arena.params.grad is not self.arena.grads:  # defensive: keep the arena authoritative
            self.arena.grads.copy_(self.arena.params.grad)
            self.arena.params.grad = self.arena.grads
        return loss.detach()

    def forward_backward_staged(self, x: torch.Tensor, y: torch.Tensor, on_bucket_ready) -> torch.Tensor:
        """Same result as forward_backward, but the backward pass runs bucket by bucket in
        gradient-ready order (BUCKET_GROUPS: the linear layer first) and ``on_bucket_ready(i)`` is
        called as soon as bucket i's gradients are in the arena — the CPU counterpart of DDP's
        autograd-hook launch (reference train_ddp.py:199): the engine starts that
        bucket's all-reduce while the convolution backward is still running."""
        a = self.arena
        v = {s.name: a.view(s.name) for s in a.specs}
        h = F.relu(F.conv2d(x, v["net.0.weight"], v["net.0.bias"], padding=1))
        h = F.relu(F.conv2d(h, v["net.2.weight"], v["net.2.bias"], padding=1))
        loss = F.cross_entropy(F.linear(h.flatten(1), v["fl.weight"], v["fl.bias"]), y)
        for i, names in enumerate(BUCKET_GROUPS):
            last = i == len(BUCKET_GROUPS) - 1
            grads = torch.autograd.grad(loss, [v[n] for n in names], retain_graph=not last)
            with torch.no_grad():
                for n, g in zip(names, grads):
                    a.grad_view(n).copy_(g)
            on_bucket_ready(i)
        return loss.detach()

    def sgd_step(self, lr: float) -> None:
        with torch.no_grad():
            self.arena.params.add_(self.arena.grads, alpha=-lr)


class SimpleCNNB200(_ModelBase):
    """sm_100a execution path: explicit forward/backward kernel sequence on one GPU.

    The step is split at the points where a gradient bucket becomes complete, so the engine
    can launch that bucket's all-reduce while the rest of backward runs
    (reference train_ddp.py:199 + DDP Reducer hooks):

        forward_loss -> backward_head   (bucket 0: fl.weight, fl.bias ready)
                     -> backward_body   (bucket 1: net.2.*, net.0.* ready)
    """

    def __init__(self, batch_cap: int, device="cuda", params=None, grads=None, init: bool = True,
                 generator: torch.Generator | None = None, alt_grads=None, precision: str | None = None):
        """``precision``:

        * ``"bf16x3"`` (default; env DDPB_PRECISION) — the reference's numerics as torch runs them
          on CUDA (reference model.py:8-16: fp32 storage, TF32 convolutions, fp32 linear layer),
          matched or exceeded everywhere: activations and activation gradients are stored at 4 bytes
          per value — as split planes x = hi + lo (two bf16 terms, a 16-bit significand) where a
          tensor-core GEMM consumes them, as plain fp32 where the linear layer does; the convolution
          GEMMs multiply split operands with three MMAs per K step (hi.hi + lo.hi + hi.lo, fp32
          accumulation: more operand bits than TF32's 11); conv1's forward is TF32 like cuDNN's; the
          linear layer is fp32 FMA like cuBLAS with TF32 off.  tests/test_precision_gpu.py measures
          the error of every op against an fp64 oracle next to torch's own CUDA path.
        * ``"bf16"`` — fast mode: bf16 activation storage and single bf16 tensor-core operands;
          NOT the reference's precision, never the default, reported separately.
        """
        import os
        from ..ops import native as nat
        self.nat = nat
        nat.lib()
        self.device = torch.device(device)
        self.batch_cap = int(batch_cap)
        self.precision = precision or os.environ.get("DDPB_PRECISION", "bf16x3")
        if self.precision not in nat.PRECISIONS:
            raise ValueError(f"precision must be one of {nat.PRECISIONS}, got {self.precision!r}")
        self.x3 = self.precision == "bf16x3"
        # Loss fused into the tail of the linear forward (one launch less).  Measured on B200 at batch 64:
        # 122.0 vs 118.6 us/step — the last block's serial tail costs more than the programmatic
        # dependent launch of a separate loss kernel, so it is opt-in (DDPB_FUSED_LOSS=1).
        self.fused_loss_max = nat.linear_fwd_f32_xent_max_batch() if os.environ.get("DDPB_FUSED_LOSS", "0") == "1" else 0
        self.arena = ParamArena(SPECS, device=self.device, params=params, grads=grads)
        if init:
            _default_init_(self.arena, generator)
        H, W = IMG_H, IMG_W
        with torch.cuda.device(self.device):
            dev = self.device
            tc_fmt = "split" if self.x3 else "bf16"   # tensors a tensor-core GEMM consumes
            self.act1 = nat.alloc_planes(C1, batch_cap, H, W, dev, tc_fmt)
            self.act2 = nat.alloc_planes(C2, batch_cap, H, W, dev, "f32" if self.x3 else "bf16")  # linear layer's input
            self.dact2 = nat.alloc_planes(C2, batch_cap, H, W, dev, tc_fmt)
            # bf16x3: dact1 feeds the exact-fp32 conv1 weight gradient (CUDA cores): plain fp32 planes
            self.dact1 = nat.alloc_planes(C1, batch_cap, H, W, dev, "f32" if self.x3 else "bf16")
            if self.x3:
                #: exact-fp32 partial sums of dact2 per channel (net.2.bias gradient), written by the
                #: linear backward; (nsplit, nbx) of the last backward_head() in self.db2_extent
                self.db2_part = nat.scratch(nat.linear_bwd_f32_bias_part_bytes(C2, batch_cap, H, W), dev)
                self.db2_extent = (0, 0)
                self.lin_partial = nat.scratch(nat.linear_fwd_f32_scratch_bytes(NCLS, C2, batch_cap, H, W), dev)
                self.lin_bwd_scr = nat.scratch(nat.linear_bwd_scratch_bytes(NCLS, C2, batch_cap, H, W), dev)
                self.lin_packed = None
                #: net.2.weight as split bf16 tensor-core tile images (forward + data gradient, hi + lo);
                #: rewritten by the fused net.2 bucket update, or by pack_weights() after any other change
                self.w2_packed = torch.zeros(nat.conv3x3_pack_split_bytes(C1, C2) // 2, dtype=torch.bfloat16,
                                             device=dev)
            else:
                self.lin_partial = nat.scratch(nat.linear_fwd_tc_scratch_bytes(NCLS, C2, batch_cap, H, W), dev)
                #: fl.weight pre-packed into bf16 tensor-core tiles (refresh with pack_weights() after
                #: every change of fl.weight; the step engine does it right after the bucket-0 update)
                self.lin_packed = torch.zeros(nat.linear_pack_bytes(NCLS, C2, H, W), dtype=torch.uint8,
                                              device=dev)
                self.w2_packed = None
            self.logits = torch.zeros(batch_cap, NCLS, device=dev)
            self.dlogits = torch.zeros(batch_cap, NCLS, device=dev)
            self.dl_packed = torch.zeros(nat.dl_packed_bytes(batch_cap), dtype=torch.uint8, device=dev)
            self.loss = torch.zeros(1, device=dev)
            self.scr_w2 = nat.scratch(nat.conv3x3_wgrad_scratch_bytes(C1, C2), dev)
            self.scr_w1 = nat.scratch(max(nat.conv1_wgrad_tc_scratch_bytes(1, C1),
                                          nat.conv1_wgrad_f32_scratch_bytes(1, C1)), dev)
            self.lr_dev = torch.full((1,), 0.01, device=dev)
        a = self.arena
        self.w1, self.b1 = a.view("net.0.weight"), a.view("net.0.bias")
        self.w2, self.b2 = a.view("net.2.weight"), a.view("net.2.bias")
        self.wf, self.bf = a.view("fl.weight"), a.view("fl.bias")
        # Gradient arenas: one per step parity when a communicator double-buffers them so that a
        # single flag round per bucket also covers buffer reuse (csrc/comm.cu); otherwise one.
        self.grad_sets = []
        for gbuf in ([a.grads] if alt_grads is None else [a.grads, alt_grads]):
            self.grad_sets.append({
                "flat": gbuf,
                "w1": a.view("net.0.weight", gbuf),  # followed in memory by net.0.bias
                "w2": a.view("net.2.weight", gbuf), "b2": a.view("net.2.bias", gbuf),
                "wf": a.view("fl.weight", gbuf), "bf": a.view("fl.bias", gbuf),
            })

    def load_state_dict(self, sd, strict: bool = True) -> None:
        super().load_state_dict(sd, strict=strict)
        self.pack_weights()  # the tensor-core tiles of fl.weight must follow the new values

    # ------------------------------------------------------------------ step pieces
    def pack_weights(self, stream=None, fl: bool = True, conv2: bool = True) -> None:
        """Refresh the tensor-core tile copies of the master weights: bf16 mode — fl.weight -> bf16
        hi+lo tiles of the linear forward/backward; bf16x3 mode — net.2.weight -> split tiles of the
        conv2 forward/data gradient (the linear layer reads the fp32 master weights directly)."""
        if self.x3:
            if conv2:
                self.nat.conv3x3_pack_split(self.w2, self.w2_packed, stream)
        elif fl:
            self.nat.linear_pack(self.wf, self.lin_packed, NCLS, C2, IMG_H, IMG_W, stream)

    def forward_loss(self, x: torch.Tensor, y: torch.Tensor, batch: int | None = None, stream=None,
                     parity: int = 0, pack: bool = True):
        nat = self.nat
        gs = self.grad_sets[parity]
        b = int(batch if batch is not None else x.shape[0])
        assert 0 < b <= self.batch_cap
        cap, H, W = self.batch_cap, IMG_H, IMG_W
        if pack:
            self.pack_weights(stream)
        if self.x3:
            nat.conv1_fwd_tc_split(x, self.w1, self.b1, self.act1, b, cap, H, W, stream)
            # act2 feeds the fp32 linear layer: plain fp32 planes
            nat.conv3x3_fwd_split(self.act1, self.w2_packed, self.b2, self.act2, C1, C2, b, cap, H, W,
                                  stream, out_f32=True)
            if b <= self.fused_loss_max:   # small batch: loss fused into the linear forward's tail
                nat.linear_fwd_f32_xent(self.act2, self.wf, self.lin_partial, self.bf, y, self.logits,
                                        self.dlogits, gs["bf"], self.loss, NCLS, C2, b, cap, H, W, stream)
                return self.loss
            S = nat.linear_fwd_f32(self.act2, self.wf, self.lin_partial, NCLS, C2, b, cap, H, W, stream)
            nat.softmax_xent(self.lin_partial, S, self.bf, y, self.logits, self.dlogits, gs["bf"],
                             self.loss, NCLS, b, stream)
            return self.loss
        nat.conv1_fwd_tc(x, self.w1, self.b1, self.act1, b, cap, H, W, stream)
        nat.conv3x3_fwd(self.act1, self.w2, self.b2, self.act2, b, cap, H, W, stream)
        S = nat.linear_fwd_tc(self.act2, self.lin_packed, self.lin_partial, NCLS, C2, b, cap, H, W, stream)
        nat.softmax_xent(self.lin_partial, S, self.bf, y, self.logits, self.dlogits, gs["bf"],
                         self.loss, NCLS, b, stream, dl_packed=self.dl_packed)
        return self.loss

    def backward_head(self, batch: int, stream=None, parity: int = 0, defer_reduce: bool = False):
        """Linear-layer backward: completes bucket 0 (fl.weight; fl.bias came from the loss kernel).
        bf16x3 mode with ``defer_reduce``: the fl.weight gradient stays as per-image-split partials and
        backward_head_reduce() sums them later (the engine does it on the bucket's own stream, off the
        critical path — only dact2 is needed by what follows on the main stream)."""
        gs = self.grad_sets[parity]
        if self.x3:
            # defer_reduce: only the dX half (dact2 + the exact bias sums) runs here; the dW half is
            # backward_head_reduce()'s job
            self.db2_extent = self.nat.linear_bwd_f32(
                self.act2, self.dlogits, self.wf, self.dact2, None if defer_reduce else gs["wf"],
                self.lin_bwd_scr, NCLS, C2, batch, self.batch_cap, IMG_H, IMG_W, stream, split_out=True,
                bias_part=self.db2_part, mode=1 if defer_reduce else 0)
            self._head_batch = batch
            return
        self.nat.linear_bwd_tc(self.act2, self.dl_packed, self.lin_packed, self.dact2, gs["wf"],
                               NCLS, C2, batch, self.batch_cap, IMG_H, IMG_W, stream)

    def backward_head_reduce(self, stream=None, parity: int = 0) -> None:
        """The deferred dW half of the linear backward: fl.weight gradient (per-image-split partials +
        their fixed-order sum) into the gradient arena.  Reads act2 and dlogits only."""
        if self.x3:
            self.nat.linear_bwd_f32(self.act2, self.dlogits, self.wf, self.dact2, self.grad_sets[parity]["wf"],
                                    self.lin_bwd_scr, NCLS, C2, self._head_batch, self.batch_cap, IMG_H, IMG_W,
                                    stream, split_out=True, mode=2)

    def backward_conv2_wgrad(self, batch: int, stream=None, parity: int = 0):
        """conv2 weight + bias gradient (independent of the dgrad chain: may run on its own stream)."""
        gs = self.grad_sets[parity]
        if self.x3:
            self.nat.conv3x3_wgrad_split(self.dact2, self.act1, gs["w2"], gs["b2"], self.scr_w2, C1, C2, batch,
                                         self.batch_cap, IMG_H, IMG_W, stream,
                                         bias_parts=(self.db2_part, *self.db2_extent))
            return
        self.nat.conv3x3_wgrad(self.dact2, self.act1, gs["w2"], gs["b2"], self.scr_w2, C1, C2, batch,
                               self.batch_cap, IMG_H, IMG_W, stream)

    def backward_conv2_wgrad_partials(self, batch: int, stream=None) -> int:
        """Like backward_conv2_wgrad but leaves the gradient as per-CTA accumulator images in
        ``self.scr_w2`` (returns their count) for the fused reduce + all-reduce + SGD kernel."""
        fn = self.nat.conv3x3_wgrad_split_partials if self.x3 else self.nat.conv3x3_wgrad_partials
        return fn(self.dact2, self.act1, self.scr_w2, C1, C2, batch, self.batch_cap, IMG_H, IMG_W, stream)

    def backward_dgrad(self, batch: int, stream=None) -> None:
        """conv2 data gradient (with ReLU'): dact1 from dact2.  The LAST reader of net.2.weight (bf16
        mode) / of its split tile images (bf16x3 mode) in a step — the engine must not let the net.2
        bucket update start before this has finished."""
        if self.x3:
            self.nat.conv3x3_dgrad_split(self.dact2, self.w2_packed, self.act1, self.dact1, C1, C2, batch,
                                         self.batch_cap, IMG_H, IMG_W, stream, out_f32=True)
            return
        self.nat.conv3x3_dgrad(self.dact2, self.w2, self.act1, self.dact1, batch, self.batch_cap,
                               IMG_H, IMG_W, stream)

    def backward_conv1_wgrad(self, x: torch.Tensor, batch: int, stream=None, parity: int = 0) -> None:
        """conv1 weight + bias gradient into the gradient arena (tensor-core kernel + split-K reduce)."""
        gs = self.grad_sets[parity]
        fn = self.nat.conv1_wgrad_f32 if self.x3 else self.nat.conv1_wgrad_tc
        fn(self.dact1, x, gs["w1"], self.scr_w1, 1, C1, batch, self.batch_cap, IMG_H, IMG_W, stream)

    def backward_conv1_wgrad_partials(self, x: torch.Tensor, batch: int, stream=None) -> int:
        """conv1 gradient left as per-CTA partials in ``self.scr_w1`` (returns their count) for the
        fused reduce + all-reduce + SGD kernel."""
        fn = self.nat.conv1_wgrad_f32_partials if self.x3 else self.nat.conv1_wgrad_tc_partials
        return fn(self.dact1, x, self.scr_w1, 1, C1, batch, self.batch_cap, IMG_H, IMG_W, stream)

    def backward_dgrad_conv1(self, x: torch.Tensor, batch: int, stream=None, parity: int = 0):
        """conv2 data gradient (with ReLU') followed by the conv1 weight + bias gradient."""
        self.backward_dgrad(batch, stream)
        self.backward_conv1_wgrad(x, batch, stream, parity)

    def backward_dgrad_conv1_partials(self, x: torch.Tensor, batch: int, stream=None) -> int:
        """Like backward_dgrad_conv1 but leaves the conv1 gradient as per-CTA partials."""
        self.backward_dgrad(batch, stream)
        return self.backward_conv1_wgrad_partials(x, batch, stream)

    def backward_body(self, x: torch.Tensor, batch: int, stream=None, parity: int = 0):
        """conv2 wgrad + dgrad, conv1 wgrad: completes bucket 1."""
        self.backward_conv2_wgrad(batch, stream, parity)
        self.backward_dgrad_conv1(x, batch, stream, parity)

    def forward_backward(self, x: torch.Tensor, y: torch.Tensor, stream=None,
                         parity: int = 0) -> torch.Tensor:
        b = int(x.shape[0])
        self.forward_loss(x, y, b, stream, parity)
        self.backward_head(b, stream, parity)
        self.backward_body(x, b, stream, parity)
        return self.loss

    def launches_per_step(self, batch: int) -> int:
        """Kernels one forward+backward launches (bench.py reports it as part of gpu_launches)."""
        from ..ops import layout as lay
        tail = 1 if lay.tiles_for(batch, IMG_H, IMG_W) * lay.TILE_M > batch * 900 else 0
        # conv1, conv2, linear, loss | (tail zero) linear bwd | wgrad+reduce, dgrad, conv1 wgrad
        # (its split-K reduce is fused into the bucket update kernel); the fp32 linear backward adds
        # the fixed-order reduce of its per-image-split dW partials
        fused_loss = 1 if (self.x3 and batch <= self.fused_loss_max) else 0
        # bf16x3: the linear backward is two kernels (dX half, dW half) + the reduce of the dW partials
        return 4 - fused_loss + (1 + tail) + 2 + 1 + 1 + (2 if self.x3 else 0)

    def sgd_step(self, lr: float | None = None, stream=None) -> None:
        if lr is not None:
            self.lr_dev.fill_(float(lr))
        n = self.arena.numel
        self.nat.sgd(self.arena.params[:n], self.arena.grads[:n], self.lr_dev, stream)

    def train_step(self, x: torch.Tensor, y: torch.Tensor, lr: float = 0.01) -> torch.Tensor:
        """Single-GPU convenience: forward, backward, SGD.  Returns the (device) loss tensor."""
        self.forward_backward(x, y)
        self.sgd_step(lr)
        return self.loss

    @torch.no_grad()
    def predict(self, x: torch.Tensor) -> torch.Tensor:
        """Forward only; returns logits [batch, 10] (a view of an internal buffer)."""
        nat = self.nat
        b = int(x.shape[0])
        cap, H, W = self.batch_cap, IMG_H, IMG_W
        y = torch.zeros(b, dtype=torch.int64, device=self.device)
        scratch_db = torch.empty(NCLS, device=self.device)
        self.pack_weights()
        if self.x3:
            nat.conv1_fwd_tc_split(x, self.w1, self.b1, self.act1, b, cap, H, W)
            nat.conv3x3_fwd_split(self.act1, self.w2_packed, self.b2, self.act2, C1, C2, b, cap, H, W)
            S = nat.linear_fwd_f32(self.act2, self.wf, self.lin_partial, NCLS, C2, b, cap, H, W)
        else:
            nat.conv1_fwd_tc(x, self.w1, self.b1, self.act1, b, cap, H, W)
            nat.conv3x3_fwd(self.act1, self.w2, self.b2, self.act2, b, cap, H, W)
            S = nat.linear_fwd_tc(self.act2, self.lin_packed, self.lin_partial, NCLS, C2, b, cap, H, W)
        nat.softmax_xent(self.lin_partial, S, self.bf, y, self.logits, self.dlogits, scratch_db,
                         self.loss, NCLS, b)
        return self.logits[:b]
