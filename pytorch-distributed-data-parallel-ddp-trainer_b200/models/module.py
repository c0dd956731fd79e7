"""``SimpleCNNModule``: the reference classifier as a real ``torch.nn.Module`` (reference model.py:4-20)
whose CUDA forward/backward are the native sm_100a kernels of this repository.

Same sub-module structure as the reference (``net = Sequential(Conv2d, ReLU, Conv2d, ReLU, Flatten)``,
``fl = Linear``), hence the same ``state_dict`` keys, the same default initialisation under a seed, and
everything ``nn.Module`` gives a user: ``.to()``, ``.parameters()``, hooks, any ``torch.optim``
optimizer, any loss on the logits, ``DistributedDataParallel`` if wanted.

On a CUDA device ``forward`` is ONE ``torch.autograd.Function`` over the kernels of the
reference-precision mode (csrc/: 3xTF32 conv1, bf16x3 conv2 on tcgen05 tensor cores, fp32 linear
layer); its ``backward`` takes whatever gradient arrives for the logits and produces the six parameter
gradients with the native backward kernels.  On CPU (or without the built library) it is the plain
torch module.  The fastest path remains the step engine (parallel/engine.py: static kernel sequence,
fused all-reduce+SGD, CUDA graph); this class is the "PyTorch stays the framework layer" entry point.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .simple_cnn import C1, C2, FEATURES, IMG_H, IMG_W, NCLS


class _Workspace:
    """Activation planes and scratch for one (device, batch capacity); grown on demand."""

    def __init__(self, nat, device, cap: int):
        H, W = IMG_H, IMG_W
        self.cap = cap
        self.act1 = nat.alloc_planes(C1, cap, H, W, device, "split")
        self.act2 = nat.alloc_planes(C2, cap, H, W, device, "f32")
        self.dact2 = nat.alloc_planes(C2, cap, H, W, device, "split")
        self.dact1 = nat.alloc_planes(C1, cap, H, W, device, "f32")
        self.lin_partial = nat.scratch(nat.linear_fwd_f32_scratch_bytes(NCLS, C2, cap, H, W), device)
        self.lin_bwd_scr = nat.scratch(nat.linear_bwd_scratch_bytes(NCLS, C2, cap, H, W), device)
        self.db2_part = nat.scratch(nat.linear_bwd_f32_bias_part_bytes(C2, cap, H, W), device)
        self.scr_w2 = nat.scratch(nat.conv3x3_wgrad_scratch_bytes(C1, C2), device)
        self.scr_w1 = nat.scratch(nat.conv1_wgrad_f32_scratch_bytes(1, C1), device)
        self.w2_packed = torch.zeros(nat.conv3x3_pack_split_bytes(C1, C2) // 2, dtype=torch.bfloat16,
                                     device=device)


class _NativeLogits(torch.autograd.Function):
    """logits = fl(relu(conv2(relu(conv1(x))))) through the native kernels."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, wf, bf, module):
        nat, ws = module._nat, module._workspace(x.device, x.shape[0])
        b, cap, H, W = x.shape[0], ws.cap, IMG_H, IMG_W
        x = x.contiguous().float()
        ws_args = [t.detach().contiguous() for t in (w1, b1, w2, b2, wf, bf)]
        w1c, b1c, w2c, b2c, wfc, bfc = ws_args
        nat.conv3x3_pack_split(w2c, ws.w2_packed)          # the weights may have changed since last call
        nat.conv1_fwd_tc_split(x, w1c, b1c, ws.act1, b, cap, H, W)
        nat.conv3x3_fwd_split(ws.act1, ws.w2_packed, b2c, ws.act2, C1, C2, b, cap, H, W, out_f32=True)
        S = nat.linear_fwd_f32(ws.act2, wfc, ws.lin_partial, NCLS, C2, b, cap, H, W)
        logits = ws.lin_partial[:S * b * NCLS].view(S, b, NCLS).sum(0) + bfc
        ctx.module, ctx.ws, ctx.batch = module, ws, b
        ctx.save_for_backward(x, wfc)
        module._live = ctx      # the workspace holds THIS forward's activations until its backward
        return logits

    @staticmethod
    def backward(ctx, grad_logits):
        module, ws, b = ctx.module, ctx.ws, ctx.batch
        if module._live is not ctx:
            raise RuntimeError("SimpleCNNModule keeps the activations of its most recent forward only: "
                               "call backward() before the next forward() on the same module")
        nat, cap, H, W = module._nat, ws.cap, IMG_H, IMG_W
        x, wfc = ctx.saved_tensors
        dl = grad_logits.contiguous().float()
        dev = dl.device
        dwf = torch.empty(NCLS, FEATURES, device=dev)
        ext = nat.linear_bwd_f32(ws.act2, dl, wfc, ws.dact2, dwf, ws.lin_bwd_scr, NCLS, C2, b, cap, H, W,
                                 split_out=True, bias_part=ws.db2_part)
        dw2 = torch.empty(C2, C1, 3, 3, device=dev)
        db2 = torch.empty(C2, device=dev)
        nat.conv3x3_wgrad_split(ws.dact2, ws.act1, dw2, db2, ws.scr_w2, C1, C2, b, cap, H, W,
                                bias_parts=(ws.db2_part, *ext))
        nat.conv3x3_dgrad_split(ws.dact2, ws.w2_packed, ws.act1, ws.dact1, C1, C2, b, cap, H, W, out_f32=True)
        g1 = torch.empty(C1 * 9 + C1, device=dev)
        nat.conv1_wgrad_f32(ws.dact1, x, g1, ws.scr_w1, 1, C1, b, cap, H, W)
        module._live = None
        return (None, g1[:C1 * 9].view(C1, 1, 3, 3), g1[C1 * 9:], dw2, db2, dwf, dl.sum(0), None)


class SimpleCNNModule(nn.Module):
    """Drop-in for the reference's ``SimpleCNN`` (reference model.py:4-20)."""

    def __init__(self, native: bool | None = None):
        """``native``: use the sm_100a kernels for CUDA inputs (default: whenever the library is built);
        False forces torch's own ops everywhere (cuDNN/cuBLAS on CUDA — the reference's path)."""
        super().__init__()
        self.net = nn.Sequential(
            nn.Conv2d(1, C1, 3, padding=1),
            nn.ReLU(),
            nn.Conv2d(C1, C2, 3, padding=1),
            nn.ReLU(),
            nn.Flatten(),
        )
        self.fl = nn.Linear(FEATURES, NCLS)
        self._native_wanted = native
        self._nat = None
        self._ws: dict = {}
        self._live = None

    # ------------------------------------------------------------------ native plumbing
    def _use_native(self, x: torch.Tensor) -> bool:
        if self._native_wanted is False or not x.is_cuda or x.shape[1:] != (1, IMG_H, IMG_W):
            return False
        if self._nat is None:
            from ..ops import native
            if not native.available():
                if self._native_wanted:
                    native.lib()  # raises with the build instruction
                return False
            native.lib()
            self._nat = native
        return True

    def _workspace(self, device, batch: int) -> _Workspace:
        ws = self._ws.get(device)
        if ws is None or ws.cap < batch:
            with torch.cuda.device(device):
                ws = _Workspace(self._nat, device, max(batch, 2 * ws.cap if ws else batch))
            self._ws[device] = ws
        return ws

    def _apply(self, fn, *a, **kw):  # .to()/.cuda()/.cpu(): workspaces belong to the old device
        self._ws.clear()
        self._live = None
        return super()._apply(fn, *a, **kw)

    # ------------------------------------------------------------------ forward
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self._use_native(x):
            c1, c2 = self.net[0], self.net[2]
            return _NativeLogits.apply(x, c1.weight, c1.bias, c2.weight, c2.bias, self.fl.weight,
                                       self.fl.bias, self)
        return self.fl(self.net(x))
