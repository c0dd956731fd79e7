"""``LinearB200`` and ``ReLUMaxPool2dB200``: ``nn.Module`` layers of any shape over the native sm_100a
kernels of csrc/gemm_tc.cu.

``LinearB200`` is an ``nn.Linear`` (same parameters, same ``state_dict`` keys, same default
initialisation) whose CUDA forward/backward are — for layers with more than 16 outputs; a classifier head
runs on exact-fp32 CUDA-core kernels instead — three tcgen05 GEMMs over operands packed once per step
as split row planes (x = hi + lo bf16, "bf16x3": above TF32 operand precision, fp32 accumulation) —
what the reference gets from cuBLAS through ``nn.Linear`` (reference model.py:16, train_ddp.py:199),
here for shapes beyond the 10-class head the step engine hard-codes: BASELINE.json config 4's
fc 32768 -> 750 -> 10.  An optional fused ReLU runs in the split-K reduction of the forward and in the
pack of the incoming gradient of the backward.

``ReLUMaxPool2dB200`` = ``nn.ReLU`` -> ``nn.MaxPool2d(2)`` in one pass each way (one byte per output
instead of int64 indices).  On CPU tensors (or without the built library) both fall back to the torch ops.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def _native():
    from ..ops import native
    return native if native.available() else None


class _LinearWorkspace:
    def __init__(self, nat, device, cap: int, in_f: int, out_f: int):
        self.cap = cap
        self.x = nat.PackedMatrix(cap, in_f, device)
        self.w = nat.PackedMatrix(out_f, in_f, device)
        self.dy = nat.PackedMatrix(cap, out_f, device)
        self.partial = None

    def scratch(self, nfloats: int, device) -> torch.Tensor:
        if self.partial is None or self.partial.numel() < nfloats:
            self.partial = torch.empty(nfloats, dtype=torch.float32, device=device)
        return self.partial


def _gemm(nat, ws, a, a_mn, b, b_mn, M, N, K, out, transposed, bias=None, relu=False):
    """out = A x B^T (+bias, ReLU) with as many K splits as fill the GPU; out is [N, M] when transposed."""
    rows, cols = (N, M) if transposed else (M, N)
    assert out.shape == (rows, cols) and out.is_contiguous()
    splits = nat.gemm_suggest_splits(M, N, K)
    if splits == 1 and bias is None and not relu:
        nat.gemm_split(a, a_mn, b, b_mn, M, N, K, out, cols, transposed)
        return
    part = ws.scratch(splits * rows * cols, out.device)
    nat.gemm_split(a, a_mn, b, b_mn, M, N, K, part, cols, transposed, splits, rows * cols)
    nat.splitk_reduce(part, splits, rows * cols, rows, cols, cols, bias, relu, out)


class _NativeLinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, layer):
        nat = layer._nat
        batch, in_f = x.shape
        out_f = weight.shape[0]
        ws = layer._workspace(x.device, batch)
        x = x.detach().contiguous().float()
        w = weight.detach().contiguous()
        ws.x.pack(x)
        if not layer._w_packed_valid:     # the weights may have changed since the last call
            ws.w.pack(w)
        layer._w_packed_valid = False     # one forward per prepack(): whoever updates the weights renews it
        y = torch.empty(batch, out_f, dtype=torch.float32, device=x.device)
        # y^T[o][b] = sum_k W[o][k] x[b][k]: A = W (K-major), B = x (K-major), stored transposed = y[b][o]
        _gemm(nat, ws, ws.w, False, ws.x, False, out_f, batch, in_f, y, True,
              None if bias is None else bias.detach(), layer.fused_relu)
        ctx.layer, ctx.ws, ctx.shape = layer, ws, (batch, in_f, out_f)
        ctx.has_bias = bias is not None
        ctx.save_for_backward(y if layer.fused_relu else None)
        layer._live = ctx
        return y

    @staticmethod
    def backward(ctx, dy):
        layer, ws = ctx.layer, ctx.ws
        if layer._live is not ctx:
            raise RuntimeError("LinearB200 keeps the packed input of its most recent forward only: "
                               "call backward() before the next forward() on the same layer")
        nat = layer._nat
        batch, in_f, out_f = ctx.shape
        (y,) = ctx.saved_tensors
        dy = dy.contiguous().float()
        ws.dy.pack(dy, y)                 # ReLU' (mask = forward output > 0) is applied while packing
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            # dx^T[k][b] = sum_o W[o][k] dy[b][o]: A = W (MN-major), B = dy (K-major), stored as dx[b][k]
            dx = torch.empty(batch, in_f, dtype=torch.float32, device=dy.device)
            _gemm(nat, ws, ws.w, True, ws.dy, False, in_f, batch, out_f, dx, True)
        if ctx.needs_input_grad[1]:
            # dW[o][k] = sum_b dy[b][o] x[b][k]: both operands MN-major (reduce over the batch rows)
            dst = layer.weight_grad_out
            if dst is not None and dst.shape == (out_f, in_f) and dst.is_contiguous() and dst.device == dy.device:
                dw = dst.view(out_f, in_f)  # a fresh alias: autograd adopts it as .grad without a copy
            else:
                dw = torch.empty(out_f, in_f, dtype=torch.float32, device=dy.device)
            _gemm(nat, ws, ws.dy, True, ws.x, True, out_f, in_f, batch, dw, False)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = torch.empty(out_f, dtype=torch.float32, device=dy.device)
            nat.colsum(dy, y, db)
        layer._live = None
        return dx, dw, db, None


class _SmallLinearFn(torch.autograd.Function):
    """A layer with at most 16 outputs (a classifier head): exact fp32 CUDA-core kernels, one launch forward
    and three backward, no packing (csrc/wide_ops.cu)."""

    @staticmethod
    def forward(ctx, x, weight, bias, layer):
        nat = layer._nat
        x = x.detach().contiguous().float()
        w = weight.detach().contiguous()
        y = torch.empty(x.shape[0], w.shape[0], dtype=torch.float32, device=x.device)
        nat.small_linear_fwd(x, w, None if bias is None else bias.detach(), y)
        ctx.layer, ctx.has_bias = layer, bias is not None
        ctx.save_for_backward(x, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        layer, nat = ctx.layer, ctx.layer._nat
        x, w = ctx.saved_tensors
        dy = dy.contiguous().float()
        out_f, in_f = w.shape
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        dst = layer.weight_grad_out
        if dst is not None and dst.shape == (out_f, in_f) and dst.is_contiguous() and dst.device == dy.device:
            dw = dst.view(out_f, in_f)
        else:
            dw = torch.empty(out_f, in_f, dtype=torch.float32, device=dy.device)
        db = torch.empty(out_f, dtype=torch.float32, device=dy.device) if ctx.has_bias else None
        nat.small_linear_bwd(dy, x, w, dx, dw, db)
        return dx, dw, db, None


class LinearB200(nn.Linear):
    """Drop-in ``nn.Linear`` (+ optional fused ReLU) on the tcgen05 GEMM of csrc/gemm_tc.cu."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, fused_relu: bool = False,
                 native: bool | None = None):
        super().__init__(in_features, out_features, bias=bias)
        self.fused_relu = fused_relu
        #: optional fp32 [out, in] buffer the native backward writes the weight gradient into (and returns
        #: as the gradient): a data-parallel engine points it at its flat gradient arena.  The buffer is
        #: OVERWRITTEN by every backward, so whoever sets it must also reset ``weight.grad`` to None before each
        #: backward (autograd then adopts the buffer as ``.grad``); it is not compatible with accumulating
        #: gradients over several backward passes.
        self.weight_grad_out = None
        self._native_wanted = native
        self._nat = None
        self._ws: dict = {}
        self._live = None
        self._w_packed_valid = False

    def prepack(self, stream=None) -> bool:
        """Pack the CURRENT weights for the next forward now (on ``stream``), instead of at the start of
        that forward: an engine that has just updated the weights calls this where the pack overlaps other
        work.  Valid for exactly one forward.  Returns False (and does nothing) before the first native
        forward, when no workspace exists yet."""
        if self._nat is None or not self.weight.is_cuda or (self.out_features <= 16 and not self.fused_relu):
            return False
        ws = self._ws.get(self.weight.device)
        if ws is None:
            return False
        ws.w.pack(self.weight.detach(), stream=stream)
        self._w_packed_valid = True
        return True

    def _use_native(self, x: torch.Tensor) -> bool:
        if self._native_wanted is False or not x.is_cuda or x.dim() != 2 or self.weight.dtype != torch.float32:
            return False
        if self._nat is None:
            nat = _native()
            if nat is None:
                if self._native_wanted:
                    from ..ops import native
                    native.lib()  # raises with the build instruction
                return False
            nat.lib()
            self._nat = nat
        return True

    def _workspace(self, device, batch: int) -> _LinearWorkspace:
        ws = self._ws.get(device)
        if ws is None or ws.cap < batch:
            with torch.cuda.device(device):
                ws = _LinearWorkspace(self._nat, device, batch, self.in_features, self.out_features)
            self._ws[device] = ws
        return ws

    def _apply(self, fn, *a, **kw):  # .to()/.cuda()/.cpu(): workspaces belong to the old device
        self._ws.clear()
        self._live = None
        self._w_packed_valid = False
        return super()._apply(fn, *a, **kw)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self._use_native(x):
            if self.out_features <= 16 and not self.fused_relu:
                return _SmallLinearFn.apply(x, self.weight, self.bias, self)
            return _NativeLinearFn.apply(x, self.weight, self.bias, self)
        y = F.linear(x, self.weight, self.bias)
        return F.relu(y) if self.fused_relu else y

    def extra_repr(self) -> str:
        return super().extra_repr() + f", fused_relu={self.fused_relu}"


class _ReLUMaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, nat):
        x = x.contiguous().float()
        b, c, h, w = x.shape
        out = torch.empty(b, c, h // 2, w // 2, dtype=torch.float32, device=x.device)
        code = torch.empty(b, c, h // 2, w // 2, dtype=torch.uint8, device=x.device)
        nat.relu_maxpool2_fwd(x, None, out, code)
        ctx.nat, ctx.in_shape = nat, x.shape
        ctx.save_for_backward(code)
        return out

    @staticmethod
    def backward(ctx, dout):
        (code,) = ctx.saved_tensors
        din = torch.empty(ctx.in_shape, dtype=torch.float32, device=dout.device)
        ctx.nat.relu_maxpool2_bwd(dout.contiguous().float(), code, din)
        return din, None


class ReLUMaxPool2dB200(nn.Module):
    """``nn.ReLU`` followed by ``nn.MaxPool2d(2)`` as one native pass (H even, W a multiple of 4)."""

    def __init__(self, native: bool | None = None):
        super().__init__()
        self._native_wanted = native

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        nat = _native() if (self._native_wanted is not False and x.is_cuda and x.dim() == 4
                            and x.shape[2] % 2 == 0 and x.shape[3] % 4 == 0) else None
        if nat is None:
            return F.max_pool2d(F.relu(x), 2)
        nat.lib()
        return _ReLUMaxPoolFn.apply(x, nat)
