"""Widened CNN (~25 M parameters, 3x32x32 input) for BASELINE.json config 4.

This configuration has NO counterpart in the reference (its only model is SimpleCNN,
reference model.py:4-20); it exists to exercise what SimpleCNN's 2 MB of gradients cannot:
several large buckets in flight, the two-shot all-reduce regime and overlap with a long backward.
``WideCNN`` runs through torch ops (cuDNN/cuBLAS): the comparison arm.  ``WideCNNB200`` is the same
network (same ``state_dict`` keys, values and initialisation) on the native kernels: conv1 and the pooling
as channels-last fp32 kernels (csrc/wide_ops.cu), the fully connected layers — 99.7 % of the parameters —
on the tcgen05 GEMM of csrc/gemm_tc.cu.  Its second convolution (64 -> 128 at 32x32) still goes through
cuDNN: the tensor-core conv kernels of csrc/conv_tc.cu keep a layer's whole weight image resident in
shared memory and are specialised for the reference's 28x28 / 32->64 layer.  The
data-parallel machinery around both (parallel/bucketed.py) is ours.

    conv(3->64,3x3,p1)+ReLU -> conv(64->128,3x3,p1)+ReLU -> maxpool2 -> fc(32768->750)+ReLU -> fc(750->10)
"""
import torch
import torch.nn as nn
import torch.nn.functional as F


class WideCNN(nn.Module):
    def __init__(self, hidden: int = 750, num_classes: int = 10):
        super().__init__()
        self.features = nn.Sequential(
            nn.Conv2d(3, 64, 3, padding=1), nn.ReLU(inplace=True),
            nn.Conv2d(64, 128, 3, padding=1), nn.ReLU(inplace=True),
            nn.MaxPool2d(2), nn.Flatten())
        self.classifier = nn.Sequential(nn.Linear(128 * 16 * 16, hidden), nn.ReLU(inplace=True),
                                        nn.Linear(hidden, num_classes))

    def forward(self, x):
        return self.classifier(self.features(x))


class _Conv1Fn(torch.autograd.Function):
    """relu(conv1(x) + b) : NCHW image -> NHWC activation, fp32 FMA kernel (csrc/wide_ops.cu)."""

    @staticmethod
    def forward(ctx, x, w, b, nat):
        x = x.contiguous().float()
        bsz, _, H, W = x.shape
        out = torch.empty(bsz, H, W, w.shape[0], dtype=torch.float32, device=x.device)
        nat.conv3x3_small_cin_fwd(x, w.detach().contiguous(), b.detach(), out)
        ctx.nat, ctx.cout = nat, w.shape[0]
        ctx.save_for_backward(x, out)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, out = ctx.saved_tensors
        dw, db = ctx.nat.conv3x3_small_cin_wgrad(dout.contiguous(), out, x, ctx.cout)
        return None, dw, db, None


class _PoolNhwcFn(torch.autograd.Function):
    """relu(maxpool2(y) + bias) on an NHWC tensor; the bias gradient is a masked column sum of dout."""

    @staticmethod
    def forward(ctx, y, bias, nat):
        y = y.contiguous()
        bsz, H, W, C = y.shape
        out = torch.empty(bsz, H // 2, W // 2, C, dtype=torch.float32, device=y.device)
        code = torch.empty(bsz, H // 2, W // 2, C, dtype=torch.uint8, device=y.device)
        nat.relu_maxpool2_nhwc_fwd(y, bias.detach(), out, code)
        ctx.nat, ctx.in_shape = nat, y.shape
        ctx.save_for_backward(code, out)
        return out

    @staticmethod
    def backward(ctx, dout):
        code, out = ctx.saved_tensors
        dout = dout.contiguous().float()
        C = out.shape[-1]
        din = torch.empty(ctx.in_shape, dtype=torch.float32, device=dout.device)
        ctx.nat.relu_maxpool2_nhwc_bwd(dout, code, din)
        db = torch.empty(C, dtype=torch.float32, device=dout.device)
        ctx.nat.colsum(dout.view(-1, C), out.view(-1, C), db)
        return din, db, None


class WideCNNB200(nn.Module):
    """``WideCNN`` on the native kernels, channels-last end to end (same ``state_dict`` as ``WideCNN``).

    conv1 + bias + ReLU: fp32 FMA kernel writing NHWC; conv2: cuDNN on the channels-last tensors (no layout
    transposes, no bias pass); bias + ReLU + max-pool: one native NHWC pass, its backward also yields
    conv2's bias gradient; fc1 (+ReLU) and fc2: tcgen05 GEMMs (models/linear.py).  The activation reaches
    fc1 flattened in (h, w, c) order, so ``classifier.0.weight`` is STORED with its columns in that order;
    ``state_dict()`` / ``load_state_dict()`` convert to and from the (c, h, w) order of ``WideCNN``.
    """

    C2, P = 128, 16  # channels and spatial size of the pooled activation

    def __init__(self, hidden: int = 750, num_classes: int = 10, native: bool | None = None):
        super().__init__()
        from .linear import LinearB200
        # indices 0/2 of both Sequentials hold the parameters, as in WideCNN
        self.features = nn.Sequential(nn.Conv2d(3, 64, 3, padding=1), nn.Identity(),
                                      nn.Conv2d(64, self.C2, 3, padding=1))
        self.classifier = nn.Sequential(LinearB200(self.C2 * self.P * self.P, hidden, fused_relu=True,
                                                   native=native),
                                        nn.Identity(),
                                        LinearB200(hidden, num_classes, native=native))
        with torch.no_grad():  # same seed => same function as WideCNN
            w = self.classifier[0].weight
            w.copy_(self._cols_to_hwc(w))
        self._native_wanted = native
        self._register_state_dict_hook(self._export_hook)
        self._register_load_state_dict_pre_hook(self._import_hook)

    # fc1 weight columns: (c, h, w) order of the reference-style flatten <-> (h, w, c) order stored here
    def _cols_to_hwc(self, w):
        o = w.shape[0]
        return w.reshape(o, self.C2, self.P * self.P).transpose(1, 2).reshape(o, -1)

    def _cols_to_chw(self, w):
        o = w.shape[0]
        return w.reshape(o, self.P * self.P, self.C2).transpose(1, 2).reshape(o, -1)

    @staticmethod
    def _export_hook(module, state_dict, prefix, local_metadata):
        key = prefix + "classifier.0.weight"
        if key in state_dict:
            state_dict[key] = module._cols_to_chw(state_dict[key]).contiguous()

    def _import_hook(self, state_dict, prefix, *args):
        key = prefix + "classifier.0.weight"
        if key in state_dict:
            state_dict[key] = self._cols_to_hwc(state_dict[key]).contiguous()

    def _nat(self, x):
        if self._native_wanted is False or not x.is_cuda or x.shape[1:] != (3, 32, 32):
            return None
        from ..ops import native
        if not native.available():
            if self._native_wanted:
                native.lib()  # raises with the build instruction
            return None
        native.lib()
        return native

    def forward(self, x):
        c1, c2 = self.features[0], self.features[2]
        nat = self._nat(x)
        if nat is not None:
            a1 = _Conv1Fn.apply(x, c1.weight, c1.bias, nat)                      # [b, 32, 32, 64] NHWC
            y2 = F.conv2d(a1.permute(0, 3, 1, 2), c2.weight, None, padding=1)    # channels-last in and out
            p = _PoolNhwcFn.apply(y2.permute(0, 2, 3, 1), c2.bias, nat)          # [b, 16, 16, 128]
        else:
            y2 = c2(F.relu(c1(x)))
            p = F.max_pool2d(F.relu(y2), 2).permute(0, 2, 3, 1)
        return self.classifier(p.reshape(p.shape[0], -1))


def param_count(m: nn.Module) -> int:
    return sum(p.numel() for p in m.parameters())
