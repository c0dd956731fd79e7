"""Widened CNN (~25 M parameters, 3x32x32 input) for BASELINE.json config 4.

This configuration has NO counterpart in the reference (its only model is SimpleCNN,
reference model.py:4-20); it exists to exercise what SimpleCNN's 2 MB of gradients cannot:
several large buckets in flight, the two-shot all-reduce regime and overlap with a long backward.
``WideCNN`` runs through torch ops (cuDNN/cuBLAS): the comparison arm.  ``WideCNNB200`` is the same
network (same ``state_dict`` keys and initialisation) with the fully connected layers — 99.7 % of the
parameters — on the tcgen05 GEMM of csrc/gemm_tc.cu and ReLU + max-pool as one native pass; its two
convolutions still go through cuDNN (the conv kernels of csrc/conv_tc.cu are specialised for the
reference's 28x28 / 32->64 layer).  The data-parallel machinery around both
(parallel/bucketed.py) is ours.

    conv(3->64,3x3,p1)+ReLU -> conv(64->128,3x3,p1)+ReLU -> maxpool2 -> fc(32768->750)+ReLU -> fc(750->10)
"""
import torch.nn as nn


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


class WideCNNB200(nn.Module):
    """``WideCNN`` with native fully connected layers and a fused ReLU + max-pool (models/linear.py)."""

    def __init__(self, hidden: int = 750, num_classes: int = 10, native: bool | None = None):
        super().__init__()
        from .linear import LinearB200, ReLUMaxPool2dB200
        # indices 0/2 of both Sequentials hold the parameters, as in WideCNN
        self.features = nn.Sequential(
            nn.Conv2d(3, 64, 3, padding=1), nn.ReLU(inplace=True),
            nn.Conv2d(64, 128, 3, padding=1), ReLUMaxPool2dB200(native=native), nn.Flatten())
        self.classifier = nn.Sequential(LinearB200(128 * 16 * 16, hidden, fused_relu=True, native=native),
                                        nn.Identity(),
                                        LinearB200(hidden, num_classes, native=native))

    def forward(self, x):
        return self.classifier(self.features(x))


def param_count(m: nn.Module) -> int:
    return sum(p.numel() for p in m.parameters())
