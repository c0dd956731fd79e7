"""Widened CNN (~25 M parameters, 3x32x32 input) for BASELINE.json config 4.

This configuration has NO counterpart in the reference (its only model is SimpleCNN,
reference model.py:4-20); it exists to exercise what SimpleCNN's 2 MB of gradients cannot:
several large buckets in flight, the two-shot all-reduce regime and overlap with a long backward.
Its forward/backward run through torch ops (cuDNN/cuBLAS) — it is a communication/optimizer
benchmark, not a kernel one; the data-parallel machinery around it (parallel/bucketed.py) is ours.

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


def param_count(m: nn.Module) -> int:
    return sum(p.numel() for p in m.parameters())
