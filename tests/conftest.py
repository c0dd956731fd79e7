"""pytest configuration: the `gpu` marker and import path.

`python -m pytest tests -m "not gpu"` must pass on a CPU-only box (gloo, world_size 2);
`python -m pytest tests -m gpu` runs on a B200.  GPU tests are skipped, not failed, when no
CUDA device is visible so a plain `pytest tests` works everywhere.
"""
import os
import sys

import pytest

# the boxes the tests run on have no network: skip torchvision's download attempt (tests that want it
# clear the variable themselves)
os.environ.setdefault("DDP_B200_NO_DOWNLOAD", "1")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_cuda = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_cuda = False
    if has_cuda:
        return
    skip = pytest.mark.skip(reason="no CUDA device visible")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)
