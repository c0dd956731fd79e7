"""Drop-in for the reference's ``model.py`` (model.py:4-20): ``from model import SimpleCNN`` gives an
``nn.Module`` with the reference's structure, state_dict keys and initialisation, usable with any torch
optimizer / loss / ``.to()`` / DDP — whose CUDA forward and backward run on this repository's sm_100a
kernels (ddp_b200/models/module.py).  The flat-arena models the step engine drives are exported too."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ddp_b200.models.module import SimpleCNNModule as SimpleCNN  # noqa: E402,F401
from ddp_b200.models.simple_cnn import SimpleCNN as SimpleCNNArena, SimpleCNNB200  # noqa: E402,F401
