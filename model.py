"""Drop-in for the reference's ``model.py`` (model.py:4-20): ``from model import SimpleCNN`` keeps working.
The implementation lives in ddp_b200/models/simple_cnn.py (flat-arena CPU model + sm_100a model)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ddp_b200.models.simple_cnn import SimpleCNN, SimpleCNNB200  # noqa: E402,F401
