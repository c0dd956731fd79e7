"""Drop-in for the reference's ``data.py``: ``get_dataloader(batch_size, world_size, rank)``
returns ``(dataloader, sampler)`` (reference data.py:6-27).  Implementation: ddp_b200/utils/data.py."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ddp_b200.utils.data import get_dataloader  # noqa: E402,F401
