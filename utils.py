"""Drop-in for the reference's ``utils.py``: ``setup(rank, world_size, backend=None)`` and
``cleanup()`` (reference utils.py:5-19).  Implementation: ddp_b200/parallel/process_group.py."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from ddp_b200.parallel.process_group import cleanup, setup  # noqa: E402,F401
