"""Importable alias for the package directory `pytorch-distributed-data-parallel-ddp-trainer_b200/`.

The framework's source lives in the directory named after the project (hyphens and all), which
Python cannot import by name.  This thin package points its ``__path__`` at that directory so
``import ddp_b200.models`` etc. resolve to the real files; nothing else lives here.
"""
import os as _os

_REAL = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                      "pytorch-distributed-data-parallel-ddp-trainer_b200")
__path__.insert(0, _REAL)

with open(_os.path.join(_REAL, "__init__.py")) as _f:
    exec(compile(_f.read(), _os.path.join(_REAL, "__init__.py"), "exec"))
del _f
