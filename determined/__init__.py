"""Drop-in alias: existing model definitions written against the reference keep their imports
(``import determined as det``, ``from determined.pytorch import PyTorchTrial, DataLoader, ...``)."""
import importlib
import sys

import determined_b200 as _impl
from determined_b200 import *  # noqa: F401,F403
from determined_b200 import __version__  # noqa: F401

for _name in ("pytorch", "experimental", "workload", "storage", "load", "constants", "utils", "ops", "models",
              "parallel", "keras"):
    try:
        _mod = importlib.import_module(f"determined_b200.{_name}")
    except Exception:  # pragma: no cover - optional pieces
        continue
    sys.modules[f"determined.{_name}"] = _mod
    globals()[_name] = _mod
util = utils.util  # noqa: F821  (reference: determined.util)
sys.modules["determined.util"] = util
