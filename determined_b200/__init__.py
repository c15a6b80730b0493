"""determined_b200 — a Blackwell-native training platform with Determined's Trial API, experiment
configuration schema, searchers, single-box master and checkpoint format. See DESIGN.md.

``import determined_b200 as det`` mirrors ``import determined as det`` of the reference
(``harness/determined/__init__.py``): ``det.pytorch``, ``det.experimental``, ``det.TrialContext``, hyperparameter
helpers ``det.Constant/Integer/Double/Categorical/Log``.
"""
from ._env import DistributedContext, EnvContext, ExperimentConfig, RendezvousInfo
from ._execution import InvalidHP
from ._hparam import Categorical, Constant, Double, Integer, Log
from . import errors
from ._train_context import NativeContext, TrialContext
from ._trial import Trial
from ._trial_controller import CallbackTrialController, LoopTrialController, TrialController
from .utils import util

__version__ = "0.13.10.dev0+b200.1"

__all__ = ["DistributedContext", "EnvContext", "ExperimentConfig", "RendezvousInfo", "InvalidHP", "Categorical",
           "Constant", "Double", "Integer", "Log", "NativeContext", "TrialContext", "Trial", "TrialController",
           "CallbackTrialController", "LoopTrialController", "errors", "__version__"]
