"""``det.experimental`` — native API and the Python SDK (reference: ``harness/determined/experimental`` and
``common/determined_common/experimental``)."""
from ._native import create, create_trial_instance, init_native, test_one_batch

__all__ = ["create", "create_trial_instance", "init_native", "test_one_batch", "Determined", "Checkpoint"]


def __getattr__(name):  # lazy: the SDK client pulls in HTTP machinery only when used
    if name in ("Determined", "Checkpoint", "ExperimentReference", "TrialReference", "Model", "ModelSortBy", "ModelOrderBy",
                "ModelFramework", "CheckpointState"):
        from . import client

        return getattr(client, name)
    raise AttributeError(name)
