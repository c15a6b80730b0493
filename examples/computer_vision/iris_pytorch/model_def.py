"""Iris species classifier as a PyTorchTrial (see determined_b200/models/classic_trials.py for the code)."""
from determined_b200.models.classic_trials import IrisTrial  # noqa: F401
