"""CIFAR-10 CNN as a PyTorchTrial (see determined_b200/models/vision_trials.py for the code)."""
from determined_b200.models.vision_trials import CIFARTrial  # noqa: F401
