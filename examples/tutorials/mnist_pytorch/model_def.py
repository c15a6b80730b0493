"""MNIST tutorial CNN as a PyTorchTrial (see determined_b200/models/vision_trials.py for the code)."""
from determined_b200.models.vision_trials import MNistTrial  # noqa: F401
