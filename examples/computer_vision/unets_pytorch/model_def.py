"""U-Net image segmentation as a PyTorchTrial (see determined_b200/models/classic_trials.py for the code)."""
from determined_b200.models.classic_trials import UNetTrial  # noqa: F401
