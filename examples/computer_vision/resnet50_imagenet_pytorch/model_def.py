"""ResNet-50 on ImageNet-shape batches as a PyTorchTrial (see determined_b200/models/vision_trials.py)."""
from determined_b200.models.vision_trials import ResNet50Trial  # noqa: F401
