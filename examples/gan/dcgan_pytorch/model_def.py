"""DCGAN (64x64 convolutional generator / discriminator) as a two-model, two-optimizer PyTorchTrial."""
from determined_b200.models.vision_trials import DCGANTrial  # noqa: F401
