"""MNIST GAN (MLP generator / discriminator) as a two-model, two-optimizer PyTorchTrial."""
from determined_b200.models.vision_trials import GANTrial  # noqa: F401
