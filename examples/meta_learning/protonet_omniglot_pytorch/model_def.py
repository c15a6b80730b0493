"""Prototypical networks on Omniglot-shape episodes as a PyTorchTrial (see determined_b200/models/protonet_trial.py)."""
from determined_b200.models.protonet_trial import OmniglotProtoNetTrial  # noqa: F401
