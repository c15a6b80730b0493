"""MNIST classifier whose data sets go through the data-layer cache (see determined_b200/models/classic_trials.py)."""
from determined_b200.models.classic_trials import DataLayerMNistTrial  # noqa: F401
