"""ALBERT SQuAD-shape (seq 384) fine-tuning as a PyTorchTrial (see determined_b200/models/squad_trials.py)."""
from determined_b200.models.squad_trials import AlbertSquadTrial  # noqa: F401
