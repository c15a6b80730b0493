"""BERT-base SQuAD-shape (seq 384) fine-tuning as a PyTorchTrial (see determined_b200/models/squad_trials.py)."""
from determined_b200.models.squad_trials import BertSquadTrial  # noqa: F401
