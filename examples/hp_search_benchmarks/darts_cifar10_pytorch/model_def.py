"""DARTS CIFAR-10 network whose genotype is the hyperparameter space (see determined_b200/models/darts_trial.py)."""
from determined_b200.models.darts_trial import DARTSCNNTrial  # noqa: F401
