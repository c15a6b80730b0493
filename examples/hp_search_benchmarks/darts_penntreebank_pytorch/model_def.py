"""DARTS recurrent-cell language model, architecture given by hyperparameters (see determined_b200/models/darts_rnn.py,
darts_rnn_trial.py)."""
from determined_b200.models.darts_rnn_trial import DARTSRNNTrial  # noqa: F401
