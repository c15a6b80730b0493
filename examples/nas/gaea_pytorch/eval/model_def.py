"""Evaluation of a searched cell: the DARTS network with the genotype given as hyperparameters (see
determined_b200/models/darts_trial.py; determined_b200.models.gaea_search.genotype_to_hparams turns the genotype the
search trial logs into these keys)."""
from determined_b200.models.darts_trial import DARTSCNNTrial  # noqa: F401
