"""BERT-base GLUE-shape fine-tuning as a PyTorchTrial (see determined_b200/models/bert_glue_trial.py)."""
from determined_b200.models.bert_glue_trial import BertGlueTrial  # noqa: F401
