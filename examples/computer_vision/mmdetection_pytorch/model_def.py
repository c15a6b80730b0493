"""Any mmdetection detector through PyTorchTrial (needs the mmdet / mmcv packages; see determined_b200/models/mmdet_trial.py)."""
from determined_b200.models.mmdet_trial import MMDetTrial  # noqa: F401
