"""DETR on the synthetic (`backend: fake`) COCO-shape back-end (see determined_b200/models/detr.py, detr_trial.py)."""
from determined_b200.models.detr_trial import DETRTrial  # noqa: F401
