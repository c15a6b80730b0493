"""Faster R-CNN fine-tuning with a stock torchvision detector (see determined_b200/models/detection_trial.py)."""
from determined_b200.models.detection_trial import ObjectDetectionTrial  # noqa: F401
