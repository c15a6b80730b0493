"""Gradient-boosted trees on Titanic-shaped data through the Trial API (see determined_b200/models/boosted_trees.py)."""
from determined_b200.models.boosted_trees import GBTTitanicTrial  # noqa: F401
