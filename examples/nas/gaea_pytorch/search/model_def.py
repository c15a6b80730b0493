"""GAEA architecture search over the DARTS cell space (see determined_b200/models/gaea_search.py)."""
from determined_b200.models.gaea_search import GAEASearchTrial  # noqa: F401
