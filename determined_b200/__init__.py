"""determined_b200 — a Blackwell-native training platform with Determined's Trial API, experiment
configuration schema, searchers, single-box master and checkpoint format. See DESIGN.md."""
__version__ = "0.1.0"
