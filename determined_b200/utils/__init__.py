"""Harness utilities (metrics packaging, JSON encoding, profiling, GPU discovery)."""
from .util import (chunks, debug_mode, get_default_master_address, is_overridden, json_encode, make_metrics,
                   preserve_random_state, sizeof_fmt, wrap_metrics, write_user_code)

__all__ = ["chunks", "debug_mode", "get_default_master_address", "is_overridden", "json_encode", "make_metrics",
           "preserve_random_state", "sizeof_fmt", "wrap_metrics", "write_user_code"]
