"""Checkpoint GC task: delete the checkpoints the master selected (``DET_DELETE`` = JSON list of
``{"uuid", "resources"}``) from the experiment's checkpoint storage.
Parity: reference ``harness/determined/exec/gc_checkpoints.py:15-91``."""
import json
import logging
import os
import sys

from determined_b200 import storage


def main() -> int:
    logging.basicConfig(level=logging.INFO, format="%(levelname)s: %(message)s")
    cfg = json.loads(os.environ["DET_EXPERIMENT_CONFIG"])
    to_delete = json.loads(os.environ.get("DET_DELETE", "[]"))
    dry_run = os.environ.get("DET_GC_DRY_RUN") == "1"
    mgr = storage.build(cfg["checkpoint_storage"])
    failed = 0
    for c in to_delete:
        meta = storage.StorageMetadata(c["uuid"], c.get("resources") or {})
        if dry_run:
            logging.info(f"dry run: would delete checkpoint {meta.storage_id}")
            continue
        try:
            mgr.delete(meta)
            logging.info(f"deleted checkpoint {meta.storage_id}")
        except FileNotFoundError:
            logging.info(f"checkpoint {meta.storage_id} already absent")
        except Exception as e:  # keep going: one bad checkpoint must not block the rest
            failed += 1
            logging.error(f"failed to delete checkpoint {meta.storage_id}: {e!r}")
    return 1 if failed else 0


if __name__ == "__main__":
    sys.exit(main())
