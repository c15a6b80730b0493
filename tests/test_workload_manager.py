"""The harness's workload manager, GC task and storage validation on their own (the intents of the reference's
``harness/tests/test_workload_manager.py``, ``test_gc_harness.py``, ``storage/test_shared_fs.py``)."""
import json
import os
import pathlib
import subprocess
import sys

import pytest

from determined_b200 import _execution, storage
from determined_b200 import workload as wl
from determined_b200.layers._workload_manager import WorkloadManager

ROOT = pathlib.Path(__file__).resolve().parent.parent


def _env(metric="loss"):
    cfg = {"searcher": {"name": "single", "metric": metric, "max_length": {"batches": 10}}, "hyperparameters": {"global_batch_size": 4}}
    env, _ = _execution.make_local_execution_env(True, cfg, hparams={"global_batch_size": 4}, limit_gpus=0)
    return env


def _drive(manager, answer):
    """Play the controller: answer every workload the manager hands down with ``answer(workload, args)``."""
    for w, args, respond in manager:
        respond(answer(w, args))


class _FailingUpload(storage.SharedFSStorageManager):
    def post_store_path(self, storage_id: str, storage_dir: str) -> None:
        raise OSError("bucket unreachable")


def test_checkpoint_upload_failure_surfaces(tmp_path):
    """A checkpoint whose upload fails must fail the workload loudly — no COMPLETED message reaches the master
    (reference test_checkpoint_upload_failure)."""
    reported = []
    stream = iter([(wl.checkpoint_workload(1, total_batches_processed=1), [], reported.append)])
    mgr = WorkloadManager(_env(), stream, _FailingUpload(str(tmp_path)))

    def answer(w, args):
        (pathlib.Path(args[0]) / "state_dict.pth").write_bytes(b"x")
        return {"framework": "torch", "format": "cloudpickle"}

    with pytest.raises(OSError, match="bucket unreachable"):
        _drive(mgr, answer)
    assert reported == []


def test_searcher_metric_must_be_a_scalar_and_present(tmp_path):
    """reference test_reject_nonscalar_searcher_metric (+ the missing-metric message)."""
    def run(vm):
        stream = iter([(wl.validation_workload(1, total_batches_processed=1), [], lambda r: None)])
        _drive(WorkloadManager(_env("loss"), stream, storage.SharedFSStorageManager(str(tmp_path))),
               lambda w, a: {"metrics": {"validation_metrics": vm, "num_inputs": 4}})

    run({"loss": 0.5, "extra": [1, 2, 3]})                 # non-scalars are fine next to a scalar searcher metric
    with pytest.raises(ValueError, match="must be a scalar"):
        run({"loss": [0.1, 0.2]})
    with pytest.raises(KeyError, match="must be one of"):
        run({"accuracy": 0.9})
    with pytest.raises(ValueError, match="NaN"):
        run({"loss": float("nan")})


def test_gc_task_deletes_and_dry_runs(tmp_path):
    """The GC task deletes exactly the checkpoints it is given; with DET_GC_DRY_RUN it only says so (reference
    test_gc_harness.py: test_delete_checkpoints, test_dry_run)."""
    mgr = storage.SharedFSStorageManager(str(tmp_path))
    ids = []
    for i in range(3):
        with mgr.store_path() as (sid, path):
            (pathlib.Path(path) / "state_dict.pth").write_text(str(i))
        ids.append(sid)
    env = dict(os.environ, PYTHONPATH=str(ROOT), DET_EXPERIMENT_CONFIG=json.dumps({"checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path)}}),
               DET_DELETE=json.dumps([{"uuid": ids[0], "resources": {"state_dict.pth": 1}}, {"uuid": ids[2], "resources": {"state_dict.pth": 1}}]))
    dry = subprocess.run([sys.executable, "-m", "determined_b200.exec.gc_checkpoints"], env=dict(env, DET_GC_DRY_RUN="1"), capture_output=True, text=True)
    assert dry.returncode == 0 and "would delete" in dry.stderr and sorted(os.listdir(tmp_path)) == sorted(ids)
    real = subprocess.run([sys.executable, "-m", "determined_b200.exec.gc_checkpoints"], env=env, capture_output=True, text=True)
    assert real.returncode == 0, real.stderr
    assert os.listdir(tmp_path) == [ids[1]]
    again = subprocess.run([sys.executable, "-m", "determined_b200.exec.gc_checkpoints"], env=env, capture_output=True, text=True)
    assert again.returncode == 0 and "already absent" in again.stderr


@pytest.mark.skipif(os.geteuid() == 0, reason="root ignores directory permissions")
def test_validate_rejects_a_read_only_directory(tmp_path):
    ro = tmp_path / "ro"
    ro.mkdir()
    ro.chmod(0o555)
    try:
        with pytest.raises(Exception):
            storage.validate_config({"type": "shared_fs", "host_path": str(ro)})
    finally:
        ro.chmod(0o755)


def test_validate_rejects_a_missing_directory(tmp_path):
    with pytest.raises(Exception):
        storage.validate_config({"type": "shared_fs", "host_path": str(tmp_path / "file-not-dir" / "x"), "propagation": "rprivate"}) if (tmp_path / "file-not-dir").write_text("x") else None
