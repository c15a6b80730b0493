"""Kill the master in the middle of an experiment, start a new one on the same data directory: the experiment is
restored from the journal (searcher events replayed, trials rolled back to their last checkpoint) and runs to
completion. Mirrors the reference's restore path (master/internal/core.go:220 restoreExperiment, experiment.go:238-294,
trial.go:1008-1033) which its e2e suite exercises by restarting the master container."""
import os
import pathlib
import signal
import subprocess
import sys
import time

from determined_b200.experimental.client import Determined

ROOT = pathlib.Path(__file__).resolve().parent.parent
MASTER = ROOT / "determined_b200" / "_lib" / "det-master"
MODEL_DIR = str(ROOT / "tests" / "fixtures" / "xor_model_def")


def start_master(tmp, port=0):
    port_file = tmp / f"port-{time.time_ns()}"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg"))
    proc = subprocess.Popen([str(MASTER), "--port", str(port), "--data-dir", str(tmp / "data"), "--artificial-slots", "2", "--repo-root", str(ROOT),
                             "--python", sys.executable, "--print-port-file", str(port_file)],
                            stdout=open(tmp / "master.log", "a"), stderr=subprocess.STDOUT, env=env, start_new_session=True)
    for _ in range(200):
        if port_file.exists() and port_file.read_text().strip():
            break
        time.sleep(0.05)
    return proc, int(port_file.read_text())


def test_experiment_survives_master_restart(tmp_path):
    from determined_b200 import _build

    _build.build_core()
    os.environ["DET_CONFIG_DIR"] = str(tmp_path / "cfg")
    proc, port = start_master(tmp_path)
    try:
        d = Determined(f"127.0.0.1:{port}")
        cfg = {"entrypoint": "model_def:XorTrial", "description": "restore",
               "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
               "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 400}}, "scheduling_unit": 10,
               "min_checkpoint_period": {"batches": 20},
               "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpts")}, "max_restarts": 3}
        exp = d.create_experiment(cfg, MODEL_DIR)
        exp_id = exp.id
        # let it make some progress (at least one checkpoint), then take the master down hard
        for _ in range(600):
            t = exp.describe()["trials"]
            if t and t[0].get("total_batches_processed", 0) >= 60:
                break
            time.sleep(0.1)
        assert t and t[0]["total_batches_processed"] >= 60
        os.killpg(proc.pid, signal.SIGKILL)     # master and the harness it spawned
        proc.wait(timeout=10)
        before = t[0]["total_batches_processed"]
    finally:
        if proc.poll() is None:
            os.killpg(proc.pid, signal.SIGKILL)
    proc2, port2 = start_master(tmp_path, port)
    try:
        d2 = Determined(f"127.0.0.1:{port2}")
        exp2 = d2.get_experiment(exp_id)
        assert exp2.describe()["state"] in ("ACTIVE", "COMPLETED")
        assert exp2.wait(timeout=240) == "COMPLETED"
        desc = exp2.describe()
        tr = desc["trials"][0]
        assert tr["state"] == "COMPLETED" and tr["total_batches_processed"] == 400 and desc["progress"] == 1.0
        assert before < 400
    finally:
        os.killpg(proc2.pid, signal.SIGTERM)
        try:
            proc2.wait(timeout=5)
        except Exception:
            os.killpg(proc2.pid, signal.SIGKILL)
