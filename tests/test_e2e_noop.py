"""Experiment lifecycle end to end with trials that do no compute (the reference's e2e_tests/tests/experiment/test_noop.py
cases): warm start from a trial / a checkpoint, RNG state restored from a checkpoint, pausing an experiment that never got
resources, cancelling many active and paused experiments at once, a step that outlasts the harness's long poll."""
import copy
import os
import pathlib
import subprocess
import sys
import time

import pytest

from determined_b200.experimental.client import Determined

ROOT = pathlib.Path(__file__).resolve().parent.parent
MASTER = ROOT / "determined_b200" / "_lib" / "det-master"
NOOP_DIR = str(ROOT / "tests" / "fixtures" / "no_op")
PT_NOOP_DIR = str(ROOT / "tests" / "fixtures" / "pytorch_no_op")


@pytest.fixture(scope="module")
def master(tmp_path_factory):
    from determined_b200 import _build

    _build.build_core()
    tmp = tmp_path_factory.mktemp("noopmaster")
    port_file = tmp / "port"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg"))
    proc = subprocess.Popen([str(MASTER), "--port", "0", "--data-dir", str(tmp / "data"), "--artificial-slots", "4", "--repo-root", str(ROOT),
                             "--python", sys.executable, "--print-port-file", str(port_file)],
                            stdout=open(tmp / "master.log", "w"), stderr=subprocess.STDOUT, env=env)
    for _ in range(100):
        if port_file.exists() and port_file.read_text().strip():
            break
        time.sleep(0.05)
    port = int(port_file.read_text())
    os.environ["DET_CONFIG_DIR"] = str(tmp / "cfg")
    yield {"url": f"127.0.0.1:{port}", "tmp": tmp}
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()


def noop_config(tmp, batches=30, **hp):
    return {"entrypoint": "model_def:NoOpTrial", "description": "noop",
            "hyperparameters": dict({"global_batch_size": 32, "metrics_progression": "decreasing", "metrics_base": 0.9, "metrics_sigma": 0}, **hp),
            "searcher": {"name": "single", "metric": "validation_error", "max_length": {"batches": batches}},
            "scheduling_unit": 10, "min_checkpoint_period": {"batches": 10}, "min_validation_period": {"batches": 10},
            "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp / "ckpts"), "save_trial_latest": 10},
            "max_restarts": 0}


def only_trial(exp):
    trials = exp.describe(metrics=True)["trials"]
    assert len(trials) == 1
    return trials[0]


def completed_checkpoints(d, trial_id):
    """(total_batches, uuid) of the trial's completed checkpoints, oldest first."""
    cks = [c for c in d.get_trial(trial_id).describe()["checkpoints"] if c["state"] == "COMPLETED"]
    return sorted((c["total_batches"], c["uuid"]) for c in cks)


def val_error(trial, i=-1):
    return trial["validations"][i]["metrics"]["validation_metrics"]["validation_error"]


def test_warm_start_from_trial_and_from_checkpoint(master):
    d = Determined(master["url"])
    cfg = noop_config(master["tmp"])
    first = d.create_experiment(cfg, NOOP_DIR)
    assert first.wait(timeout=120) == "COMPLETED"
    t1 = only_trial(first)
    assert len(t1["steps"]) == 3 and val_error(t1) == pytest.approx(0.9 ** 3)
    cks = completed_checkpoints(d, t1["id"])
    assert [c[0] for c in cks] == [10, 20, 30]

    # source_trial_id: continue from the source trial's LATEST checkpoint
    cfg2 = copy.deepcopy(cfg)
    cfg2["searcher"]["source_trial_id"] = t1["id"]
    second = d.create_experiment(cfg2, NOOP_DIR)
    assert second.wait(timeout=120) == "COMPLETED"
    t2 = only_trial(second)
    assert t2["warm_start_checkpoint"]["uuid"] == cks[-1][1]
    assert len(t2["steps"]) == 3 and val_error(t2) == pytest.approx(0.9 ** 6)

    # source_checkpoint_uuid: continue from exactly that checkpoint
    cfg3 = copy.deepcopy(cfg)
    cfg3["searcher"]["source_checkpoint_uuid"] = cks[0][1]
    third = d.create_experiment(cfg3, NOOP_DIR)
    assert third.wait(timeout=120) == "COMPLETED"
    t3 = only_trial(third)
    assert t3["warm_start_checkpoint"]["uuid"] == cks[0][1]
    assert len(t3["steps"]) == 3 and val_error(t3) == pytest.approx(0.9 ** 4)


def test_pytorch_cpu_rng_state_is_restored_from_a_checkpoint(master):
    d = Determined(master["url"])
    cfg = {"entrypoint": "model_def:NoopPyTorchTrial", "description": "rng", "hyperparameters": {"global_batch_size": 8},
           "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 3}},
           "scheduling_unit": 1, "min_checkpoint_period": {"batches": 1}, "min_validation_period": {"batches": 1},
           "checkpoint_storage": {"type": "shared_fs", "host_path": str(master["tmp"] / "ckpts"), "save_trial_latest": 10},
           "reproducibility": {"experiment_seed": 7}, "max_restarts": 0}
    first = d.create_experiment(cfg, PT_NOOP_DIR)
    assert first.wait(timeout=180) == "COMPLETED"
    t1 = only_trial(first)
    assert len(t1["validations"]) >= 3
    cks = completed_checkpoints(d, t1["id"])
    assert cks[0][0] == 1
    cfg2 = copy.deepcopy(cfg)
    cfg2["searcher"]["source_checkpoint_uuid"] = cks[0][1]
    cfg2["reproducibility"]["experiment_seed"] = 1234          # the restored state, not the seed, must decide the draws
    second = d.create_experiment(cfg2, PT_NOOP_DIR)
    assert second.wait(timeout=180) == "COMPLETED"
    t2 = only_trial(second)
    assert t2["warm_start_checkpoint"]["uuid"] == cks[0][1]
    v1 = sorted(t1["validations"], key=lambda v: v["total_batches"])
    v2 = sorted(t2["validations"], key=lambda v: v["total_batches"])
    for step in range(2):
        for metric in ("np_rand", "rand_rand", "torch_rand"):
            a = v1[step + 1]["metrics"]["validation_metrics"][metric]
            b = v2[step]["metrics"]["validation_metrics"][metric]
            assert a == b, f"step {step}, {metric}: {a} != {b}"


def test_pause_of_an_experiment_without_trials(master):
    d = Determined(master["url"])
    cfg = noop_config(master["tmp"])
    cfg["resources"] = {"slots_per_trial": 4096}               # can never be scheduled: no trial gets a container
    exp = d.create_experiment(cfg, NOOP_DIR)
    time.sleep(1.0)
    exp.pause()
    assert exp.describe()["state"] == "PAUSED"
    exp.activate()
    for _ in range(5):
        assert exp.describe()["state"] == "ACTIVE"
        time.sleep(0.3)
    assert all(t.get("total_batches_processed", 0) == 0 for t in exp.describe()["trials"])
    exp.cancel()
    assert exp.wait(timeout=30) == "CANCELED"


@pytest.mark.parametrize("paused", [False, True])
def test_cancel_ten_experiments(master, paused):
    d = Determined(master["url"])
    cfg = noop_config(master["tmp"], batches=3000, training_batch_seconds=0.05)
    exps = [d.create_experiment(cfg, NOOP_DIR, activate=not paused) for _ in range(10)]
    if not paused:
        time.sleep(2.0)     # some are running, the rest are queued
    for e in exps:
        e.cancel()
    for e in exps:
        assert e.wait(timeout=90) == "CANCELED"
    # nothing is left holding slots: a fresh experiment runs to completion
    after = d.create_experiment(noop_config(master["tmp"], batches=10), NOOP_DIR)
    assert after.wait(timeout=120) == "COMPLETED"


def test_a_training_step_longer_than_the_long_poll(master):
    d = Determined(master["url"])
    cfg = noop_config(master["tmp"], batches=1, training_batch_seconds=25, validation_seconds=0)     # the harness's poll is answered every 20 s
    cfg["scheduling_unit"] = 1
    exp = d.create_experiment(cfg, NOOP_DIR)
    assert exp.wait(timeout=180) == "COMPLETED"
    assert only_trial(exp)["total_batches_processed"] == 1


def test_boosted_trees_trial_runs_under_the_master(master):
    """A trial that is not a neural network (a CallbackTrialController over boosting rounds) goes through the same master /
    harness path: scheduling units of boosting rounds, validations, a checkpoint, the searcher metric."""
    import yaml

    d = Determined(master["url"])
    ex = ROOT / "examples" / "decision_trees" / "gbt_titanic_pytorch"
    cfg = yaml.safe_load((ex / "const.yaml").read_text())
    cfg["hyperparameters"]["n_trees"] = 20
    cfg["searcher"]["max_length"] = {"batches": 20}
    cfg["scheduling_unit"] = 5
    cfg["min_validation_period"] = {"batches": 10}
    cfg["checkpoint_storage"] = {"type": "shared_fs", "host_path": str(master["tmp"] / "ckpts")}
    exp = d.create_experiment(cfg, str(ex))
    assert exp.wait(timeout=180) == "COMPLETED"
    t = only_trial(exp)
    assert t["total_batches_processed"] == 20 and len(t["validations"]) >= 2
    last = sorted(t["validations"], key=lambda v: v["total_batches"])[-1]["metrics"]["validation_metrics"]
    assert last["n_trees"] == 20 and last["accuracy"] > 0.7
    ck = d.get_trial(t["id"]).select_checkpoint(latest=True)
    assert "boosted_trees.pt" in ck.resources
