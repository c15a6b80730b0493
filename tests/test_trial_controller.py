"""Controller hierarchy (reference harness/determined/_trial_controller.py): the callback-style loop that the no_op chaos
trial uses, driven in-process by a hand-written workload stream; configuration checks of the base class."""
import json
import pathlib
import sys

import pytest

import determined_b200 as det
from determined_b200 import _execution, load
from determined_b200 import workload as wl

ROOT = pathlib.Path(__file__).resolve().parent.parent
NO_OP_DIR = ROOT / "tests" / "fixtures" / "no_op"


def _env(hp, **cfg):
    config = {"searcher": {"name": "single", "metric": "validation_error", "max_length": {"batches": 100}}, "hyperparameters": hp}
    config.update(cfg)
    env, dist = _execution.make_local_execution_env(True, config, hparams=dict(hp), limit_gpus=0, seed=5)
    return env, dist


def _drive(env, dist, stream_fn, load_path=None):
    import importlib.util

    # by path, under a private name: other test modules of the same xdist worker import their own `model_def`
    spec = importlib.util.spec_from_file_location("no_op_model_def", NO_OP_DIR / "model_def.py")
    md = importlib.util.module_from_spec(spec)
    sys.modules["no_op_model_def"] = md
    spec.loader.exec_module(md)
    results = []

    def stream():
        icpt = wl.WorkloadResponseInterceptor()
        for w, args in stream_fn():
            yield from icpt.send(w, args)
            results.append(icpt.result())

    c = load.load_controller_from_trial(md.NoOpTrial, env, stream(), load_path, None, dist)
    assert isinstance(c, det.CallbackTrialController) and isinstance(c, det.TrialController)
    c.run()
    return c, results


def test_callback_controller_runs_all_workload_kinds_and_resumes(tmp_path):
    env, dist = _env({"global_batch_size": 32, "metrics_base": 0.5, "metrics_progression": "decreasing"})
    ck = tmp_path / "ck"

    def stream():
        yield wl.train_workload(1, num_batches=10, total_batches_processed=0), []
        yield wl.train_workload(2, num_batches=10, total_batches_processed=10), []
        yield wl.validation_workload(2, total_batches_processed=20), []
        yield wl.checkpoint_workload(total_batches_processed=20), [ck]
        yield wl.terminate_workload(), []

    c, res = _drive(env, dist, stream)
    assert res[0]["metrics"]["num_inputs"] == 320 and len(res[0]["metrics"]["batch_metrics"]) == 10
    assert res[0]["metrics"]["avg_metrics"]["loss"] == pytest.approx(0.5)        # base ** 1 trained step
    assert res[1]["metrics"]["avg_metrics"]["loss"] == pytest.approx(0.25)
    assert res[2]["metrics"]["validation_metrics"]["validation_error"] == pytest.approx(0.25)
    assert isinstance(res[4], wl.Skipped)
    assert json.loads((ck / "no_op_checkpoint").read_text()) == {"1": 1, "2": 1}

    def stream2():
        yield wl.train_workload(3, num_batches=10, total_batches_processed=20), []
        yield wl.terminate_workload(), []

    c2, res2 = _drive(env, dist, stream2, load_path=ck)
    assert c2.steps_trained() == 3 and res2[0]["metrics"]["avg_metrics"]["loss"] == pytest.approx(0.125)


def test_chaos_and_scripted_failures_raise():
    env, dist = _env({"global_batch_size": 32, "chaos_probability": 1.0})
    with pytest.raises(Exception, match="CHAOS"):
        _drive(env, dist, lambda: iter([(wl.train_workload(1, num_batches=1, total_batches_processed=0), [])]))
    env, dist = _env({"global_batch_size": 32, "fail_on_first_validation": "boom"})
    with pytest.raises(Exception, match="boom"):
        _drive(env, dist, lambda: iter([(wl.validation_workload(1, total_batches_processed=0), [])]))


def test_base_class_rejects_configurations_the_interface_cannot_honour():
    env, dist = _env({"global_batch_size": 32}, optimizations={"mixed_precision": "O1"})
    with pytest.raises(det.errors.InvalidConfigurationException, match="mixed precision"):
        _drive(env, dist, lambda: iter([]))
    env, dist = _env({"global_batch_size": 32}, optimizations={"average_training_metrics": True})
    with pytest.raises(det.errors.InvalidConfigurationException, match="average_training_metrics"):
        _drive(env, dist, lambda: iter([]))


def test_pytorch_controller_is_a_loop_controller():
    from determined_b200.pytorch import PyTorchTrial, PyTorchTrialController

    assert issubclass(PyTorchTrialController, det.LoopTrialController) and issubclass(PyTorchTrial, det.Trial)
