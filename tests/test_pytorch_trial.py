"""PyTorchTrial controller tests on CPU: exact-gradient check, validation cadence, checkpoint→restore
equivalence, reproducibility, reducers, callbacks, gradient clipping, aggregation, legacy API, local test mode.
Test ideas follow the reference's ``harness/tests/experiment/pytorch/test_pytorch_trial.py``."""
import pathlib

import numpy as np
import pytest
import torch

from determined_b200 import experimental
from tests.fixtures import pytorch_models as M
from tests.harness_utils import TrainAndValidate, make_controller

XOR_HP = {"hidden_size": 4, "learning_rate": 0.05, "global_batch_size": 4, "shuffle": False}


def test_onevar_exact_gradients():
    """w' = w - lr * dMSE/dw must hold batch after batch (SGD on y = 2x, w0 = 0)."""
    tv = TrainAndValidate()
    hp = {"learning_rate": 0.01, "global_batch_size": 4}
    make_controller(M.OneVarTrial, hp, tv.send(steps=8, validation_freq=4)).run()
    data = M._OneVarData()
    w = 0.0
    for i, m in enumerate(tv.training_metrics):
        xs = np.array([float(data[j][0]) for j in range(4 * i, 4 * i + 4)])
        ys = 2.0 * xs
        loss = np.mean((w * xs - ys) ** 2)
        grad = np.mean(2 * (w * xs - ys) * xs)
        assert m["w_before"] == pytest.approx(w, abs=1e-5)
        assert m["loss"] == pytest.approx(loss, rel=1e-4, abs=1e-6)
        w = w - 0.01 * grad
        assert m["w_after"] == pytest.approx(w, abs=1e-5)
    assert len(tv.validation_metrics) == 2
    assert tv.validation_metrics[1]["val_loss"] < tv.validation_metrics[0]["val_loss"]


def test_xor_multi_validation_and_metrics_shape():
    tv = TrainAndValidate()
    make_controller(M.XorTrial, XOR_HP, tv.send(steps=6, validation_freq=2, scheduling_unit=3)).run()
    assert len(tv.training_metrics) == 18 and len(tv.validation_metrics) == 3
    for m in tv.training_metrics:
        assert set(m) == {"loss", "lr"} and isinstance(m["loss"], float)
    for v in tv.validation_metrics:
        assert set(v) == {"loss", "binary_error"}
    # per-batch LR scheduler (StepLR 5, 0.9) stepped by the harness every batch
    lrs = [m["lr"] for m in tv.training_metrics]
    assert lrs[0] == pytest.approx(0.05) and lrs[5] == pytest.approx(0.045) and lrs[10] == pytest.approx(0.0405)
    for avg in tv.avg_metrics:
        assert "loss" in avg


def test_checkpoint_restore_equivalence(tmp_path: pathlib.Path):
    """A: train 1 step, checkpoint, restore in a fresh controller, train 1 more. B: train 2 steps. Same metrics."""
    ckpt = tmp_path / "ckpt"
    a1 = TrainAndValidate()
    make_controller(M.XorTrial, XOR_HP, a1.send(steps=1, validation_freq=1, scheduling_unit=5, checkpoint_dir=ckpt)).run()
    assert (ckpt / "state_dict.pth").exists() and (ckpt / "code").is_dir()
    assert a1.checkpoint_response["framework"].startswith("torch-") and a1.checkpoint_response["format"] == "cloudpickle"
    a2 = TrainAndValidate()
    make_controller(M.XorTrial, XOR_HP, a2.send(steps=1, validation_freq=1, scheduling_unit=5, initial_step_id=2,
                                                batches_done=5), load_path=ckpt, initial_batches=5).run()
    b = TrainAndValidate()
    make_controller(M.XorTrial, XOR_HP, b.send(steps=2, validation_freq=1, scheduling_unit=5)).run()
    a_train = a1.training_metrics + a2.training_metrics
    assert len(a_train) == len(b.training_metrics) == 10
    for x, y in zip(a_train, b.training_metrics):
        assert x["loss"] == pytest.approx(y["loss"], rel=1e-6)
        assert x["lr"] == pytest.approx(y["lr"])
    for x, y in zip(a1.validation_metrics + a2.validation_metrics, b.validation_metrics):
        assert x["loss"] == pytest.approx(y["loss"], rel=1e-6)


def test_checkpoint_keys_and_legacy_load(tmp_path: pathlib.Path):
    ckpt = tmp_path / "c"
    tv = TrainAndValidate()
    make_controller(M.XorTrialCallbacks, XOR_HP, tv.send(steps=1, checkpoint_dir=ckpt)).run()
    import cloudpickle

    sd = torch.load(str(ckpt / "state_dict.pth"), weights_only=False)
    assert set(sd) >= {"models_state_dict", "optimizers_state_dict", "lr_schedulers_state_dict", "callbacks", "rng_state"}
    assert set(sd["rng_state"]) >= {"cpu_rng_state", "np_rng_state", "random_rng_state"}
    assert sd["callbacks"]["counter"]["validation_steps_ended"] == 1
    # legacy layout: singular keys under determined/state_dict.pth
    legacy = tmp_path / "legacy" / "determined"
    legacy.mkdir(parents=True)
    torch.save({"model_state_dict": sd["models_state_dict"][0], "optimizer_state_dict": sd["optimizers_state_dict"][0]},
               str(legacy / "state_dict.pth"))
    tv2 = TrainAndValidate()
    c = make_controller(M.XorTrial, XOR_HP, tv2.send(steps=1), load_path=tmp_path / "legacy")
    for k, v in c.context.models[0].state_dict().items():
        assert torch.equal(v, sd["models_state_dict"][0][k])


def test_reproducibility_same_seed_same_metrics():
    hp = dict(XOR_HP, shuffle=True)
    runs = []
    for _ in range(2):
        tv = TrainAndValidate()
        make_controller(M.XorTrial, hp, tv.send(steps=3, scheduling_unit=4), seed=1234).run()
        runs.append(tv)
    for x, y in zip(runs[0].training_metrics, runs[1].training_metrics):
        assert x["loss"] == y["loss"]
    tv = TrainAndValidate()
    make_controller(M.XorTrial, hp, tv.send(steps=3, scheduling_unit=4), seed=4321).run()
    assert any(x["loss"] != y["loss"] for x, y in zip(runs[0].training_metrics, tv.training_metrics))


def test_per_metric_and_custom_reducers():
    tv = TrainAndValidate()
    make_controller(M.XorTrialPerMetricReducers, XOR_HP, tv.send(steps=1)).run()
    tv2 = TrainAndValidate()
    make_controller(M.XorTrial, XOR_HP, tv2.send(steps=1)).run()
    assert tv.validation_metrics[0]["loss"] == pytest.approx(tv2.validation_metrics[0]["loss"])
    assert tv.validation_metrics[0]["binary_error"] >= tv2.validation_metrics[0]["binary_error"]
    tv3 = TrainAndValidate()
    make_controller(M.XorTrialCustomReducer, XOR_HP, tv3.send(steps=1)).run()
    assert tv3.validation_metrics[0]["loss"] == pytest.approx(tv2.validation_metrics[0]["loss"], rel=1e-5)


def test_evaluate_full_dataset_and_exactly_one_eval_method():
    tv = TrainAndValidate()
    make_controller(M.XorTrialFullDataset, XOR_HP, tv.send(steps=1)).run()
    assert "loss" in tv.validation_metrics[0]

    class Both(M.XorTrial):
        def evaluate_full_dataset(self, data_loader):
            return {}

    with pytest.raises(RuntimeError, match="exactly one"):
        make_controller(Both, XOR_HP, iter([]))


def test_callbacks_fire():
    tv = TrainAndValidate()
    c = make_controller(M.XorTrialCallbacks, XOR_HP, tv.send(steps=2, validation_freq=1))
    c.run()
    assert c.trial.counter.validation_steps_started == 2 and c.trial.counter.validation_steps_ended == 2


def test_grad_clipping_changes_updates():
    hp = dict(XOR_HP, learning_rate=0.5)
    free, clipped = TrainAndValidate(), TrainAndValidate()
    make_controller(M.XorTrialGradClipping, hp, free.send(steps=4)).run()
    make_controller(M.XorTrialGradClipping, dict(hp, clip_norm=1e-4), clipped.send(steps=4)).run()
    assert free.training_metrics[0]["loss"] == pytest.approx(clipped.training_metrics[0]["loss"])
    assert free.training_metrics[-1]["loss"] != pytest.approx(clipped.training_metrics[-1]["loss"], rel=1e-4)


def test_aggregation_frequency_steps_every_nth_batch():
    hp = {"learning_rate": 0.01, "global_batch_size": 4}
    tv = TrainAndValidate()
    make_controller(M.OneVarTrial, hp, tv.send(steps=1, scheduling_unit=4, validation_freq=0),
                    exp_config={"optimizations": {"aggregation_frequency": 2, "average_aggregated_gradients": True}}).run()
    m = tv.training_metrics
    assert m[0]["w_after"] == m[0]["w_before"]           # batch 0 only accumulates
    assert m[1]["w_after"] != m[1]["w_before"]           # batch 1 closes the window
    assert m[2]["w_after"] == m[2]["w_before"]
    data = M._OneVarData()
    xs = [np.array([float(data[j][0]) for j in range(4 * i, 4 * i + 4)]) for i in range(2)]
    g = np.mean([np.mean(2 * (0.0 * x - 2 * x) * x) for x in xs])  # averaged over the two micro-batches at w = 0
    assert m[1]["w_after"] == pytest.approx(-0.01 * g, rel=1e-5)


def test_legacy_single_model_api():
    tv = TrainAndValidate()
    make_controller(M.LegacyXorTrial, {"global_batch_size": 4}, tv.send(steps=3)).run()
    assert len(tv.training_metrics) == 3 and "loss" in tv.validation_metrics[0]


def test_context_accessors():
    tv = TrainAndValidate()
    c = make_controller(M.XorTrial, XOR_HP, tv.send(steps=1))
    ctx = c.context
    assert ctx.get_per_slot_batch_size() == 4 and ctx.get_global_batch_size() == 4
    assert ctx.get_hparam("hidden_size") == 4 and ctx.get_trial_id() == 1 and ctx.get_experiment_id() == 1
    assert ctx.distributed.get_rank() == 0 and ctx.distributed.get_size() == 1
    with pytest.raises(KeyError):
        ctx.get_hparam("nope")
    assert ctx.get_model() is c.trial.model and ctx.get_optimizer() is c.trial.opt
    ctx.set_stop_requested(True)
    c.run()


def test_stop_requested_is_reported():
    class Stopper(M.XorTrial):
        def train_batch(self, batch, epoch_idx, batch_idx):
            self.context.set_stop_requested(True)
            return super().train_batch(batch, epoch_idx, batch_idx)

    from determined_b200 import workload as wl

    got = {}
    icpt = wl.WorkloadResponseInterceptor()

    def stream():
        yield from icpt.send(wl.train_workload(1, num_batches=1), [])
        got["r"] = icpt.metrics_result()
        yield from icpt.send(wl.terminate_workload(), [])

    make_controller(Stopper, XOR_HP, stream()).run()
    assert got["r"]["stop_requested"] is True


def test_test_one_batch_local_mode():
    cfg = {"hyperparameters": XOR_HP, "searcher": {"name": "single", "metric": "loss", "max_length": {"batches": 10}}}
    res = experimental.create(M.XorTrial, cfg, local=True, test=True)
    assert "state_dict.pth" in res["checkpoint_files"] and "code" in res["checkpoint_files"]
    with pytest.raises(NotImplementedError):
        experimental.create(M.XorTrial, cfg, local=True, test=False)


def test_native_parallel_wraps_data_parallel(monkeypatch):
    """`resources.native_parallel`: one process, nn.DataParallel around the model, whole global batch per process,
    aggregation_frequency must stay 1 (reference _pytorch_context.py:135-144, _env_context.py:89)."""
    from determined_b200.pytorch import PyTorchTrialContext

    monkeypatch.setattr(PyTorchTrialContext, "native_parallel_gpus", lambda self: 2)
    cfg = {"resources": {"slots_per_trial": 2, "native_parallel": True}}
    tv = TrainAndValidate()
    c = make_controller(M.XorTrial, XOR_HP, tv.send(steps=2), exp_config=cfg)
    assert isinstance(c.context.models[0], torch.nn.DataParallel)
    assert c.context.get_per_slot_batch_size() == c.context.get_global_batch_size() == 4
    assert type(c.context.optimizers[0]).__name__ != "FusedOptimizer"
    c.run()
    assert len(tv.training_metrics) == 2 and "loss" in tv.validation_metrics[0]
    cfg["optimizations"] = {"aggregation_frequency": 2}
    with pytest.raises(ValueError, match="native_parallel"):
        make_controller(M.XorTrial, XOR_HP, tv.send(steps=1), exp_config=cfg)


def test_native_parallel_off_on_cpu():
    tv = TrainAndValidate()
    c = make_controller(M.XorTrial, XOR_HP, tv.send(steps=1), exp_config={"resources": {"native_parallel": True}})
    assert c.context.native_parallel_gpus() == 0 and not isinstance(c.context.models[0], torch.nn.DataParallel)


def test_nonscalar_validation_metrics_pass_through():
    """A validation that returns an array next to its scalars (reference test_xor_nonscalar_validation)."""
    tv = TrainAndValidate()
    make_controller(M.XorTrialNonScalarValidation, XOR_HP, tv.send(steps=4, validation_freq=4)).run()
    v = tv.validation_metrics[0]
    assert 0.0 <= v["binary_error"] <= 1.0
    assert np.asarray(v["predictions"]).shape == (32, 1)


def test_restore_into_a_different_architecture_fails(tmp_path: pathlib.Path):
    """A checkpoint of hidden_size 4 must not load into hidden_size 3 (reference test_restore_invalid_checkpoint)."""
    ckpt = tmp_path / "ckpt"
    make_controller(M.XorTrial, XOR_HP, TrainAndValidate().send(steps=1, validation_freq=1, checkpoint_dir=ckpt)).run()
    with pytest.raises(RuntimeError, match="size mismatch"):
        make_controller(M.XorTrial, dict(XOR_HP, hidden_size=3), TrainAndValidate().send(steps=1), load_path=ckpt)


def test_workloads_of_varying_length():
    """RUN_STEP workloads of 1, 2, ... 9 batches: each answers with exactly that many batch metrics and the batch index
    keeps counting across them (reference test_variable_workload_size)."""
    from determined_b200 import workload as wl

    seen = []

    def stream():
        icpt = wl.WorkloadResponseInterceptor()
        total = 0
        for step_id in range(1, 10):
            yield from icpt.send(wl.train_workload(step_id, num_batches=step_id, total_batches_processed=total), [])
            bm = icpt.result()["metrics"]["batch_metrics"]
            assert len(bm) == step_id
            seen.extend(bm)
            total += step_id
        yield from icpt.send(wl.terminate_workload(10), [])

    make_controller(M.XorTrial, XOR_HP, stream()).run()
    assert len(seen) == 45 and all(np.isfinite(m["loss"]) for m in seen)


def test_create_trial_instance_outside_training(tmp_path: pathlib.Path):
    """`det.experimental.create_trial_instance`: a trial object with restored weights, no workloads involved."""
    ckpt = tmp_path / "ckpt"
    tv = TrainAndValidate()
    c = make_controller(M.XorTrial, XOR_HP, tv.send(steps=2, validation_freq=2, checkpoint_dir=ckpt))
    c.run()
    trial = experimental.create_trial_instance(M.XorTrial, str(ckpt), config={"hyperparameters": XOR_HP}, hparams=XOR_HP)
    for a, b in zip(trial.model.parameters(), c.trial.model.parameters()):
        assert torch.equal(a.detach().cpu(), b.detach().cpu())
