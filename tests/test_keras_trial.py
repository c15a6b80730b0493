"""TFKerasTrial on the workload loop, driven with a NumPy model that implements the Keras model protocol (train_on_batch /
test_on_batch / save_weights / load_weights) — TensorFlow is not installed here. Mirrors the intents of the reference's
harness/tests/experiment/keras tests: training metrics per batch, `val_`-prefixed validation metrics, callbacks, checkpoint /
resume equivalence, data kinds."""
import pathlib

import numpy as np
import pytest

from determined_b200 import errors
from determined_b200 import workload as wl
from determined_b200.keras import TFKerasTrial, TFKerasTrialController
from tests import harness_utils as utils


class NumpyLogReg:
    """Softmax regression with plain SGD behind the Keras model protocol."""

    metrics_names = ["loss", "accuracy"]

    def __init__(self, features: int, classes: int, lr: float) -> None:
        self.w = np.zeros((features, classes))
        self.lr = lr
        self.stop_training = False

    def _forward(self, x, y):
        z = x @ self.w
        z -= z.max(1, keepdims=True)
        p = np.exp(z)
        p /= p.sum(1, keepdims=True)
        loss = -np.log(p[np.arange(len(y)), y] + 1e-12).mean()
        return p, loss, (p.argmax(1) == y).mean()

    def train_on_batch(self, x, y, sample_weight=None, class_weight=None, return_dict=False):
        p, loss, acc = self._forward(x, y)
        p[np.arange(len(y)), y] -= 1
        self.w -= self.lr * x.T @ p / len(y)
        return {"loss": loss, "accuracy": acc} if return_dict else [loss, acc]

    def test_on_batch(self, x, y, sample_weight=None, return_dict=False):
        _, loss, acc = self._forward(x, y)
        return {"loss": loss, "accuracy": acc} if return_dict else [loss, acc]

    def save_weights(self, path):
        np.save(path + ".npy", self.w)

    def load_weights(self, path):
        self.w = np.load(path + ".npy")


def blobs(n, seed):
    rng = np.random.RandomState(seed)
    y = rng.randint(0, 3, n)
    centres = np.array([[2.0, 0.0, 0.0, 1.0], [0.0, 2.0, 0.0, -1.0], [0.0, 0.0, 2.0, 0.0]])
    return centres[y] + 0.5 * rng.randn(n, 4), y


class Recorder:
    def __init__(self):
        self.events = []

    def set_model(self, model):
        self.events.append("set_model")

    def on_train_begin(self, logs=None):
        self.events.append("train_begin")

    def on_epoch_begin(self, epoch, logs=None):
        self.events.append(f"epoch_begin:{epoch}")

    def on_epoch_end(self, epoch, logs=None):
        self.events.append(f"epoch_end:{epoch}")

    def on_train_batch_end(self, batch, logs=None):
        self.events.append(f"batch_end:{batch}")

    def on_test_end(self, logs=None):
        self.events.append("test_end:" + ",".join(sorted(logs)))

    def on_train_end(self, logs=None):
        self.events.append("train_end")


class BlobTrial(TFKerasTrial):
    recorder = None

    def build_model(self):
        return self.context.wrap_model(NumpyLogReg(4, 3, float(self.context.get_hparam("learning_rate"))))

    def build_training_data_loader(self):
        return self.context.wrap_dataset(blobs(64, 1))

    def build_validation_data_loader(self):
        return blobs(40, 2)

    def keras_callbacks(self):
        return [self.recorder] if self.recorder is not None else []


HP = {"global_batch_size": 16, "learning_rate": 0.5}
CFG = {"records_per_epoch": 64, "searcher": {"name": "single", "metric": "val_loss", "max_length": {"batches": 100}}}


def test_trains_validates_and_fires_callbacks():
    BlobTrial.recorder = Recorder()
    tv = utils.TrainAndValidate()
    controller = utils.make_controller(BlobTrial, HP, tv.send(3, validation_freq=3, scheduling_unit=4), exp_config=CFG)
    assert isinstance(controller, TFKerasTrialController)
    controller.run()
    assert len(tv.training_metrics) == 12 and set(tv.training_metrics[0]) == {"loss", "accuracy"}
    assert tv.avg_metrics[-1]["loss"] < tv.avg_metrics[0]["loss"]
    v = tv.validation_metrics[-1]
    assert set(v) == {"val_loss", "val_accuracy"} and v["val_accuracy"] > 0.9
    ev = BlobTrial.recorder.events
    assert ev[:3] == ["set_model", "train_begin", "epoch_begin:0"] and ev[-1] == "train_end"
    # 64 records / batch 16 = 4 batches per epoch: epochs 0, 1, 2 were opened, 0 and 1 closed
    assert [e for e in ev if e.startswith("epoch")] == ["epoch_begin:0", "epoch_end:0", "epoch_begin:1", "epoch_end:1", "epoch_begin:2"]
    assert [e for e in ev if e.startswith("batch_end")] == [f"batch_end:{i}" for i in range(12)]
    assert "test_end:val_accuracy,val_loss" in ev
    BlobTrial.recorder = None


def test_checkpoint_and_resume_match_an_uninterrupted_run(tmp_path):
    straight = utils.TrainAndValidate()
    utils.make_controller(BlobTrial, HP, straight.send(6, validation_freq=6, scheduling_unit=1), exp_config=CFG).run()
    first = utils.TrainAndValidate()
    utils.make_controller(BlobTrial, HP, first.send(3, validation_freq=3, scheduling_unit=1, checkpoint_dir=tmp_path / "ck"), exp_config=CFG).run()
    assert (tmp_path / "ck" / "determined-keras-model-weights.npy").exists() and (tmp_path / "ck" / "determined-keras-trial-state.pkl").exists()
    second = utils.TrainAndValidate()
    utils.make_controller(BlobTrial, HP, second.send(3, validation_freq=3, initial_step_id=4, scheduling_unit=1, batches_done=3), exp_config=CFG,
                          load_path=tmp_path / "ck", initial_batches=3).run()
    resumed = first.training_metrics + second.training_metrics
    assert [m["loss"] for m in resumed] == pytest.approx([m["loss"] for m in straight.training_metrics])      # same batches in the same order
    assert second.validation_metrics[-1]["val_loss"] == pytest.approx(straight.validation_metrics[-1]["val_loss"])


def test_sequence_and_iterable_inputs_and_stop_training():
    class Seq:
        def __init__(self, n, seed):
            self.x, self.y = blobs(n, seed)

        def __len__(self):
            return len(self.x) // 8

        def __getitem__(self, i):
            return self.x[i * 8:(i + 1) * 8], self.y[i * 8:(i + 1) * 8]

    class SeqTrial(BlobTrial):
        def build_training_data_loader(self):
            return Seq(32, 1)

        def build_validation_data_loader(self):
            x, y = blobs(24, 2)
            return [(x[:12], y[:12]), (x[12:], y[12:])]          # any iterable of batches

        def build_model(self):
            m = super().build_model()
            m.stop_training = True                                   # what EarlyStopping sets
            return m

    tv = utils.TrainAndValidate()
    interceptor = wl.WorkloadResponseInterceptor()

    def stream():
        yield from interceptor.send(wl.train_workload(1, num_batches=6), [])
        tv.training_metrics.append(interceptor.metrics_result())
        yield from interceptor.send(wl.validation_workload(1, total_batches_processed=6), [])
        tv.validation_metrics.append(interceptor.metrics_result()["metrics"])
        yield from interceptor.send(wl.terminate_workload(), [])

    utils.make_controller(SeqTrial, HP, stream(), exp_config={"searcher": CFG["searcher"]}).run()
    assert tv.training_metrics[0]["stop_requested"] is True            # model.stop_training becomes the trial's stop request
    assert len(tv.training_metrics[0]["metrics"]["batch_metrics"]) == 6     # 4 batches per pass over the Sequence: it wrapped around
    assert tv.validation_metrics[0]["num_inputs"] == 24


def test_misuse_is_refused():
    class NoWrap(BlobTrial):
        def build_model(self):
            return NumpyLogReg(4, 3, 0.1)

    with pytest.raises(errors.InvalidModelException):
        utils.make_controller(NoWrap, HP, iter(()), exp_config=CFG)

    class NotAModel(BlobTrial):
        def build_model(self):
            return self.context.wrap_model(object())

    with pytest.raises(errors.InvalidModelException):
        utils.make_controller(NotAModel, HP, iter(()), exp_config=CFG)
    class TwoSlots:          # stands in for the communicator of a two-slot trial
        size, rank = 2, 0

    with pytest.raises(errors.InvalidConfigurationException, match="one slot per trial"):
        utils.make_controller(BlobTrial, HP, iter(()), exp_config=CFG, comm=TwoSlots())


def test_reference_import_path():
    from determined.keras import TFKerasTrial as Aliased

    assert Aliased is TFKerasTrial
