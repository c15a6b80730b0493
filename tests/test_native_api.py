"""Native (script-style) API: ``init_native(controller_cls=, native_context_cls=)`` hands the script a context, the
script's train call unwinds it under the harness, and the controller is then built with ``from_native``
(reference ``experimental/_native.py:118-292``, ``load/_load_implementation.py:117-200``)."""
import pathlib
import textwrap

import determined_b200 as det
from determined_b200 import _execution, experimental, load
from determined_b200 import workload as wl

SCRIPT = textwrap.dedent('''
    import sys
    sys.path.insert(0, {root!r})
    from tests.test_native_api import CountingContext, CountingController
    from determined_b200 import experimental
    ctx = experimental.init_native(controller_cls=CountingController, native_context_cls=CountingContext,
                                   config={{"hyperparameters": {{"global_batch_size": 4}}}})
    ctx.scale = 3
    ctx.fit()
    raise SystemExit("unreachable under the harness")
''')


class CountingContext(det.NativeContext):
    scale = 1

    def fit(self) -> None:
        self._run_train_fn()


class CountingController(det.LoopTrialController):
    """The smallest controller that can answer a workload stream from a native context."""

    @staticmethod
    def pre_execute_hook(env, dist=None):
        CountingController.hook_calls = getattr(CountingController, "hook_calls", 0) + 1

    @classmethod
    def from_trial(cls, *a, **k):
        raise NotImplementedError

    @classmethod
    def from_native(cls, context, env, workloads, load_path=None, rendezvous_info=None, **kwargs):
        return cls(context, env, workloads, load_path, rendezvous_info)

    def run(self) -> None:
        for w, args, respond in self.workloads:
            if w.kind == wl.Workload.Kind.RUN_STEP:
                respond({"metrics": {"batch_metrics": [{"loss": float(self.context.scale)}] * w.num_batches,
                                     "avg_metrics": {"loss": float(self.context.scale)}}, "stop_requested": False})
            elif w.kind == wl.Workload.Kind.COMPUTE_VALIDATION_METRICS:
                respond({"metrics": {"validation_metrics": {"loss": 0.5 * self.context.scale}, "num_inputs": 1},
                         "stop_requested": False})
            elif w.kind == wl.Workload.Kind.CHECKPOINT_MODEL:
                path = pathlib.Path(args[0])
                path.mkdir(parents=True, exist_ok=True)
                (path / "native.txt").write_text(str(self.context.scale))
                respond({"framework": "native", "format": "txt"})
            elif w.kind == wl.Workload.Kind.TERMINATE:
                respond({})
                return


def test_native_script_is_unwound_and_controller_built_from_context(tmp_path):
    root = str(pathlib.Path(__file__).resolve().parent.parent)
    script = tmp_path / "train.py"
    script.write_text(SCRIPT.format(root=root))
    cfg = {"hyperparameters": {"global_batch_size": 4}, "internal": {"native": {"command": [str(script)]}},
           "searcher": {"name": "single", "metric": "loss", "max_length": {"batches": 4}}}
    env, dist = _execution.make_local_execution_env(True, cfg, limit_gpus=1)
    impl = load.load_native_implementation(env, dist, [str(script)])
    assert isinstance(impl, load.NativeImplementation) and impl.context.scale == 3
    seen = {}
    inter = wl.WorkloadResponseInterceptor()

    def stream():
        yield from inter.send(wl.train_workload(1, num_batches=2), [])
        seen["train"] = inter.result()
        yield from inter.send(wl.terminate_workload(1), [])

    controller = load.load_controller_from_trial(impl, env, stream(), None, None, dist)
    assert isinstance(controller, CountingController)
    controller.run()
    assert seen["train"]["metrics"]["avg_metrics"]["loss"] == 3.0


def test_native_local_test_mode_runs_at_the_train_call():
    ctx = experimental.init_native(controller_cls=CountingController, native_context_cls=CountingContext,
                                   config={"hyperparameters": {"global_batch_size": 4}}, local=True, test=True)
    assert isinstance(ctx, CountingContext) and not ctx._test_results
    ctx.scale = 2
    ctx.fit()
    assert ctx._test_results["checkpoint_files"] == ["native.txt"]


def test_trial_api_through_init_native():
    from tests.fixtures import pytorch_models as M

    out = experimental.init_native(trial_def=M.XorTrial, local=True, test=True, config={
        "hyperparameters": {"hidden_size": 4, "learning_rate": 0.05, "global_batch_size": 4, "shuffle": False}})
    assert "state_dict.pth" in out["checkpoint_files"]
