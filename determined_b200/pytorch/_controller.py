"""``PyTorchTrialController`` — runs a ``PyTorchTrial`` against a stream of workloads.

Parity: reference ``harness/determined/pytorch/_pytorch_trial.py:24-766`` (controller init order, the
RUN_STEP / COMPUTE_VALIDATION_METRICS / CHECKPOINT_MODEL / TERMINATE loop, metric packaging, checkpoint
layout and legacy load paths, RNG save/restore) and ``_trial_controller.py:14-315``.

B200-first differences:

* the training loop never synchronises with the host per batch: per-batch metrics stay on the device and are
  fetched with ONE device→host copy at the end of the workload (the reference does ``metric.cpu().numpy()``
  for every metric of every batch, ``:393-395``);
* cross-rank metric averaging / validation reduction use the peer-memory metric kernel (K8/K9) instead of
  ZMQ + pickle; custom ``MetricReducer`` state still travels as pickled objects over the control plane;
* initial parameter / optimizer-state broadcast is a peer pull over the symmetric arena (K2);
* optional CUDA-graph replay of ``train_batch`` (``context.experimental.enable_cuda_graphs()``).
"""
from __future__ import annotations

import logging
import pathlib
import pickle
import random
from typing import Any, Dict, Iterator, List, Optional, Tuple, Union, cast

import numpy as np
import torch

from .. import workload as wl
from .._env import DistributedContext, EnvContext, RendezvousInfo
from .._trial_controller import LoopTrialController
from ..ops import functional as F
from ..parallel.fused_optimizer import FusedOptimizer
from ..utils import fast_start, util
from . import _callback, _data, _reducer
from ._context import PyTorchTrialContext
from ._graph import GraphedTrainBatch
from ._lr_scheduler import LRScheduler
from ._trial import PyTorchTrial

try:  # cloudpickle keeps closures in callbacks picklable, as the reference format requires
    import cloudpickle as _pickle_module
except Exception:  # pragma: no cover
    _pickle_module = pickle

log = logging.getLogger("determined.pytorch")

_REDUCER_OPS = {_reducer.Reducer.AVG: F.RED_AVG, _reducer.Reducer.SUM: F.RED_SUM, _reducer.Reducer.MAX: F.RED_MAX,
                _reducer.Reducer.MIN: F.RED_MIN}


def _to_scalar_tensor(v: Any, device: torch.device) -> Optional[torch.Tensor]:
    if isinstance(v, torch.Tensor):
        return v.detach().reshape(()).float() if v.numel() == 1 else None
    if isinstance(v, (int, float, np.integer, np.floating)):
        return torch.tensor(float(v), dtype=torch.float32, device=device)
    return None


class PyTorchTrialController(LoopTrialController):
    def __init__(self, trial_inst: PyTorchTrial, context: PyTorchTrialContext, env: EnvContext,
                 workloads: wl.Stream, load_path: Optional[pathlib.Path] = None,
                 rendezvous_info: Optional[RendezvousInfo] = None) -> None:
        if not isinstance(trial_inst, PyTorchTrial):
            raise TypeError("PyTorchTrialController needs a PyTorchTrial instance")
        super().__init__(context, env, workloads, load_path, rendezvous_info)
        self.trial = trial_inst
        self.callbacks = self.trial.build_callbacks()
        self._apply_backwards_compatibility()
        if not context.models:
            raise RuntimeError("the trial must wrap at least one model with context.wrap_model() in __init__")
        if not context.optimizers:
            raise RuntimeError("the trial must wrap at least one optimizer with context.wrap_optimizer() in __init__")
        self._check_evaluate_implementation()
        self._set_data_loaders()
        self._load()
        self._broadcast_initial_state()
        self._graphed: Optional[GraphedTrainBatch] = None
        if context.device.type == "cuda":
            # the dropout kernels hash (seed, step counter, element index); the counter starts at a point fixed by the
            # trial seed and the batches already trained, so a resumed trial does not redraw the masks it began with
            F.set_step_seed(env.trial_seed + env.initial_workload.total_batches_processed, context.device)
        if context.experimental.cuda_graphs and context.device.type == "cuda":
            self._graphed = GraphedTrainBatch(self.trial, context)

    # ---- construction helpers --------------------------------------------------------------------------
    @staticmethod
    def pre_execute_hook(env: EnvContext, dist: Optional[DistributedContext] = None) -> None:
        """Seed python / numpy / torch with the trial seed: every rank gets the SAME seed (so model init is
        identical) and sees different data only through the batch sharding (reference ``:75-95``)."""
        seed = env.trial_seed
        random.seed(seed)
        np.random.seed(seed % (2 ** 32))
        torch.manual_seed(seed)

    @classmethod
    def from_trial(cls, trial_inst: PyTorchTrial, context: PyTorchTrialContext, env: EnvContext, workloads: wl.Stream,
                   load_path: Optional[pathlib.Path] = None, rendezvous_info: Optional[RendezvousInfo] = None,
                   **_: Any) -> "PyTorchTrialController":
        return cls(trial_inst, context, env, workloads, load_path, rendezvous_info)

    @staticmethod
    def supports_mixed_precision() -> bool:
        return True

    @staticmethod
    def supports_averaging_training_metrics() -> bool:
        return True

    def _apply_backwards_compatibility(self) -> None:
        """Trials written against the deprecated single-model API (``build_model`` / ``optimizer`` /
        ``create_lr_scheduler``) get their objects wrapped here (reference ``:97-174``)."""
        if self.context.models:
            return
        if not util.is_overridden(self.trial.build_model, PyTorchTrial):
            return
        model = self.context.wrap_model(self.trial.build_model())
        opt = self.context.wrap_optimizer(self.trial.optimizer(model))
        sched = self.trial.create_lr_scheduler(opt)
        if sched is not None:
            if isinstance(sched, LRScheduler):
                self.context.lr_schedulers.append(sched)
            else:
                self.context.wrap_lr_scheduler(sched, LRScheduler.StepMode.STEP_EVERY_BATCH)

        trial = self.trial
        ctx = self.context
        user_train = trial.train_batch

        def train_batch(batch, epoch_idx, batch_idx, _user=user_train):  # old signature: (batch, model, epoch, idx)
            out = _user(batch, model, epoch_idx, batch_idx)
            loss = out["loss"] if isinstance(out, dict) else out
            ctx.backward(loss)
            clip = None
            for cb in self.callbacks.values():
                if util.is_overridden(cb.on_before_optimizer_step, _callback.PyTorchCallback):
                    clip = cb.on_before_optimizer_step if clip is None else clip
            ctx.step_optimizer(opt, clip_grads=clip)
            return out

        trial.train_batch = train_batch  # type: ignore[assignment]
        if util.is_overridden(trial.evaluate_batch, PyTorchTrial):
            user_eval = trial.evaluate_batch
            trial.evaluate_batch = lambda batch, _u=user_eval: _u(batch, model)  # type: ignore[assignment]
        if util.is_overridden(trial.evaluate_full_dataset, PyTorchTrial):
            user_full = trial.evaluate_full_dataset
            trial.evaluate_full_dataset = lambda dl, _u=user_full: _u(dl, model)  # type: ignore[assignment]

    def _check_evaluate_implementation(self) -> None:
        has_batch = util.is_overridden(self.trial.evaluate_batch, PyTorchTrial) or "evaluate_batch" in self.trial.__dict__
        has_full = (util.is_overridden(self.trial.evaluate_full_dataset, PyTorchTrial)
                    or "evaluate_full_dataset" in self.trial.__dict__)
        if has_batch == has_full:
            raise RuntimeError("Please define exactly one of evaluate_batch() or evaluate_full_dataset(): "
                               f"evaluate_batch {'is' if has_batch else 'is not'} defined, "
                               f"evaluate_full_dataset {'is' if has_full else 'is not'} defined")
        self._evaluate_batch_defined = has_batch

    def _set_data_loaders(self) -> None:
        skip = self.env.initial_workload.total_batches_processed
        nrep, rank = self.comm.size, self.comm.rank
        train = self.trial.build_training_data_loader()
        if not isinstance(train, _data.DataLoader):
            raise TypeError("build_training_data_loader() must return a determined.pytorch.DataLoader")
        self.training_loader = train.get_data_loader(repeat=True, skip=skip, num_replicas=nrep, rank=rank)
        self.training_iterator: Iterator = iter(self.training_loader)
        self._train_len = len(train) // max(nrep, 1) if len(train) else None
        self.context._epoch_len = self._epoch_len()
        val = self.trial.build_validation_data_loader()
        if not isinstance(val, _data.DataLoader):
            raise TypeError("build_validation_data_loader() must return a determined.pytorch.DataLoader")
        if self._evaluate_batch_defined:
            self.validation_loader = val.get_data_loader(repeat=False, skip=0, num_replicas=nrep, rank=rank)
        else:
            self.validation_loader = val.get_data_loader(repeat=False, skip=0) if self.is_chief else None

    def _epoch_len(self) -> Optional[int]:
        rpe = self.env.experiment_config.get_records_per_epoch()
        if rpe:
            return max(rpe // max(self.env.global_batch_size, 1), 1)
        return self._train_len

    def _broadcast_initial_state(self) -> None:
        """Rank 0's parameters, buffers and optimizer state become everyone's (K2)."""
        if self.comm.size <= 1:
            return
        covered = set()
        for opt in self.context.optimizers:
            if isinstance(opt, FusedOptimizer) and opt.fused:
                self.comm.broadcast_tensors(opt.param_buckets() + opt.state_buckets(), root=0)
                opt.refresh_shadows()
                steps = self.comm.broadcast_object([g.step_count for g in opt.groups], 0)
                for g, s in zip(opt.groups, steps):
                    g.step_count = s
                for grp in opt.param_groups:
                    covered.update(id(p) for p in grp["params"])
            else:
                sd = self.comm.broadcast_object(opt.state_dict() if self.is_chief else None, 0)
                if not self.is_chief:
                    opt.load_state_dict(sd)
        # buffers (BatchNorm statistics, ...) and parameters no fused optimizer owns
        rest = [t for t in self.context._main_model.state_dict(keep_vars=True).values()
                if isinstance(t, torch.Tensor) and id(t) not in covered]
        self.comm.broadcast_tensors([t.data for t in rest], root=0)

    # ---- main loop ----------------------------------------------------------------------------------------------
    def run(self) -> None:
        for w, args, respond in self.workloads:
            if w.kind == wl.Workload.Kind.RUN_STEP:
                respond(self._train_for_step(w.step_id, w.num_batches, w.total_batches_processed))
            elif w.kind == wl.Workload.Kind.COMPUTE_VALIDATION_METRICS:
                respond(self._compute_validation_metrics())
            elif w.kind == wl.Workload.Kind.CHECKPOINT_MODEL:
                if len(args) != 1:
                    raise ValueError("CHECKPOINT_MODEL expects exactly one argument: the directory to write")
                path = pathlib.Path(args[0])
                if self.is_chief:
                    path.mkdir(parents=True, exist_ok=True)
                    respond(self._save(path))
                else:
                    respond(wl.Skipped())
            elif w.kind == wl.Workload.Kind.TERMINATE:
                respond({} if self.is_chief else wl.Skipped())
                break
            else:
                raise AssertionError(f"unexpected workload kind {w.kind}")

    def get_epoch_idx(self, batch_id: int) -> int:
        """Epoch a batch index falls into (reference ``_pytorch_trial.py:289-290``)."""
        n = self.context._epoch_len
        return batch_id // n if n else 0

    # ---- training -----------------------------------------------------------------------------------------------------
    def _auto_step_lr_schedulers(self, batch_idx: int) -> None:
        for sched in self.context.lr_schedulers:
            mode = sched._step_mode
            if mode == LRScheduler.StepMode.STEP_EVERY_BATCH:
                if (batch_idx + 1) % sched._frequency == 0:
                    sched.step()
            elif mode == LRScheduler.StepMode.STEP_EVERY_EPOCH:
                if self.context._epoch_len and (batch_idx + 1) % (self.context._epoch_len * sched._frequency) == 0:
                    sched.step()

    def _train_for_step(self, step_id: int, num_batches: int, total_batches_processed: int) -> wl.Response:
        if step_id <= 0 or num_batches <= 0:
            raise ValueError("step_id and num_batches must be positive")
        ctx = self.context
        for m in ctx.models:
            # same effect as the reference's unconditional model.train() per step, without paying Module.__setattr__ on
            # every submodule when nothing left training mode (the call sits in the host gap between two workloads)
            if not all(sub.training for sub in m.modules()):
                m.train()
        for name, (red, for_train, _) in ctx.reducers.items():
            if for_train:
                red.reset()
        for cb in self.callbacks.values():
            cb.on_train_step_start(step_id)
        epoch_len = ctx._epoch_len
        start = total_batches_processed
        per_batch: List[Dict[str, Any]] = []
        num_inputs = 0
        for batch_idx in range(start, start + num_batches):
            batch = next(self.training_iterator)
            num_inputs += _data.data_length(batch)
            ctx._set_current_batch(batch_idx, epoch_len)
            epoch_idx = batch_idx // epoch_len if epoch_len else 0
            self.context.experimental._set_step_info(epoch_idx, batch_idx)
            if self._graphed is not None:
                tr = self._graphed.run(batch, epoch_idx, batch_idx)
            else:
                batch = ctx.to_device(batch)
                with ctx.autocast():
                    tr = self.trial.train_batch(batch=batch, epoch_idx=epoch_idx, batch_idx=batch_idx)
                if ctx.device.type == "cuda":
                    F.advance_step_seed(ctx.device)     # fresh dropout masks for the next batch
            if isinstance(tr, torch.Tensor):
                tr = {"loss": tr}
            if not isinstance(tr, dict) or "loss" not in tr:
                raise ValueError("train_batch() must return the loss tensor or a dict of metrics that includes 'loss'")
            per_batch.append(tr)
            self._auto_step_lr_schedulers(batch_idx)
            if fast_start.FIRST_BATCH_TS is None:
                fast_start.mark_first_batch(ctx.device)
        ctx._current_batch_idx = None
        num_inputs *= self.comm.size
        metrics = self._finalize_training_metrics(per_batch)
        extra = self._reduce_custom(for_training=True)
        if not self.is_chief:
            return wl.Skipped()
        out = util.make_metrics(num_inputs, metrics)
        out["avg_metrics"].update(extra)
        for cb in self.callbacks.values():
            cb.on_train_step_end(step_id, out)
        return {"metrics": out, "stop_requested": ctx.get_stop_requested()}

    def _finalize_training_metrics(self, per_batch: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        """One device→host transfer for the whole workload; optional cross-rank mean on the device."""
        dev = self.context.device
        names = sorted({k for d in per_batch for k in d})
        scalar_names = []
        for n in names:
            if all(n in d and _to_scalar_tensor(d[n], dev) is not None for d in per_batch):
                scalar_names.append(n)
        out: List[Dict[str, Any]] = [dict() for _ in per_batch]
        if scalar_names:
            mat = torch.stack([torch.stack([cast(torch.Tensor, _to_scalar_tensor(d[n], dev)).to(dev)
                                            for n in scalar_names]) for d in per_batch])
            if self.env.experiment_config.averaging_training_metrics_enabled() and self.comm.size > 1:
                mat = self.comm.reduce_metrics(mat.contiguous(), F.RED_AVG)
            host = mat.cpu().numpy()
            for i in range(len(per_batch)):
                for j, n in enumerate(scalar_names):
                    out[i][n] = float(host[i, j])
        for i, d in enumerate(per_batch):
            for n, v in d.items():
                if n in scalar_names:
                    continue
                out[i][n] = v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v
        return out

    def _reduce_custom(self, for_training: bool) -> Dict[str, Any]:
        out: Dict[str, Any] = {}
        for name, (red, tr, va) in self.context.reducers.items():
            if not (tr if for_training else va):
                continue
            per_slot = red.per_slot_reduce()
            if isinstance(per_slot, torch.Tensor):
                per_slot = per_slot.detach().cpu()
            gathered = self.comm.allgather_object(per_slot)
            if self.is_chief:
                val = red.cross_slot_reduce(gathered)
                if isinstance(val, dict):
                    out.update(val)
                else:
                    out[name] = val
        return out

    # ---- validation ---------------------------------------------------------------------------------------------------------
    def _compute_validation_metrics(self) -> wl.Response:
        ctx = self.context
        for m in ctx.models:
            m.eval()
        for cb in self.callbacks.values():
            cb.on_validation_start()
        for name, (red, _, va) in ctx.reducers.items():
            if va:
                red.reset()
        num_inputs = 0
        metrics: Dict[str, Any] = {}
        with torch.no_grad():
            if self._evaluate_batch_defined:
                keys: Optional[List[str]] = None
                rows: List[List[Any]] = []
                for batch in cast(Any, self.validation_loader):
                    num_inputs += _data.data_length(batch)
                    batch = ctx.to_device(batch)
                    for cb in self.callbacks.values():
                        cb.on_validation_step_start()
                    with ctx.autocast():
                        vm = self.trial.evaluate_batch(batch=batch)
                    if not isinstance(vm, dict):
                        raise ValueError("evaluate_batch() must return a dict of metrics")
                    for cb in self.callbacks.values():
                        cb.on_validation_step_end(vm)
                    if keys is None:
                        keys = list(vm.keys())
                    elif set(keys) != set(vm.keys()):
                        raise ValueError(f"validation metric names changed between batches: {keys} vs {list(vm)}")
                    rows.append([vm[k] for k in keys])
                if self.comm.size > 1:
                    # every rank runs the same collectives: a rank whose shard held no batch adopts the metric names of
                    # the first rank that saw one (DistributedBatchSampler deals whole batches, so shards can be empty)
                    all_keys = self.comm.allgather_object(keys)
                    seen = [k for k in all_keys if k is not None]
                    for k in seen[1:]:
                        if set(k) != set(seen[0]):
                            raise ValueError(f"validation metric names differ between ranks: {seen[0]} vs {k}")
                    keys = seen[0] if seen else None
                if keys is None:
                    # reference: check.gt(len(validation_loader), 0) (_pytorch_trial.py:446)
                    raise ValueError("the validation data loader produced no batches on any rank")
                metrics = self._reduce_validation(keys, rows)
                num_inputs = int(sum(self.comm.allgather_object(num_inputs)))
            else:
                if self.is_chief:
                    metrics = self.trial.evaluate_full_dataset(data_loader=self.validation_loader)
                    if not isinstance(metrics, dict):
                        raise ValueError("evaluate_full_dataset() must return a dict of metrics")
                    metrics = {k: (float(v.detach().cpu()) if isinstance(v, torch.Tensor) and v.numel() == 1 else
                                   (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else v))
                               for k, v in metrics.items()}
                    # the reference reports per-slot batch size x loader length here (_pytorch_trial.py:519-540)
                    num_inputs = self.context.get_per_slot_batch_size() * len(cast(Any, self.validation_loader))
        metrics.update(self._reduce_custom(for_training=False))
        if self.comm.size > 1 and any(util.is_overridden(cb.on_validation_end, _callback.PyTorchCallback)
                                      for cb in self.callbacks.values()):
            metrics = self.comm.broadcast_object(metrics if self.is_chief else None, 0)
        for cb in self.callbacks.values():
            cb.on_validation_end(metrics)
        if not self.is_chief:
            return wl.Skipped()
        return {"metrics": {"num_inputs": num_inputs, "validation_metrics": metrics},
                "stop_requested": ctx.get_stop_requested()}

    def _reducer_for(self, name: str) -> Union[_reducer.Reducer, _reducer.MetricReducer]:
        red = self.trial.evaluation_reducer()
        if isinstance(red, dict):
            if name not in red:
                raise KeyError(f"evaluation_reducer() has no entry for validation metric {name!r}")
            return red[name]
        return red

    def _reduce_validation(self, keys: List[str], rows: List[List[Any]]) -> Dict[str, Any]:
        """Per-rank reduce over this rank's batches, then a weighted cross-rank reduce (weights = batches per
        rank for AVG) on the device; MetricReducer entries gather their per-rank state as objects."""
        dev = self.context.device
        n_local = len(rows)
        counts = self.comm.allgather_object(n_local)
        out: Dict[str, Any] = {}
        if sum(counts) == 0:
            return out
        enum_names = [k for k in keys if isinstance(self._reducer_for(k), _reducer.Reducer)]
        if enum_names:
            idx = [keys.index(k) for k in enum_names]
            if n_local:
                mat = torch.stack([torch.stack([cast(torch.Tensor, _to_scalar_tensor(r[i], dev)).to(dev) if
                                                _to_scalar_tensor(r[i], dev) is not None else
                                                torch.as_tensor(np.asarray(r[i], dtype=np.float32).mean(), device=dev)
                                                for i in idx]) for r in rows]).cpu().numpy()
            else:
                mat = np.zeros((0, len(idx)), dtype=np.float32)
            local = np.zeros(len(idx), dtype=np.float32)
            for j, k in enumerate(enum_names):
                red = cast(_reducer.Reducer, self._reducer_for(k))
                if n_local:
                    local[j] = _reducer._reduce_metrics(red, mat[:, j], [1] * n_local if red == _reducer.Reducer.AVG else None)
                else:
                    local[j] = {_reducer.Reducer.MAX: -np.inf, _reducer.Reducer.MIN: np.inf}.get(red, 0.0)
            if self.comm.size > 1:
                by_op: Dict[int, List[int]] = {}
                for j, k in enumerate(enum_names):
                    by_op.setdefault(_REDUCER_OPS[cast(_reducer.Reducer, self._reducer_for(k))], []).append(j)
                final = np.zeros_like(local)
                for op, js in by_op.items():
                    vec = torch.tensor(local[js], dtype=torch.float32, device=dev)
                    res = self.comm.reduce_metrics(vec, op, weights=[float(c) for c in counts] if op == F.RED_AVG else None)
                    final[js] = res.cpu().numpy()
                local = final
            for j, k in enumerate(enum_names):
                out[k] = float(local[j])
        for k in keys:
            red = self._reducer_for(k)
            if isinstance(red, _reducer.MetricReducer):
                red.reset()
                i = keys.index(k)
                for r in rows:
                    v = r[i]
                    red.accumulate(v.detach().cpu() if isinstance(v, torch.Tensor) else v)
                gathered = self.comm.allgather_object(red.per_slot_reduce())
                if self.is_chief:
                    out[k] = red.cross_slot_reduce(gathered)
        return out

    # ---- checkpoints -----------------------------------------------------------------------------------------------------------
    @staticmethod
    def _compact(sd: Any) -> Any:
        """Detach + copy to host so views into flat buckets serialise as their own small storages."""
        if isinstance(sd, torch.Tensor):
            return sd.detach().cpu().clone()
        if isinstance(sd, dict):
            return {k: PyTorchTrialController._compact(v) for k, v in sd.items()}
        if isinstance(sd, (list, tuple)):
            return type(sd)(PyTorchTrialController._compact(v) for v in sd)
        return sd

    def _save(self, path: pathlib.Path) -> wl.Response:
        """``<path>/state_dict.pth`` (cloudpickle) + ``<path>/code/`` — the reference layout (``:713-766``)."""
        path.mkdir(parents=True, exist_ok=True)
        util.write_user_code(path)
        rng_state: Dict[str, Any] = {"cpu_rng_state": torch.random.get_rng_state(), "np_rng_state": np.random.get_state(),
                                     "random_rng_state": random.getstate()}
        if torch.cuda.is_available() and self.context.device.type == "cuda":
            rng_state["gpu_rng_state"] = torch.cuda.get_rng_state(self.context.device)
        checkpoint: Dict[str, Any] = {
            "models_state_dict": [self._compact(m.state_dict()) for m in self.context.models],
            "optimizers_state_dict": [self._compact(o.state_dict()) for o in self.context.optimizers],
            "lr_schedulers_state_dict": [s.state_dict() for s in self.context.lr_schedulers],
            "callbacks": {name: cb.state_dict() for name, cb in self.callbacks.items()},
            "rng_state": rng_state,
        }
        if self.context._loss_scaler is not None:
            checkpoint["amp_state"] = self.context._loss_scaler.state_dict()
        torch.save(checkpoint, str(path / "state_dict.pth"), pickle_module=_pickle_module)
        for cb in self.callbacks.values():
            cb.on_checkpoint_end(str(path))
        return {"framework": f"torch-{torch.__version__}", "format": "cloudpickle"}

    def _load(self) -> None:
        if not self.load_path:
            return
        base = pathlib.Path(self.load_path)
        ckpt_path = None
        for cand in ("state_dict.pth", "determined/state_dict.pth", "pedl/state_dict.pth", "checkpoint.pt"):
            if (base / cand).exists():
                ckpt_path = base / cand
                break
        if ckpt_path is None:
            raise FileNotFoundError(f"no PyTorch checkpoint found under {base}")
        checkpoint = torch.load(str(ckpt_path), map_location="cpu", weights_only=False, pickle_module=pickle)
        ctx = self.context
        if "model_state_dict" in checkpoint:  # legacy single-model layout
            if len(ctx.models) > 1:
                raise RuntimeError("checkpoint holds one model but the trial wraps several")
            ctx.models[0].load_state_dict(checkpoint["model_state_dict"])
        else:
            for m, sd in zip(ctx.models, checkpoint["models_state_dict"]):
                m.load_state_dict(sd)
        if "optimizer_state_dict" in checkpoint:
            ctx.optimizers[0].load_state_dict(checkpoint["optimizer_state_dict"])
        else:
            for o, sd in zip(ctx.optimizers, checkpoint.get("optimizers_state_dict", [])):
                o.load_state_dict(sd)
        for o in ctx.optimizers:
            if isinstance(o, FusedOptimizer) and o.fused:
                o.refresh_shadows()
        if "lr_scheduler" in checkpoint and ctx.lr_schedulers:
            ctx.lr_schedulers[0].load_state_dict(checkpoint["lr_scheduler"])
        else:
            for s, sd in zip(ctx.lr_schedulers, checkpoint.get("lr_schedulers_state_dict", [])):
                s.load_state_dict(sd)
        for name, cb in self.callbacks.items():
            if name in checkpoint.get("callbacks", {}):
                cb.load_state_dict(checkpoint["callbacks"][name])
        if "amp_state" in checkpoint and ctx._loss_scaler is not None:
            ctx._loss_scaler.load_state_dict(checkpoint["amp_state"])
        rng = checkpoint.get("rng_state")
        if rng:
            np.random.set_state(rng["np_rng_state"])
            random.setstate(rng["random_rng_state"])
            torch.random.set_rng_state(rng["cpu_rng_state"])
            if "gpu_rng_state" in rng and ctx.device.type == "cuda":
                torch.cuda.set_rng_state(rng["gpu_rng_state"], device=ctx.device)
