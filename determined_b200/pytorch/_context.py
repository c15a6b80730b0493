"""``PyTorchTrialContext`` — model / optimizer wrapping, backward, optimizer stepping.

API parity with reference ``harness/determined/pytorch/_pytorch_context.py:38-571`` (``wrap_model``,
``wrap_optimizer``, ``wrap_lr_scheduler``, ``configure_apex_amp``, ``backward``, ``step_optimizer``,
``to_device``, ``is_epoch_start/end``, deprecated ``get_model/get_optimizer/get_lr_scheduler``).

What is different underneath (B200-first):

* no Horovod — one process per GPU, identity from the launcher; ``wrap_optimizer`` returns a
  :class:`~determined_b200.parallel.fused_optimizer.FusedOptimizer` whose gradient bucket lives in the
  peer-mapped symmetric arena, so ``step_optimizer`` is: fused all-reduce kernel (sum + average over ranks and
  aggregation_frequency + loss-scale unscale + non-finite flag + grad-norm partial) → fused update kernel
  (clip coefficient, skip-on-overflow, zero-grad, bf16 shadow refresh);
* no apex — ``configure_apex_amp`` / ``optimizations.mixed_precision`` map to native bf16 autocast (O1/O2/O3,
  no loss scaling needed) or fp16 with a device-resident dynamic loss scaler; the reference's restriction
  ``aggregation_frequency == 1 and num_losses == 1`` under AMP + distributed (``:336-352``) is lifted;
* ``native_parallel: true`` keeps the reference's meaning (``:135-144``): one process over the container's GPUs,
  ``wrap_model`` returns ``nn.DataParallel(model)``, the process sees the whole global batch, and
  ``aggregation_frequency`` must be 1. The stock torch optimizer is kept on that path (replica gradients are
  already reduced onto device 0 by DataParallel; there is no cross-process collective to fuse).
"""
from __future__ import annotations

import contextlib
import logging
from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence, Tuple, Union

import torch
import torch.nn as nn

from .._env import DistributedContext, EnvContext
from .._train_context import TrialContext
from ..parallel.comm import Communicator
from ..parallel.fused_optimizer import FusedOptimizer
from . import _callback, _data, _reducer
from ._lr_scheduler import LRScheduler

log = logging.getLogger("determined.pytorch")


class _LossScaler:
    """fp16 dynamic loss scaling with all state on the device (K5)."""

    def __init__(self, device: torch.device, init_scale: float = 65536.0, growth: float = 2.0, backoff: float = 0.5,
                 interval: int = 2000) -> None:
        self.state = torch.tensor([init_scale, 0.0, 1.0 / init_scale], dtype=torch.float32, device=device)
        self.growth, self.backoff, self.interval = growth, backoff, interval

    def scale_tensor(self) -> torch.Tensor:
        return self.state[0:1]

    def state_dict(self) -> Dict[str, Any]:
        return {"state": self.state.detach().cpu()}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.state.copy_(sd["state"])


class _Experimental:
    """``context.experimental`` — B200-specific opt-ins, plus the data-layer decorators the reference hangs here
    (``harness/determined/_experimental_context.py``: ``cache_train_dataset`` / ``cache_validation_dataset``)."""

    def __init__(self, parent: "PyTorchTrialContext") -> None:
        self._parent = parent
        self._data_layer: Any = None
        self.cuda_graphs = False
        self.graph_warmup_batches = 3
        self._step_dev: Optional[torch.Tensor] = None
        self._step_host: Optional[torch.Tensor] = None

    def _layer(self) -> Any:
        if self._data_layer is None:
            from .. import data_layer

            self._data_layer = data_layer.from_context(self._parent)
        return self._data_layer

    def cache_train_dataset(self, dataset_id: str, dataset_version: str, shuffle: bool = False,
                            skip_shuffle_at_epoch_end: bool = False) -> Any:
        """Decorator: build the training data set once per (id, version) and serve the cached copy afterwards — to this trial
        positioned at its resume offset / shard (``determined_b200/data_layer.py``)."""
        return self._layer().cache_train_dataset(dataset_id, dataset_version, shuffle, skip_shuffle_at_epoch_end)

    def cache_validation_dataset(self, dataset_id: str, dataset_version: str, shuffle: bool = False) -> Any:
        return self._layer().cache_validation_dataset(dataset_id, dataset_version, shuffle)

    def step_info(self) -> torch.Tensor:
        """Device tensor ``[epoch_idx, batch_idx]`` (fp32) of the batch being trained, refreshed by the controller BEFORE
        every ``train_batch`` — also before graph replays, where the Python arguments of ``train_batch`` are frozen at
        their capture-time values. Schedules that depend on training progress (drop-path ramps, loss-weight warm-ups)
        read it instead of ``epoch_idx`` when CUDA graphs are on."""
        if self._step_dev is None:
            dev = self._parent.device
            self._step_dev = torch.zeros(2, dtype=torch.float32, device=dev)
            self._step_host = torch.zeros(2, dtype=torch.float32, pin_memory=(dev.type == "cuda"))
        return self._step_dev

    def _set_step_info(self, epoch_idx: int, batch_idx: int) -> None:
        if self._step_dev is None:
            return
        self._step_host[0], self._step_host[1] = float(epoch_idx), float(batch_idx)
        self._step_dev.copy_(self._step_host, non_blocking=True)

    def enable_cuda_graphs(self, warmup_batches: int = 3) -> None:
        """Capture ``train_batch`` (forward, backward, all-reduce, optimizer) into CUDA graphs after a few eager
        batches and replay it afterwards. Requires static shapes and a ``train_batch`` without host syncs."""
        self.cuda_graphs = True
        self.graph_warmup_batches = max(int(warmup_batches), 3)

    def enable_grad_overlap(self, chunk_mb: int = 32, reserved_sms: int = 0) -> None:
        """Overlap the gradient all-reduce with backward (multi-slot trials on GPUs): slices of the flat gradient
        bucket are reduced on a side stream as soon as backward has finished writing them. Only for trials that make
        ONE ``context.backward`` call per optimizer step — see :mod:`determined_b200.parallel.overlap`."""
        p = self._parent
        if p.comm.size > 1 and p.device.type == "cuda" and getattr(p.comm, "p2p", False) and p._overlap is None:
            from ..parallel.overlap import GradOverlap

            import os

            reserved = int(os.environ.get("DET_OVERLAP_SMS", reserved_sms))
            p._overlap = GradOverlap(p.comm, chunk_bytes=int(os.environ.get("DET_OVERLAP_CHUNK_MB", chunk_mb)) << 20,
                                     reserved_sms=reserved, comm_blocks=int(os.environ.get("DET_OVERLAP_BLOCKS", 64)),
                                     cta_threads=int(os.environ.get("DET_OVERLAP_CTA_THREADS", 128)))
            for o in p.optimizers:
                if isinstance(o, FusedOptimizer):
                    p._overlap.add_optimizer(o)


class PyTorchTrialContext(TrialContext):
    def __init__(self, env: EnvContext, dist: Optional[DistributedContext] = None,
                 comm: Optional[Communicator] = None) -> None:
        super().__init__(env, dist)
        self.comm: Communicator = comm or Communicator()
        self.models: List[nn.Module] = []
        self.optimizers: List[torch.optim.Optimizer] = []
        self.lr_schedulers: List[LRScheduler] = []
        self.reducers: Dict[str, Tuple[_reducer.MetricReducer, bool, bool]] = {}
        self._epoch_len: Optional[int] = None
        self._current_batch_idx: Optional[int] = None
        self._loss_ids: Dict[int, int] = {}
        self._last_backward_batch_idx: Optional[int] = None
        self._main_model = nn.Module()
        self._use_amp = False
        self._amp_dtype: Optional[torch.dtype] = None
        self._loss_scaler: Optional[_LossScaler] = None
        self._scaler_divisor_applied = False
        self.experimental = _Experimental(self)
        self._overlap: Any = None     # parallel.overlap.GradOverlap when enabled
        self._ag_stream: Any = None   # side stream of the pipelined all-gather (created on first use)
        self.warned_types: set = set()

        opt_cfg = env.experiment_config.get_optimizations_config()
        self.aggregation_frequency = int(opt_cfg.get("aggregation_frequency", 1))
        self.average_aggregated_gradients = bool(opt_cfg.get("average_aggregated_gradients", True))
        self.gradient_compression = bool(opt_cfg.get("gradient_compression", False))
        if self.gradient_compression and hasattr(self.comm, "compress_gradients"):
            # Inside one NVSwitch domain the fused fp32 all-reduce is ~4 % of a BERT-base step and stays fp32; the flag
            # takes effect on the NCCL path (trials spanning nodes, or DET_DISABLE_P2P=1).
            self.comm.compress_gradients = True
        if self.aggregation_frequency < 1:
            raise ValueError("optimizations.aggregation_frequency must be >= 1")
        self.device = self._init_device()
        mp = opt_cfg.get("mixed_precision", "O0")
        if mp != "O0":
            self._configure_mixed_precision(mp)

    # ---- device / identity -----------------------------------------------------------------------
    def _init_device(self) -> torch.device:
        if self.env.use_gpu and torch.cuda.is_available():
            idx = self.distributed.get_local_rank() if self.distributed.get_size() > 1 else 0
            idx = idx % max(torch.cuda.device_count(), 1)
            torch.cuda.set_device(idx)
            return torch.device("cuda", idx)
        return torch.device("cpu")

    @property
    def n_gpus(self) -> int:
        return len(self.env.container_gpus)

    def _configure_mixed_precision(self, level: str, dtype: Optional[torch.dtype] = None) -> None:
        if level not in ("O0", "O1", "O2", "O3"):
            raise ValueError(f"mixed_precision must be one of O0..O3, got {level!r}")
        if level == "O0":
            self._use_amp = False
            return
        self._use_amp = True
        self._amp_dtype = dtype or torch.bfloat16
        if self._amp_dtype == torch.float16:
            # the scaler also exists on CPU (autocast stays off there): the skip-on-overflow protocol between ranks is
            # host-side logic and is tested with gloo
            self._loss_scaler = _LossScaler(self.device)

    @contextlib.contextmanager
    def autocast(self):
        """Autocast region matching the trial's mixed-precision setting (no-op for O0 or on CPU)."""
        if self._use_amp and self.device.type == "cuda":
            with torch.autocast("cuda", dtype=self._amp_dtype):
                yield
        else:
            yield

    # ---- wrapping ------------------------------------------------------------------------------------
    def wrap_model(self, model: nn.Module) -> nn.Module:
        """Move the model to this trial's device and register it (checkpointing, broadcast)."""
        if not isinstance(model, nn.Module):
            raise TypeError("wrap_model expects a torch.nn.Module")
        model = model.to(self.device)
        if self.native_parallel_gpus() > 1:
            if self.aggregation_frequency != 1:
                raise ValueError("aggregation_frequency > 1 needs the multi-process path: drop `resources.native_parallel` "
                                 "(reference _pytorch_context.py:136-142)")
            own = sorted({type(m).__name__ for m in model.modules() if type(m).__module__.startswith("determined_b200.ops")})
            if own:
                # the in-tree modules write parameter gradients in place from their backward (into the flat bucket) instead
                # of returning them through autograd; nn.DataParallel's replicas are non-leaf copies, so those writes would
                # never reach the real parameters
                raise ValueError("resources.native_parallel (nn.DataParallel) works with stock torch.nn modules; this model "
                                 f"uses determined_b200.ops modules ({', '.join(own[:4])}) — drop native_parallel and use the "
                                 "default one-process-per-GPU data parallelism")
            model = nn.DataParallel(model, device_ids=list(range(self.native_parallel_gpus())))
            log.debug("model wrapped for native parallel training")
        self._main_model.add_module(f"model_{len(self.models)}", model)
        self.models.append(model)
        return model

    def native_parallel_gpus(self) -> int:
        """GPUs this one process drives under ``resources.native_parallel`` (0 when the path is off)."""
        import os

        if not self.env.experiment_config.native_parallel_enabled() or self.comm.size > 1 or self.device.type != "cuda":
            return 0
        want = int(os.environ.get("DET_NATIVE_PARALLEL_GPUS", "0") or 0) or max(self.env.experiment_config.slots_per_trial(), 1)
        return min(want, torch.cuda.device_count())

    def wrap_optimizer(self, optimizer: torch.optim.Optimizer, backward_passes_per_step: int = 1,
                       fused: Optional[bool] = None) -> torch.optim.Optimizer:
        """Adopt a torch optimizer. On CUDA this flattens its parameters into fp32 buckets inside the symmetric
        arena and swaps in the fused update; elsewhere the optimizer is used as is. ``backward_passes_per_step``
        is accepted for API compatibility (several ``backward`` calls per batch accumulate into the same
        bucket; communication happens once, in ``step_optimizer``). ``fused=False`` (an extension) keeps the stock
        torch optimizer and per-tensor collectives — the comparison arm of the benchmarks."""
        if not isinstance(optimizer, torch.optim.Optimizer):
            raise TypeError("wrap_optimizer expects a torch.optim.Optimizer")
        if backward_passes_per_step < 1:
            raise ValueError("backward_passes_per_step must be >= 1")
        if self.device.type == "cuda" and fused is not False and self.native_parallel_gpus() <= 1:
            dev = self.device

            def alloc(n: int, dtype: torch.dtype, kind: str) -> torch.Tensor:
                return self.comm.alloc(n, dtype, kind, dev)

            wrapped: torch.optim.Optimizer = FusedOptimizer(optimizer, alloc=alloc)
            if self._overlap is not None:
                self._overlap.add_optimizer(wrapped)
        else:
            wrapped = optimizer
        self.optimizers.append(wrapped)
        return wrapped

    def wrap_lr_scheduler(self, lr_scheduler: Any, step_mode: LRScheduler.StepMode, frequency: int = 1) -> Any:
        """Register a torch LR scheduler together with when the harness should step it. Returns the scheduler
        (so ``MANUAL_STEP`` users can call ``.step()`` themselves)."""
        opt = getattr(lr_scheduler, "optimizer", None)
        if opt is not None and not any(opt is o or opt is getattr(o, "inner", None) for o in self.optimizers):
            raise ValueError("wrap the optimizer with context.wrap_optimizer() before its LR scheduler")
        wrapped = LRScheduler(lr_scheduler, step_mode, frequency)
        self.lr_schedulers.append(wrapped)
        return lr_scheduler

    def wrap_reducer(self, reducer: Union[_reducer.MetricReducer, Callable[[List[Any]], Any]], name: Optional[str] = None,
                     for_training: bool = True, for_validation: bool = True) -> _reducer.MetricReducer:
        """Register a custom cross-rank metric (extension; see :class:`MetricReducer`). Call
        ``reducer.accumulate(value)`` from ``train_batch`` / ``evaluate_batch``; the harness reduces it at the
        end of the workload and reports it under ``name``."""
        if not isinstance(reducer, _reducer.MetricReducer):
            if not callable(reducer):
                raise TypeError("reducer must be a MetricReducer or a function over the list of values")
            reducer = _reducer.make_metric_reducer(reducer)
        name = name or f"reducer_{len(self.reducers)}"
        if name in self.reducers:
            raise ValueError(f"a reducer named {name!r} is already registered")
        self.reducers[name] = (reducer, for_training, for_validation)
        return reducer

    def configure_apex_amp(self, models: Union[nn.Module, List[nn.Module]],
                           optimizers: Union[torch.optim.Optimizer, List[torch.optim.Optimizer]],
                           enabled: Optional[bool] = True, opt_level: Optional[str] = "O1",
                           cast_model_type: Optional[torch.dtype] = None, patch_torch_functions: Optional[bool] = None,
                           keep_batchnorm_fp32: Optional[Union[bool, str]] = None,
                           master_weights: Optional[bool] = None, loss_scale: Optional[Union[float, str]] = None,
                           cast_model_outputs: Optional[torch.dtype] = None, num_losses: Optional[int] = 1,
                           verbosity: Optional[int] = 1, min_loss_scale: Optional[float] = None,
                           max_loss_scale: Optional[float] = 2.0 ** 24) -> Tuple:
        """Same signature as the reference (which forwards to ``apex.amp.initialize``). Here it switches on
        native mixed precision: parameters stay fp32 masters, compute runs in bf16 (or fp16 when
        ``cast_model_type=torch.float16``, with dynamic loss scaling kept on the device). Works with any
        ``aggregation_frequency`` and any number of losses, distributed or not."""
        if enabled:
            dtype = torch.float16 if cast_model_type == torch.float16 else torch.bfloat16
            self._configure_mixed_precision(opt_level or "O1", dtype)
            if self._loss_scaler is not None and isinstance(loss_scale, (int, float)):
                self._loss_scaler.state[0] = float(loss_scale)
        return models, optimizers

    # ---- batch bookkeeping ---------------------------------------------------------------------------
    def _set_current_batch(self, batch_idx: int, epoch_len: Optional[int]) -> None:
        self._current_batch_idx = batch_idx
        self._epoch_len = epoch_len

    def current_train_batch(self) -> int:
        if self._current_batch_idx is None:
            raise RuntimeError("no training batch is in progress")
        return self._current_batch_idx

    def _should_communicate_and_update(self) -> bool:
        if self._current_batch_idx is None:
            raise RuntimeError("step_optimizer / backward called outside train_batch")
        return (self._current_batch_idx + 1) % self.aggregation_frequency == 0

    def is_epoch_start(self) -> bool:
        if self._current_batch_idx is None:
            raise RuntimeError("is_epoch_start() is only valid during training")
        if self._epoch_len is None:
            raise RuntimeError("is_epoch_start() needs records_per_epoch or a sized training loader")
        return self._current_batch_idx % self._epoch_len == 0

    def is_epoch_end(self) -> bool:
        if self._current_batch_idx is None:
            raise RuntimeError("is_epoch_end() is only valid during training")
        if self._epoch_len is None:
            raise RuntimeError("is_epoch_end() needs records_per_epoch or a sized training loader")
        return self._current_batch_idx % self._epoch_len == self._epoch_len - 1

    def to_device(self, data: _data.TorchData) -> _data.TorchData:
        return _data.to_device(data, self.device, self.warned_types)

    # ---- backward / step -------------------------------------------------------------------------------
    def backward(self, loss: torch.Tensor, gradient: Optional[torch.Tensor] = None, retain_graph: bool = False,
                 create_graph: bool = False) -> None:
        """``loss.backward()`` with loss scaling when fp16 is on. Gradients accumulate into the optimizer's
        bucket; nothing is communicated until ``step_optimizer``. May be called several times per batch."""
        if self._loss_scaler is not None:
            loss = loss * self._loss_scaler.scale_tensor()
        if self._overlap is not None and self._should_communicate_and_update():
            scaler = self._loss_scaler
            self._overlap.arm(1.0 / self._grad_divisor(), scaler.state[2:3] if scaler is not None else None,
                              (lambda o: o.found_inf_dev) if scaler is not None else None)

        loss.backward(gradient=gradient, retain_graph=retain_graph, create_graph=create_graph)

    def _grad_divisor(self) -> float:
        d = float(self.comm.size)
        if self.average_aggregated_gradients:
            d *= float(self.aggregation_frequency)
        return d

    def step_optimizer(self, optimizer: torch.optim.Optimizer,
                       clip_grads: Optional[Callable[[Iterable[torch.Tensor]], None]] = None,
                       auto_zero_grads: bool = True) -> None:
        """Communicate, average, (clip), update, (zero) — only on batches that close an aggregation window.

        ``clip_grads`` receives the parameters with unscaled, averaged gradients, as in the reference
        (``_pytorch_context.py:524-536``). Passing ``det.pytorch.ClipGradsL2Norm(x)`` keeps the whole step on the
        fused path (norm computed by the all-reduce / bucket kernels, coefficient folded into the update)."""
        if optimizer not in self.optimizers:
            raise ValueError("step_optimizer() got an optimizer that was not wrapped with wrap_optimizer()")
        if not self._should_communicate_and_update():
            return
        divisor = self._grad_divisor()
        fused = isinstance(optimizer, FusedOptimizer) and optimizer.fused
        max_norm = 0.0
        slow_clip = None
        if clip_grads is not None:
            if isinstance(clip_grads, _callback.ClipGradsL2Norm) and fused:
                max_norm = clip_grads.clip_value
            else:
                slow_clip = clip_grads
        if fused:
            self._step_fused(optimizer, divisor, max_norm, slow_clip, auto_zero_grads)
        else:
            self._step_plain(optimizer, divisor, slow_clip, auto_zero_grads)

    def _step_fused(self, opt: FusedOptimizer, divisor: float, max_norm: float, slow_clip, auto_zero: bool) -> None:
        scaler = self._loss_scaler
        inv_scale_dev = scaler.state[2:3] if scaler is not None else None  # 1/(loss_scale) maintained on device
        if scaler is not None:
            # state[2] holds 1/(scale*divisor_ref); keep divisor_ref == 1 and fold the divisor on the host side
            pass
        found_inf = opt.found_inf_dev if scaler is not None else None
        applied_all = True
        # the collective also leaves the sum of squares of the whole reduced bucket (grad-norm clipping) when asked to
        want_sumsq = max_norm > 0 and self.comm.size > 1 and slow_clip is None
        if self._overlap is not None and self._overlap.armed_for(opt):
            # slices were reduced on the side stream while backward ran; flush the rest and join
            applied_all = self._overlap.finish(opt)
            want_sumsq = want_sumsq and self._overlap.collects_sumsq
        else:
            if want_sumsq:
                opt._scratch[0:1].zero_()
            pipelines: Dict[int, Any] = {}
            for g in opt.grad_buckets():
                bounds = self.comm.pipeline_plan(g) if (slow_clip is None and auto_zero) else None
                if bounds is not None:
                    # reduce-scatter + statistics now; the all-gather travels slice by slice on a side stream while the
                    # optimizer updates the slices that have already landed (csrc/comm/fused_allreduce.cu, "Pipelined form")
                    self.comm.reduce_scatter_stats(g, len(bounds), 1.0 / divisor, scale_dev=inv_scale_dev, found_inf_dev=found_inf,
                                                   sumsq_dev=opt._scratch[0:1] if want_sumsq else None)
                    pipelines[g.data_ptr()] = (bounds, self._start_allgather(g, len(bounds)))
                    continue
                applied = self.comm.allreduce_grad_bucket(g, 1.0 / divisor, scale_dev=inv_scale_dev, found_inf_dev=found_inf,
                                                          sumsq_dev=opt._scratch[0:1] if want_sumsq else None)
                applied_all = applied_all and applied
            opt.slice_pipelines = pipelines
        sumsq_ready = bool(want_sumsq and applied_all)
        if applied_all and self.comm.size > 1:
            opt.grad_scale_dev.fill_(1.0)
        else:
            # single process (or NCCL fallback): averaging / unscale rides along in the update kernel
            if inv_scale_dev is not None:
                torch.mul(inv_scale_dev, 1.0 / divisor, out=opt.grad_scale_dev)
            else:
                opt.grad_scale_dev.fill_(1.0 / divisor)
        if slow_clip is not None:
            # arbitrary user callback: materialise unscaled, averaged gradients first
            for g in opt.grad_buckets():
                g.mul_(opt.grad_scale_dev)
            opt.grad_scale_dev.fill_(1.0)
            slow_clip([p for grp in opt.param_groups for p in grp["params"]])
        if scaler is not None and found_inf is not None and not (applied_all and self.comm.size > 1):
            # the collective did not scan for inf/NaN (single process, or the library all-reduce fallback: several nodes,
            # DET_DISABLE_P2P, no arena): scan here, then agree across ranks so that every rank skips or steps together
            from ..ops import functional as F

            sc = opt._scratch
            sc[0:1].zero_()
            for g in opt.grad_buckets():
                F.grad_sumsq_finite(g, opt.grad_scale_dev, sc[0:1], found_inf)
            if self.comm.size > 1:
                self.comm.allreduce_max_int(found_inf)
        try:
            opt.step(max_grad_norm=max_norm, zero_grad=auto_zero, sumsq_ready=sumsq_ready)
        finally:
            opt.slice_pipelines = {}
        if scaler is not None:
            from ..ops import functional as F

            F.loss_scale_update(scaler.state, opt.found_inf_dev, scaler.growth, scaler.backoff, scaler.interval, 1.0)

    def _start_allgather(self, g: torch.Tensor, chunks: int) -> Callable[[int], None]:
        """Launch this rank's sends of a bucket's pipelined all-gather on the side stream (chunk 0, 1, ... back to back) and
        return what the optimizer runs in front of each chunk's update: a stream dependency on the own send of that chunk
        (which reads the chunk) plus a one-warp wait for the other ranks' shares of it."""
        comm = self.comm
        main = torch.cuda.current_stream(self.device)
        if self._ag_stream is None:
            self._ag_stream = torch.cuda.Stream(self.device, priority=-1)
        side = self._ag_stream
        side.wait_stream(main)                 # the reduce-scatter (and everything before it) is in front of the sends
        sent = []
        with torch.cuda.stream(side):
            for c in range(chunks):
                comm.allgather_send(g, chunks, c)
                ev = torch.cuda.Event()
                ev.record(side)
                sent.append(ev)

        def before_chunk(c: int) -> None:
            main.wait_event(sent[c])           # the last one also joins the side stream back (ends its capture in a graph)
            comm.wait_chunk(c)
        return before_chunk

    def _step_plain(self, opt: torch.optim.Optimizer, divisor: float, slow_clip, auto_zero: bool) -> None:
        params = [p for grp in opt.param_groups for p in grp["params"] if p.grad is not None]
        if self.comm.size > 1:
            self.comm.allreduce_tensors([p.grad for p in params], 1.0 / divisor)
        elif divisor != 1.0:
            for p in params:
                p.grad.div_(divisor)
        skip = False
        if self._loss_scaler is not None:
            inv = 1.0 / float(self._loss_scaler.state[0])
            for p in params:
                p.grad.mul_(inv)
            # stock-optimizer path: overflow check on the host (gradients are already summed over ranks, so every rank
            # sees the same values; the MAX reduce guards against rank-local non-determinism in the collective)
            bad = torch.zeros(1, dtype=torch.int32, device=self.device)
            for p in params:
                if not bool(torch.isfinite(p.grad).all()):
                    bad.fill_(1)
                    break
            if self.comm.size > 1:
                self.comm.allreduce_max_int(bad)
            skip = bool(bad.item())
            st = self._loss_scaler.state
            if skip:
                st[0] = torch.clamp(st[0] * self._loss_scaler.backoff, min=1.0)
                st[1] = 0.0
            else:
                st[1] = st[1] + 1.0
                if float(st[1]) >= self._loss_scaler.interval:
                    st[0] = st[0] * self._loss_scaler.growth
                    st[1] = 0.0
            st[2] = 1.0 / st[0]
        if slow_clip is not None and not skip:
            slow_clip([p for grp in opt.param_groups for p in grp["params"]])
        if not skip:
            opt.step()
        if auto_zero:
            opt.zero_grad()

    # ---- deprecated single-object getters ------------------------------------------------------------------
    def get_model(self) -> nn.Module:
        if len(self.models) != 1:
            raise RuntimeError("get_model() requires exactly one wrapped model")
        return self.models[0]

    def get_optimizer(self) -> torch.optim.Optimizer:
        if len(self.optimizers) != 1:
            raise RuntimeError("get_optimizer() requires exactly one wrapped optimizer")
        return self.optimizers[0]

    def get_lr_scheduler(self) -> Optional[LRScheduler]:
        if len(self.lr_schedulers) > 1:
            raise RuntimeError("get_lr_scheduler() requires at most one wrapped LR scheduler")
        return self.lr_schedulers[0] if self.lr_schedulers else None
