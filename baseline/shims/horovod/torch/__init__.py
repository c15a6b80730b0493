"""``horovod.torch`` stand-in over ``torch.distributed`` / NCCL, for the reference arm at N > 1.

The reference's only data-parallel transport is Horovod (``harness/determined/horovod.py``,
``pytorch/_pytorch_context.py:152-204, 479-545``, ``pytorch/_pytorch_trial.py:35-95``). Horovod itself is a C++/MPI/NCCL
package that cannot be built without a network, so this module re-states the slice of its Python API that the reference
calls, on the transport Horovod would use on one NVSwitch box anyway (NCCL):

* ``init / size / rank / local_rank / local_size``
* ``DistributedOptimizer(opt, named_parameters, compression, backward_passes_per_step)`` with Horovod's behaviour:
  a hook on every parameter's gradient accumulator; after ``backward_passes_per_step`` backward passes the gradient is
  handed to the communication layer; ready gradients are packed into fusion buffers of up to ``HOROVOD_FUSION_THRESHOLD``
  bytes (default 64 MB, Horovod's and the reference's default) and all-reduced asynchronously while backward continues;
  ``synchronize()`` drains, averages and unpacks; ``skip_synchronize()``; ``step()`` synchronizes unless skipped.
* ``broadcast_parameters``, ``broadcast_optimizer_state``, ``broadcast_object``, ``allreduce``, ``Compression``.

One process per GPU, launched by ``torch.distributed.run`` (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* from the
environment) instead of ``horovodrun`` + ssh.
"""
from __future__ import annotations

import contextlib
import io
import os
import pickle
from typing import Any, Dict, Iterable, List, Optional, Tuple

import torch
import torch.distributed as dist

Average, Sum, Adasum = 0, 1, 2
_FUSION_BYTES = int(os.environ.get("HOROVOD_FUSION_THRESHOLD", str(64 << 20)))


def init(*_args: Any, **_kwargs: Any) -> None:
    if dist.is_initialized():
        return
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group("gloo")


def shutdown() -> None:
    if dist.is_initialized():
        dist.destroy_process_group()


def is_initialized() -> bool:
    return dist.is_initialized()


def size() -> int:
    return dist.get_world_size()


def rank() -> int:
    return dist.get_rank()


def local_rank() -> int:
    return int(os.environ.get("LOCAL_RANK", str(rank())))


def local_size() -> int:
    return int(os.environ.get("LOCAL_WORLD_SIZE", str(size())))


def mpi_threads_supported() -> bool:
    return False


def nccl_built() -> bool:
    return True


class _NoneCompressor:
    @staticmethod
    def compress(tensor: torch.Tensor) -> Tuple[torch.Tensor, Any]:
        return tensor, None

    @staticmethod
    def decompress(tensor: torch.Tensor, ctx: Any) -> torch.Tensor:
        return tensor


class _FP16Compressor:
    @staticmethod
    def compress(tensor: torch.Tensor) -> Tuple[torch.Tensor, Any]:
        if tensor.dtype.is_floating_point:
            return tensor.to(torch.float16), tensor.dtype
        return tensor, None

    @staticmethod
    def decompress(tensor: torch.Tensor, ctx: Any) -> torch.Tensor:
        return tensor.to(ctx) if ctx is not None else tensor


class Compression:
    none = _NoneCompressor
    fp16 = _FP16Compressor


def allreduce_(tensor: torch.Tensor, average: Optional[bool] = None, name: Optional[str] = None, op: Optional[int] = None) -> torch.Tensor:
    avg = (op == Average) if op is not None else (True if average is None else average)
    dist.all_reduce(tensor)
    if avg:
        tensor.div_(size())
    return tensor


def allreduce(tensor: torch.Tensor, average: Optional[bool] = None, name: Optional[str] = None,
              compression: Any = Compression.none, op: Optional[int] = None) -> torch.Tensor:
    t, ctx = compression.compress(tensor.clone())
    return compression.decompress(allreduce_(t, average=average, name=name, op=op), ctx)


def allgather(tensor: torch.Tensor, name: Optional[str] = None) -> torch.Tensor:
    out = [torch.empty_like(tensor) for _ in range(size())]
    dist.all_gather(out, tensor)
    return torch.cat(out, dim=0)


def broadcast_(tensor: torch.Tensor, root_rank: int, name: Optional[str] = None) -> torch.Tensor:
    dist.broadcast(tensor, src=root_rank)
    return tensor


def broadcast(tensor: torch.Tensor, root_rank: int, name: Optional[str] = None) -> torch.Tensor:
    return broadcast_(tensor.clone(), root_rank, name)


def broadcast_object(obj: Any, root_rank: int = 0, name: Optional[str] = None) -> Any:
    box = [obj]
    dist.broadcast_object_list(box, src=root_rank)
    return box[0]


def broadcast_parameters(params: Any, root_rank: int) -> None:
    if isinstance(params, dict):
        items = sorted(params.items())
    else:
        items = list(params)
    for _, p in items:
        if isinstance(p, torch.Tensor):
            on_dev = p if (p.is_cuda or not torch.cuda.is_available()) else p.cuda()
            dist.broadcast(on_dev.data if on_dev is p else on_dev, src=root_rank)
            if on_dev is not p:
                p.data.copy_(on_dev)


def broadcast_optimizer_state(optimizer: torch.optim.Optimizer, root_rank: int) -> None:
    """Horovod walks the state dict, wraps scalars into tensors and broadcasts everything from the root. Same effect:
    hyper-parameters and non-tensor state travel as one pickled object, tensors are broadcast in place."""
    sd = optimizer.state_dict()
    meta = {"param_groups": sd["param_groups"],
            "state": {k: {n: (None if isinstance(v, torch.Tensor) else v) for n, v in st.items()} for k, st in sd["state"].items()},
            "shapes": {k: {n: (tuple(v.shape), str(v.dtype)) for n, v in st.items() if isinstance(v, torch.Tensor)}
                       for k, st in sd["state"].items()}}
    meta = broadcast_object(meta, root_rank)
    if rank() != root_rank:
        # materialise missing state so that tensors have somewhere to land
        params = [p for g in optimizer.param_groups for p in g["params"]]
        for k, shapes in meta["shapes"].items():
            st = optimizer.state[params[int(k)]]
            for n, (shape, dtype) in shapes.items():
                if n not in st or not isinstance(st[n], torch.Tensor) or tuple(st[n].shape) != shape:
                    st[n] = torch.zeros(shape, dtype=getattr(torch, dtype.split(".")[-1]), device=params[int(k)].device)
            for n, v in meta["state"][k].items():
                if v is not None:
                    st[n] = v
        for g, saved in zip(optimizer.param_groups, meta["param_groups"]):
            for key, v in saved.items():
                if key != "params":
                    g[key] = v
    params = [p for g in optimizer.param_groups for p in g["params"]]
    for k in sorted(meta["shapes"], key=lambda x: int(x)):
        st = optimizer.state[params[int(k)]]
        for n in sorted(meta["shapes"][k]):
            t = st[n]
            if t.is_cuda or not torch.cuda.is_available():
                dist.broadcast(t, src=root_rank)
            else:
                d = t.cuda()
                dist.broadcast(d, src=root_rank)
                t.copy_(d)


class _DistributedOptimizer(torch.optim.Optimizer):
    def __init__(self, params: Any, named_parameters: Optional[Iterable[Tuple[str, torch.Tensor]]], compression: Any,
                 backward_passes_per_step: int = 1, op: int = Average) -> None:
        super(self.__class__, self).__init__(params)
        self._compression = compression
        named = list(named_parameters) if named_parameters is not None else [
            (f"allreduce.noname.{i}", v) for i, g in enumerate(self.param_groups) for v in g["params"]]
        self._parameter_names = {v: k for k, v in sorted(named)}
        self.backward_passes_per_step = backward_passes_per_step
        self._allreduce_delay = {v: self.backward_passes_per_step for _, v in sorted(named)}
        self.op = op
        self._handles: Dict[torch.Tensor, Any] = {}
        self._grad_accs: List[Any] = []
        self._requires_update = set()
        self._synchronized = False
        self._should_synchronize = True
        self._pending: List[torch.Tensor] = []          # gradients ready for the next fusion buffer
        self._pending_bytes = 0
        self._inflight: List[Tuple[Any, torch.Tensor, List[torch.Tensor], Any]] = []
        if size() > 1:
            self._register_hooks()

    # ---- Horovod's hook machinery -------------------------------------------------------------------------------
    def _register_hooks(self) -> None:
        for group in self.param_groups:
            for p in group["params"]:
                if p.requires_grad:
                    p.grad = p.data.new(p.size()).zero_()
                    self._requires_update.add(p)
                    p_tmp = p.expand_as(p)
                    grad_acc = p_tmp.grad_fn.next_functions[0][0]
                    grad_acc.register_hook(self._make_hook(p))
                    self._grad_accs.append(grad_acc)

    def _make_hook(self, p: torch.Tensor):
        def hook(*ignore: Any) -> None:
            if p in self._handles and self._handles[p] is not None:
                if self._allreduce_delay[p] <= 0:
                    raise AssertionError(
                        "Gradients were computed more than backward_passes_per_step times before call to step(). "
                        "Increase backward_passes_per_step to accumulate gradients locally.")
            assert not p.grad.requires_grad
            assert self._allreduce_delay[p] > 0
            self._allreduce_delay[p] -= 1
            if self._allreduce_delay[p] == 0:
                self._enqueue(p)
        return hook

    # ---- tensor fusion: ready gradients -> <= 64 MB buffers -> asynchronous NCCL all-reduce -----------------------
    def _enqueue(self, p: torch.Tensor) -> None:
        self._handles[p] = True
        self._pending.append(p)
        self._pending_bytes += p.grad.numel() * p.grad.element_size()
        if self._pending_bytes >= _FUSION_BYTES:
            self._flush()

    def _flush(self) -> None:
        if not self._pending:
            return
        grads = [p.grad for p in self._pending]
        flat = torch._utils._flatten_dense_tensors(grads)
        wire, ctx = self._compression.compress(flat)
        work = dist.all_reduce(wire, async_op=True)
        self._inflight.append((work, wire, grads, ctx))
        self._pending, self._pending_bytes = [], 0

    def synchronize(self) -> None:
        if size() > 1:
            missing = self._requires_update - set(self._handles.keys())
            for p in missing:
                self._enqueue(p)
            self._flush()
            for work, wire, grads, ctx in self._inflight:
                work.wait()
                flat = self._compression.decompress(wire, ctx)
                if self.op == Average:
                    flat = flat / size()
                for g, synced in zip(grads, torch._utils._unflatten_dense_tensors(flat, grads)):
                    g.copy_(synced)
            self._inflight = []
            for p in self._handles:
                self._allreduce_delay[p] = self.backward_passes_per_step
            self._handles.clear()
        self._synchronized = True

    @contextlib.contextmanager
    def skip_synchronize(self):
        self._should_synchronize = False
        try:
            yield
        finally:
            self._should_synchronize = True

    def step(self, closure: Any = None):
        if self._should_synchronize:
            if self._synchronized:
                import warnings
                warnings.warn("optimizer.step() called without optimizer.skip_synchronize() context after "
                              "optimizer.synchronize().")
            self.synchronize()
        self._synchronized = False
        return super(self.__class__, self).step(closure)

    def zero_grad(self, *args: Any, **kwargs: Any):
        if self._handles:
            raise AssertionError("optimizer.zero_grad() was called after loss.backward() but before optimizer.step() or "
                                 "optimizer.synchronize().")
        # hooks hold on to p.grad: keep the tensors (torch >= 2 would otherwise drop them to None)
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is not None:
                    p.grad.detach_()
                    p.grad.zero_()


def DistributedOptimizer(optimizer: torch.optim.Optimizer, named_parameters: Any = None, compression: Any = Compression.none,
                         backward_passes_per_step: int = 1, op: int = Average) -> torch.optim.Optimizer:
    """Same construction as Horovod's: a dynamic subclass of the wrapped optimizer's class, so ``isinstance`` checks and
    LR schedulers keep working."""
    cls = type(optimizer.__class__.__name__, (optimizer.__class__,), dict(_DistributedOptimizer.__dict__))
    return cls(optimizer.param_groups, named_parameters, compression, backward_passes_per_step, op)
