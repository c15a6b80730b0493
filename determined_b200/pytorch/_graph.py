"""CUDA-graph replay of ``train_batch`` (opt-in: ``context.experimental.enable_cuda_graphs()``).

The reference's hot loop pays Python + launch overhead for every kernel of every batch
(``_pytorch_trial.py:348-412``). On B200 a BERT-base micro-batch is ~1000 launches of 5–40 µs kernels, so the
loop is launch-bound unless the whole batch — forward, backward, gradient all-reduce, fused optimizer — is
captured once and replayed. Two graphs are kept: one for batches that only accumulate gradients and one for
batches that close an ``aggregation_frequency`` window (communicate + update).

Contract for a graphed trial: static batch shapes, no host synchronisation and no Python side effects inside
``train_batch`` (it is not re-executed on replay); hyper-parameters that change per step (learning rate, Adam
bias corrections, dropout seed) are read from device memory that is refreshed outside the graph.
"""
from __future__ import annotations

import gc
from typing import Any, Dict, List, Tuple

import torch

from ..ops import _native
from ..ops import functional as F
from ..parallel import fused_optimizer as fo


def _flatten(x: Any, out: List[torch.Tensor]) -> Any:
    """Replace every tensor leaf by its index in ``out``; returns the skeleton."""
    if isinstance(x, torch.Tensor):
        out.append(x)
        return ("__leaf__", len(out) - 1)
    if isinstance(x, dict):
        return {k: _flatten(v, out) for k, v in x.items()}
    if isinstance(x, (list, tuple)):
        return type(x)(_flatten(v, out) for v in x)
    return x


def _unflatten(skel: Any, leaves: List[torch.Tensor]) -> Any:
    if isinstance(skel, tuple) and len(skel) == 2 and skel[0] == "__leaf__":
        return leaves[skel[1]]
    if isinstance(skel, dict):
        return {k: _unflatten(v, leaves) for k, v in skel.items()}
    if isinstance(skel, (list, tuple)):
        return type(skel)(_unflatten(v, leaves) for v in skel)
    return skel


class GraphedTrainBatch:
    def __init__(self, trial: Any, context: Any) -> None:
        self.trial = trial
        self.context = context
        self.seen = 0
        self.graphs: Dict[bool, Tuple[torch.cuda.CUDAGraph, Dict[str, torch.Tensor], List[fo.FusedOptimizer]]] = {}
        self.static_leaves: List[torch.Tensor] = []
        self.skeleton: Any = None
        self.replays = 0
        self.kernels_per_graph: Dict[bool, int] = {}
        self.replayed_launches = 0  # kernels of this library executed through graph replays

    def _eager(self, host_batch: Any, epoch_idx: int, batch_idx: int) -> Any:
        batch = self.context.to_device(host_batch)
        with self.context.autocast():
            out = self.trial.train_batch(batch=batch, epoch_idx=epoch_idx, batch_idx=batch_idx)
        F.advance_step_seed(self.context.device)
        return out

    def _stage(self, host_batch: Any) -> None:
        leaves: List[torch.Tensor] = []
        skel = _flatten(host_batch, leaves)
        if self.skeleton is None:
            self.skeleton = skel
            self.static_leaves = [torch.empty_like(t, device=self.context.device) for t in leaves]
        if len(leaves) != len(self.static_leaves):
            raise RuntimeError("CUDA-graph mode needs the same batch structure on every step")
        for dst, src in zip(self.static_leaves, leaves):
            if dst.shape != src.shape:
                raise RuntimeError(f"CUDA-graph mode needs static batch shapes (got {tuple(src.shape)} after "
                                   f"{tuple(dst.shape)}); use drop_last=True")
            dst.copy_(src, non_blocking=True)

    def run(self, host_batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        ctx = self.context
        update = ctx._should_communicate_and_update()
        warm = ctx.experimental.graph_warmup_batches * max(ctx.aggregation_frequency, 1)
        if self.seen < warm:
            self.seen += 1
            out = self._eager(host_batch, epoch_idx, batch_idx)
            return {"loss": out} if isinstance(out, torch.Tensor) else out
        self._stage(host_batch)
        if update not in self.graphs:
            self._capture(update, epoch_idx, batch_idx)
        else:
            graph, outputs, opts = self.graphs[update]
            for o in opts:
                o.host_prepare()
            graph.replay()
            self.replays += 1
            self.replayed_launches += self.kernels_per_graph[update]
        _, outputs, _ = self.graphs[update]
        return {k: v.clone() for k, v in outputs.items()}

    def _capture(self, update: bool, epoch_idx: int, batch_idx: int) -> None:
        ctx = self.context
        gc.collect()  # drop autograd graphs of eager batches: their grad accumulators are bound to other streams
        torch.cuda.synchronize(ctx.device)
        stepped: List[fo.FusedOptimizer] = []
        observer = stepped.append
        fo.STEP_OBSERVERS.append(observer)
        graph = torch.cuda.CUDAGraph()
        n0 = _native.launch_count
        batch = _unflatten(self.skeleton, self.static_leaves)
        try:
            # thread_local: the DataLoader's prefetch thread pins host memory (cudaHostAlloc) while this thread captures;
            # under the default global mode that call invalidates the capture (cudaErrorStreamCaptureInvalidated)
            with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                with ctx.autocast():
                    out = self.trial.train_batch(batch=batch, epoch_idx=epoch_idx, batch_idx=batch_idx)
                F.advance_step_seed(ctx.device)
        finally:
            fo.STEP_OBSERVERS.remove(observer)
        if isinstance(out, torch.Tensor):
            out = {"loss": out}
        outputs = {k: v.detach() for k, v in out.items() if isinstance(v, torch.Tensor)}
        self.graphs[update] = (graph, outputs, stepped)
        self.kernels_per_graph[update] = _native.launch_count - n0  # counted once at capture = first replay
        # capture only records; run the batch for real
        graph.replay()
        self.replays += 1
