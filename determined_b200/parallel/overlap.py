"""GradOverlap — start the all-reduce of a slice of the flat gradient bucket as soon as backward has finished writing
it, on a side stream, while the rest of backward keeps computing.

This is what Horovod's per-parameter hooks + background fusion thread do for the reference
(``hvd.DistributedOptimizer`` created at ``harness/determined/pytorch/_pytorch_context.py:192-198``, drained by
``optimizer.synchronize()`` at ``:462`` / ``:522``). Differences that come from the bucket design:

* gradients already live in ONE flat fp32 bucket per parameter group, in registration order, so a "fusion buffer" is
  just a contiguous slice ``[start, start + n)`` — nothing is copied or packed;
* readiness comes from the autograd nodes of :mod:`determined_b200.ops` (``note_param_use`` in forward,
  ``notify_grads_written`` at the end of backward): a chunk is launched when every pending use of every parameter in it
  has reported — a parameter used twice in one graph (a GAN discriminator run on real and fake batches) counts twice;
* the collective is the fused peer-memory kernel (sum + scale + unscale + non-finite flag) launched as many SMALL CTAs
  (128 threads, 8 K registers, ~130 B of shared memory) on a high-priority side stream: a CTA of that size fits next to
  a resident GEMM CTA (576 threads x 96 registers, 226 KB of shared memory) on the same SM, so backward keeps all 148
  SMs and the collective's loads and switch reductions ride along in its issue slots. Reserving SMs instead
  (``reserved_sms > 0`` -> ``det_set_sm_limit``) remains possible but costs the persistent GEMMs a whole wave.

Contract: between ``arm()`` (called by ``context.backward`` on the batch that closes an aggregation window) and
``finish(optimizer)`` (called by ``context.step_optimizer``) every gradient of that optimizer must be produced by ONE
``backward`` call. Trials that call ``backward`` several times per optimizer step must not enable overlap.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Tuple

import torch

from ..ops import functional as F
from ..ops import modules as M
from .fused_optimizer import FusedOptimizer


class _Chunk:
    __slots__ = ("opt", "bucket", "start", "n", "params", "pending", "launched")

    def __init__(self, opt: FusedOptimizer, bucket: torch.Tensor, start: int, n: int, params: List[torch.Tensor]) -> None:
        self.opt, self.bucket, self.start, self.n, self.params = opt, bucket, start, n, params
        self.pending = 0
        self.launched = False


class GradOverlap:
    def __init__(self, comm: Any, chunk_bytes: int = 32 << 20, reserved_sms: int = 0, comm_blocks: int = 64,
                 cta_threads: int = 128) -> None:
        self.comm = comm
        self.chunk_elems = max(chunk_bytes // 4, 1)
        self.reserved_sms, self.comm_blocks, self.cta_threads = reserved_sms, comm_blocks, cta_threads
        self.chunks: List[_Chunk] = []
        self.by_param: Dict[int, _Chunk] = {}
        self.known_opts: List[FusedOptimizer] = []
        self.open = False
        self.args: Tuple[float, Optional[torch.Tensor], Optional[Callable[[FusedOptimizer], Optional[torch.Tensor]]]] = (1.0, None, None)
        self.stream: Optional[torch.cuda.Stream] = None
        self.launch_log: List[Tuple[int, int]] = []   # (start, n) in launch order, for tests / diagnostics
        self.collects_sumsq = True   # every slice's collective adds the slice's sum of squares to opt._scratch[0]
        if self._on_notify not in M.GRAD_OBSERVERS:
            M.GRAD_OBSERVERS.append(self._on_notify)

    def close(self) -> None:
        if self._on_notify in M.GRAD_OBSERVERS:
            M.GRAD_OBSERVERS.remove(self._on_notify)

    # ---- layout --------------------------------------------------------------------------------------------------
    def add_optimizer(self, opt: FusedOptimizer) -> None:
        """Cut each of the optimizer's gradient buckets into chunks of whole parameters, in bucket order."""
        if any(o is opt for o in self.known_opts) or not opt.fused:
            return
        self.known_opts.append(opt)
        for grp in opt.groups:
            cur: List[torch.Tensor] = []
            cur_start = 0
            for i, (p, off) in enumerate(zip(grp.params, grp.offsets)):
                if not cur:
                    cur_start = off
                cur.append(p)
                end = grp.offsets[i + 1] if i + 1 < len(grp.params) else grp.total
                if end - cur_start >= self.chunk_elems or i + 1 == len(grp.params):
                    ch = _Chunk(opt, grp.flat_g, cur_start, end - cur_start, cur)
                    self.chunks.append(ch)
                    for q in cur:
                        self.by_param[id(q)] = ch
                    cur = []

    # ---- window --------------------------------------------------------------------------------------------------
    def arm(self, scale: float, scale_dev: Optional[torch.Tensor], found_inf_of: Optional[Callable[[FusedOptimizer], Optional[torch.Tensor]]]) -> None:
        """Called right before ``loss.backward()`` on a communicating batch: (re)compute what every unlaunched chunk is
        still waiting for from the forward-side use counts."""
        self.args = (scale, scale_dev, found_inf_of)
        if not self.open:
            for o in self.known_opts:   # grad-norm accumulator of this window (before any slice is launched)
                o._scratch[0:1].zero_()
        for ch in self.chunks:
            if not ch.launched:
                # Early launch only when EVERY trainable parameter of the chunk is tracked by a notifying autograd node
                # in this graph. A parameter with no recorded use may be unused — or may get its gradient from plain
                # torch autograd, which never notifies; such chunks wait for finish().
                uses = [M.pending_uses(p) for p in ch.params if p.requires_grad]
                ch.pending = sum(uses) if uses and all(u > 0 for u in uses) else -1
        if not self.open:
            self.open = True
            dev = self.chunks[0].bucket.device if self.chunks else None
            if dev is not None and dev.type == "cuda":
                if self.stream is None:
                    self.stream = torch.cuda.Stream(device=dev, priority=-1)
                if self.reserved_sms > 0:
                    F.set_sm_limit(max(F.device_sm_count(dev) - self.reserved_sms, 8))

    def _on_notify(self, params: List[torch.Tensor]) -> None:
        if not self.open:
            return
        for p in params:
            ch = self.by_param.get(id(p))
            if ch is None or ch.launched or ch.pending <= 0:
                continue
            ch.pending -= 1
            if ch.pending == 0:
                self._launch(ch)

    def _launch(self, ch: _Chunk, overlapped: bool = True) -> None:
        scale, scale_dev, found_inf_of = self.args
        found_inf = found_inf_of(ch.opt) if found_inf_of is not None else None
        view = ch.bucket[ch.start:ch.start + ch.n]
        sumsq = ch.opt._scratch[0:1] if self.collects_sumsq else None
        ch.launched = True
        self.launch_log.append((ch.start, ch.n))
        if self.stream is not None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            self.stream.wait_event(ev)
            with torch.cuda.stream(self.stream):
                # small grid while backward is still running next to it; full width for the final flush
                if overlapped:
                    self.comm.allreduce_grad_bucket(view, scale, scale_dev=scale_dev, found_inf_dev=found_inf, sumsq_dev=sumsq,
                                                    max_blocks=self.comm_blocks, cta_threads=self.cta_threads)
                else:
                    self.comm.allreduce_grad_bucket(view, scale, scale_dev=scale_dev, found_inf_dev=found_inf, sumsq_dev=sumsq)
        else:
            self.comm.allreduce_grad_bucket(view, scale, scale_dev=scale_dev, found_inf_dev=found_inf, sumsq_dev=sumsq)

    def armed_for(self, opt: FusedOptimizer) -> bool:
        return self.open and any(o is opt for o in self.known_opts)

    def finish(self, opt: FusedOptimizer) -> bool:
        """Launch whatever of ``opt`` is still unlaunched (parameters that got no gradient this window, or whose
        gradients came from plain torch autograd), wait for the side stream, and reopen the optimizer's chunks for
        the next window. Returns True when every slice was reduced by the fused collective."""
        mine = [ch for ch in self.chunks if ch.opt is opt]
        for ch in mine:
            if not ch.launched:
                self._launch(ch, overlapped=False)
        if self.stream is not None:
            torch.cuda.current_stream().wait_stream(self.stream)
        for ch in mine:
            ch.launched = False
            ch.pending = 0
        if not any(ch.launched for ch in self.chunks):
            self.open = False
            if self.stream is not None and self.reserved_sms > 0:
                F.set_sm_limit(0)
        return True
