"""FusedOptimizer — adopts a stock ``torch.optim`` optimizer and runs its update as ONE kernel per
parameter group on flat fp32 buckets (K6 in SURVEY §2.9), keeping the torch object as the source of
truth for hyper-parameters so LR schedulers, ``state_dict`` and checkpoints stay interchangeable
with plain PyTorch.

What changes underneath (CUDA only):

* every parameter's ``.data`` becomes a view into a flat fp32 *parameter bucket*, its ``.grad`` a view
  into a flat fp32 *gradient bucket* (the target of the wgrad kernels' atomic accumulation and of the
  fused all-reduce), and ``._det_shadow`` a view into a bf16 *shadow bucket* that the tensor-core
  kernels read and the update kernel refreshes;
* ``step()`` = (optional fused grad-norm + clip coefficient) → fused update kernel, which also
  applies the gradient scale (1/loss_scale · 1/aggregation_frequency · clip), skips on overflow and
  zeroes the gradient bucket in the same pass;
* per-step hyper-parameters travel through a pinned host buffer → device copy, so a captured CUDA
  graph replays with fresh learning rates and bias corrections.

Reference call sites replaced: ``optimizer.step()`` / ``zero_grad()`` / ``_average_gradients`` /
``clip_grads`` in ``harness/determined/pytorch/_pytorch_context.py:470-545``.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Tuple

import torch

from ..ops import functional as F

_ALIGN = 64  # elements; keeps every parameter 256-byte (fp32) / 128-byte (bf16) aligned

# Callables notified with the optimizer after every fused ``step()`` — the CUDA-graph runner uses this to learn
# which optimizers a captured batch stepped, so it can refresh their host-side hyper-parameters on replay.
STEP_OBSERVERS: List[Callable[["FusedOptimizer"], None]] = []


def _variant_of(opt: torch.optim.Optimizer) -> Optional[int]:
    name = type(opt).__name__
    groups = opt.param_groups
    if name == "Adam" and not any(g.get("amsgrad") or g.get("maximize") for g in groups):
        return F.OPT_ADAM
    if name == "AdamW" and not any(g.get("amsgrad") or g.get("maximize") for g in groups):
        return F.OPT_ADAMW
    if name == "SGD" and not any(g.get("maximize") for g in groups):
        return F.OPT_SGD
    if name == "RMSprop" and not any(g.get("centered") or g.get("maximize") for g in groups):
        return F.OPT_RMSPROP
    if name == "Adadelta" and not any(g.get("maximize") for g in groups):
        return F.OPT_ADADELTA
    return None


class _Group:
    """Flat buckets of one param group."""

    def __init__(self, params: List[torch.Tensor], alloc: Callable[[int, torch.dtype, str], torch.Tensor]):
        self.params = params
        self.offsets: List[int] = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() + _ALIGN - 1) // _ALIGN * _ALIGN
        self.total = max(off, _ALIGN)
        self.flat_p = alloc(self.total, torch.float32, "param")
        self.flat_g = alloc(self.total, torch.float32, "grad")
        self.flat_s1 = alloc(self.total, torch.float32, "state")
        self.flat_s2 = alloc(self.total, torch.float32, "state")
        self.flat_shadow = alloc(self.total, torch.bfloat16, "shadow")
        with torch.no_grad():
            for p, o in zip(params, self.offsets):
                n = p.numel()
                self.flat_p[o:o + n].copy_(p.detach().reshape(-1).float())
                old_grad = p.grad
                p.data = self.flat_p[o:o + n].view(p.shape)
                gview = self.flat_g[o:o + n].view(p.shape)
                if old_grad is not None:
                    gview.copy_(old_grad.float())
                p.grad = gview
                p._det_shadow = self.flat_shadow[o:o + n].view(p.shape)
            F.cast_to_bf16(self.flat_p, self.flat_shadow)
        self.hyper_host = torch.zeros(12, dtype=torch.float32).pin_memory()
        self.hyper_dev = torch.zeros(12, dtype=torch.float32, device=self.flat_p.device)
        self.step_count = 0

    def view(self, flat: torch.Tensor, i: int) -> torch.Tensor:
        o, p = self.offsets[i], self.params[i]
        return flat[o:o + p.numel()].view(p.shape)


class FusedOptimizer(torch.optim.Optimizer):
    def __init__(self, inner: torch.optim.Optimizer,
                 alloc: Optional[Callable[[int, torch.dtype, str], torch.Tensor]] = None) -> None:
        self.inner = inner
        super().__init__(inner.param_groups, inner.defaults)
        self.variant = _variant_of(inner)
        params = [p for g in inner.param_groups for p in g["params"]]
        on_cuda = len(params) > 0 and all(p.is_cuda and p.dtype == torch.float32 for p in params)
        self.fused = bool(self.variant is not None and on_cuda and F.native_available())
        self.groups: List[_Group] = []
        self.grad_scale_dev: Optional[torch.Tensor] = None  # device scalar multiplied into grads at update
        self.found_inf_dev: Optional[torch.Tensor] = None
        self._scratch: Optional[torch.Tensor] = None
        self.last_grad_norm_dev: Optional[torch.Tensor] = None
        # gradient bucket address -> (slice bounds per owner rank, callable run before a slice is updated); set by the trial
        # context for the duration of one step when that bucket's all-gather is pipelined with the update
        self.slice_pipelines: Dict[int, Any] = {}
        if self.fused:
            dev = params[0].device
            if alloc is None:
                def alloc(n: int, dtype: torch.dtype, kind: str) -> torch.Tensor:
                    return torch.zeros(n, dtype=dtype, device=dev)
            for g in inner.param_groups:
                self.groups.append(_Group(list(g["params"]), alloc))
            self.grad_scale_dev = torch.ones(1, dtype=torch.float32, device=dev)
            self.found_inf_dev = torch.zeros(1, dtype=torch.int32, device=dev)
            self._scratch = torch.zeros(4, dtype=torch.float32, device=dev)  # [sumsq, -, scale_out, norm]
            if inner.state:
                self._import_state(inner.state_dict())

    # ---- gradient plumbing ---------------------------------------------------------------------
    def grad_buckets(self) -> List[torch.Tensor]:
        return [g.flat_g for g in self.groups]

    def param_buckets(self) -> List[torch.Tensor]:
        return [g.flat_p for g in self.groups]

    def state_buckets(self) -> List[torch.Tensor]:
        out = []
        for g in self.groups:
            out += [g.flat_s1, g.flat_s2]
        return out

    def refresh_shadows(self) -> None:
        """Re-cast bf16 shadows after parameters were written from outside (load / broadcast)."""
        for g in self.groups:
            F.cast_to_bf16(g.flat_p, g.flat_shadow)

    def zero_grad(self, set_to_none: bool = False) -> None:  # noqa: D401 - torch signature
        if not self.fused:
            return self.inner.zero_grad(set_to_none=set_to_none)
        for g in self.groups:
            g.flat_g.zero_()

    # ---- step ------------------------------------------------------------------------------------
    def _fill_hyper(self, grp: _Group, pg: Dict[str, Any]) -> None:
        h = grp.hyper_host
        h.zero_()
        t = grp.step_count
        v = self.variant
        h[0] = float(pg["lr"])
        h[4] = float(pg.get("weight_decay", 0.0))
        if v in (F.OPT_ADAM, F.OPT_ADAMW):
            b1, b2 = pg["betas"]
            h[1], h[2], h[3] = float(b1), float(b2), float(pg["eps"])
            h[5], h[6] = 1.0 - float(b1) ** t, 1.0 - float(b2) ** t
        elif v == F.OPT_SGD:
            h[1] = float(pg.get("momentum", 0.0))
            h[2] = 1.0 if pg.get("nesterov") else 0.0
            h[7] = float(pg.get("dampening", 0.0))
            h[8] = 1.0 if t == 1 else 0.0
        elif v == F.OPT_RMSPROP:
            h[1], h[2], h[3] = float(pg["alpha"]), float(pg.get("momentum", 0.0)), float(pg["eps"])
        elif v == F.OPT_ADADELTA:
            h[1], h[3] = float(pg["rho"]), float(pg["eps"])

    def host_prepare(self) -> None:
        """Host half of a step: advance step counters and stage hyper-parameters in pinned memory.
        Safe to call outside a CUDA graph right before replaying a captured ``device_step``."""
        for grp, pg in zip(self.groups, self.inner.param_groups):
            grp.step_count += 1
            self._fill_hyper(grp, pg)

    def device_step(self, max_grad_norm: float = 0.0, zero_grad: bool = True, sumsq_ready: bool = False) -> None:
        """Device half of a step (capturable): hyper H2D, optional fused clip, fused update. ``sumsq_ready``: the
        gradient all-reduce already left the bucket's sum of squares in ``_scratch[0]`` (and raised ``found_inf_dev``),
        so the bucket is not read again."""
        assert self.fused
        scale = self.grad_scale_dev
        if max_grad_norm and max_grad_norm > 0:
            sc = self._scratch
            if not sumsq_ready:
                sc[0:1].zero_()
                for grp in self.groups:
                    F.grad_sumsq_finite(grp.flat_g, scale, sc[0:1], self.found_inf_dev)
            F.clip_coef(sc[0:1], max_grad_norm, scale, sc[2:4])
            scale = sc[2:3]
            self.last_grad_norm_dev = sc[3:4]
        for grp in self.groups:
            grp.hyper_dev.copy_(grp.hyper_host, non_blocking=True)
            pipe = self.slice_pipelines.get(grp.flat_g.data_ptr())
            if pipe is None:
                F.fused_optimizer_step(self.variant, grp.flat_p, grp.flat_g, grp.flat_s1,
                                       grp.flat_s2, grp.flat_shadow, grp.hyper_dev, scale, self.found_inf_dev, zero_grad)
                continue
            # the bucket's all-gather is still arriving, one owner's slice at a time (parallel/comm.py, pipelined
            # all-reduce): update each slice as soon as it is here, under the transfer of the next one
            bounds, before_chunk = pipe
            for owner, (lo, hi) in enumerate(bounds):
                before_chunk(owner)
                if hi > lo:
                    cut = lambda t: None if t is None else t[lo:hi]      # noqa: E731
                    F.fused_optimizer_step(self.variant, grp.flat_p[lo:hi], grp.flat_g[lo:hi], cut(grp.flat_s1), cut(grp.flat_s2),
                                           cut(grp.flat_shadow), grp.hyper_dev, scale, self.found_inf_dev, zero_grad)

    @torch.no_grad()
    def step(self, closure=None, max_grad_norm: float = 0.0, zero_grad: bool = False, sumsq_ready: bool = False):  # type: ignore[override]
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        if not self.fused:
            self.inner.step()
            return loss
        self.host_prepare()
        self.device_step(max_grad_norm=max_grad_norm, zero_grad=zero_grad, sumsq_ready=sumsq_ready)
        for observer in STEP_OBSERVERS:
            observer(self)
        return loss

    # ---- state dict in stock torch format -----------------------------------------------------------
    def _state_names(self) -> Tuple[Optional[str], Optional[str]]:
        v = self.variant
        if v in (F.OPT_ADAM, F.OPT_ADAMW):
            return "exp_avg", "exp_avg_sq"
        if v == F.OPT_SGD:
            return "momentum_buffer", None
        if v == F.OPT_RMSPROP:
            return "square_avg", "momentum_buffer"
        if v == F.OPT_ADADELTA:
            return "square_avg", "acc_delta"
        return None, None

    def state_dict(self) -> Dict[str, Any]:  # type: ignore[override]
        if not self.fused:
            return self.inner.state_dict()
        n1, n2 = self._state_names()
        state: Dict[int, Dict[str, Any]] = {}
        groups_out = []
        idx = 0
        for grp, pg in zip(self.groups, self.inner.param_groups):
            ids = []
            for i, _ in enumerate(grp.params):
                if grp.step_count > 0:
                    st: Dict[str, Any] = {"step": torch.tensor(float(grp.step_count))}
                    if n1:
                        st[n1] = grp.view(grp.flat_s1, i).detach().clone()
                    if n2:
                        st[n2] = grp.view(grp.flat_s2, i).detach().clone()
                    state[idx] = st
                ids.append(idx)
                idx += 1
            packed = {k: v for k, v in pg.items() if k != "params"}
            packed["params"] = ids
            groups_out.append(packed)
        return {"state": state, "param_groups": groups_out}

    def _import_state(self, sd: Dict[str, Any]) -> None:
        n1, n2 = self._state_names()
        idx = 0
        for grp in self.groups:
            steps = 0
            for i, _ in enumerate(grp.params):
                st = sd["state"].get(idx, sd["state"].get(str(idx)))
                if st:
                    if n1 and st.get(n1) is not None:
                        grp.view(grp.flat_s1, i).copy_(st[n1])
                        if self.variant == F.OPT_SGD:
                            # stock torch SGD state is the momentum buffer alone (no "step"): a restored buffer means the
                            # first update already happened, so the kernel must not re-seed the buffer with the gradient
                            steps = max(steps, 1)
                    if n2 and st.get(n2) is not None:
                        grp.view(grp.flat_s2, i).copy_(st[n2])
                    s = st.get("step", 0)
                    steps = max(steps, int(s.item() if torch.is_tensor(s) else s))
                idx += 1
            grp.step_count = steps

    @torch.no_grad()
    def load_state_dict(self, sd: Dict[str, Any]) -> None:  # type: ignore[override]
        if not self.fused:
            return self.inner.load_state_dict(sd)
        for pg, saved in zip(self.inner.param_groups, sd["param_groups"]):
            for k, v in saved.items():
                if k != "params":
                    pg[k] = v
        for grp in self.groups:
            grp.flat_s1.zero_()
            grp.flat_s2.zero_()
        self._import_state(sd)
