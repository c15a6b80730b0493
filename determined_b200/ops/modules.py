"""``nn.Module`` drop-ins and autograd functions built on the sm_100a kernels.

Conventions
-----------
* Parameters stay fp32 ("master" weights, what checkpoints store). Tensor-core kernels read a bf16
  *shadow* of each matrix: when the parameter was adopted by
  :class:`determined_b200.parallel.fused_optimizer.FusedOptimizer` the shadow is a view into the
  optimizer's bf16 bucket and is refreshed by the fused update kernel itself; otherwise it is cast on
  demand and cached per parameter version.
* Weight/bias gradients are ACCUMULATED by the kernels straight into ``param.grad`` (fp32; a view
  into the flat gradient bucket when the optimizer is fused): wgrad GEMMs use split-K
  ``red.global.add`` epilogues, bias/LayerNorm gradients use fp32 atomics. The autograd functions
  therefore return ``None`` for parameters — gradient accumulation over
  ``aggregation_frequency`` micro-batches costs no extra kernel.
* Activations are bf16 [tokens, features].
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn

from . import functional as F


def _use_native(t: torch.Tensor) -> bool:
    return t.is_cuda


def shadow_of(p: torch.Tensor) -> torch.Tensor:
    """bf16 copy of a matrix parameter for the tensor-core kernels."""
    sh = getattr(p, "_det_shadow", None)
    if sh is not None:
        return sh
    cache = getattr(p, "_det_shadow_cache", None)
    if cache is not None and cache[0] == p._version and cache[1].device == p.device:
        return cache[1]
    with torch.no_grad():
        sh = p.detach().to(torch.bfloat16).contiguous()
    try:
        p._det_shadow_cache = (p._version, sh)
    except Exception:  # pragma: no cover
        pass
    return sh


def grad_buffer(p: torch.Tensor) -> torch.Tensor:
    """fp32 gradient accumulator of ``p`` (allocated zeroed on first use)."""
    g = p.grad
    if g is None:
        g = torch.zeros_like(p, dtype=torch.float32, memory_format=torch.contiguous_format)
        p.grad = g
    return g


# ------------------------------------------------------------------------------------------------
# Gradient-readiness notifications (what Horovod gets from per-parameter autograd hooks)
# ------------------------------------------------------------------------------------------------
# The autograd nodes in this package write parameter gradients themselves and return ``None`` for them, so torch's
# post-accumulate hooks never fire. Instead every node reports (a) in forward, which trainable parameters it used, and
# (b) at the end of backward, that their gradients for this use are complete. ``parallel.overlap.GradOverlap`` turns
# that into "this slice of the flat gradient bucket is final → start its all-reduce now".
_PARAM_USES: dict = {}
GRAD_OBSERVERS: list = []


def note_param_use(ctx, *params: Optional[torch.Tensor]) -> None:
    """Forward side (called from ``autograd.Function.forward``): count one pending gradient contribution per trainable
    parameter. No-op without observers, and when this forward will never be differentiated (``torch.no_grad`` or no
    input requiring grad — grad mode itself is always off inside ``Function.forward``, hence the ctx test)."""
    if not GRAD_OBSERVERS or not any(ctx.needs_input_grad):
        return
    for p in params:
        if p is not None and p.requires_grad:
            _PARAM_USES[id(p)] = _PARAM_USES.get(id(p), 0) + 1


def pending_uses(p: torch.Tensor) -> int:
    return _PARAM_USES.get(id(p), 0)


def notify_grads_written(*params: Optional[torch.Tensor]) -> None:
    """Backward side: the contribution counted by :func:`note_param_use` has been written (or skipped because the
    parameter no longer requires grad)."""
    if not GRAD_OBSERVERS:
        return
    done = []
    for p in params:
        if p is None:
            continue
        n = _PARAM_USES.get(id(p), 0)
        if n > 0:
            _PARAM_USES[id(p)] = n - 1
            done.append(p)
    if done:
        for obs in GRAD_OBSERVERS:
            obs(done)


def reset_param_uses() -> None:
    _PARAM_USES.clear()


# ------------------------------------------------------------------------------------------------
# Parameter gradients on a side stream
# ------------------------------------------------------------------------------------------------
# In a layer's backward the data-gradient kernels form a chain (each needs the previous one's output) while the weight /
# bias gradients hang off that chain as leaves: nothing in backward reads them. Launched on a second stream they become
# independent work the GPU can start as soon as SMs drain from a chain kernel's tail, instead of every kernel paying its
# own ramp-up and tail in sequence. ``side_grads(device)`` opens such a window; inside it ``accumulate_wgrad`` /
# ``accumulate_bgrad`` run on the side stream after whatever the main stream has enqueued so far, and leaving the window
# makes the main stream wait for them (so every tensor they read is still referenced by the caller's frame). Works the
# same under CUDA-graph capture, where the two streams become parallel branches of the graph.
import contextlib
import os

_SIDE_STREAMS: dict = {}
_SIDE_ACTIVE: list = []
_SIDE_ENABLED = os.environ.get("DET_SIDE_GRADS", "1") == "1"


@contextlib.contextmanager
def side_grads(device: torch.device):
    if not _SIDE_ENABLED or device.type != "cuda":
        yield
        return
    side = _SIDE_STREAMS.get(device)
    if side is None:
        side = _SIDE_STREAMS[device] = torch.cuda.Stream(device=device)
    _SIDE_ACTIVE.append(side)
    try:
        yield
    finally:
        _SIDE_ACTIVE.pop()
        torch.cuda.current_stream(device).wait_stream(side)


def _on_side(fn) -> None:
    if not _SIDE_ACTIVE:
        fn()
        return
    side = _SIDE_ACTIVE[-1]
    side.wait_stream(torch.cuda.current_stream(side.device))
    with torch.cuda.stream(side):
        fn()


def accumulate_wgrad(weight: torch.Tensor, dy: torch.Tensor, x: torch.Tensor) -> None:
    """weight.grad[N_out, K_in] += dy[T, N_out]^T @ x[T, K_in] (both operands MN-major, split-K atomics)."""
    g = grad_buffer(weight)
    n_out, k_in = weight.shape
    _on_side(lambda: F.gemm(dy, x, g, n_out, k_in, dy.shape[0], a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD))


def accumulate_bgrad(bias: torch.Tensor, dy: torch.Tensor) -> None:
    g = grad_buffer(bias)
    _on_side(lambda: F.colsum_accum(dy, g))


# ------------------------------------------------------------------------------------------------
# Linear
# ------------------------------------------------------------------------------------------------
class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, gelu: bool):
        T, K = x.shape
        N = weight.shape[0]
        w = shadow_of(weight)
        out = torch.empty(T, N, dtype=torch.bfloat16, device=x.device)
        z = None
        if gelu:
            z = torch.empty_like(out) if (x.requires_grad or weight.requires_grad) else None
            F.gemm(x, w, out, T, N, K, epi=F.EPI_BIAS_GELU, bias=bias, aux_out=z)
        else:
            F.gemm(x, w, out, T, N, K, epi=F.EPI_BIAS if bias is not None else F.EPI_NONE, bias=bias)
        ctx.save_for_backward(x, z)
        ctx.weight, ctx.bias, ctx.gelu = weight, bias, gelu
        note_param_use(ctx, weight, bias)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, z = ctx.saved_tensors
        weight, bias = ctx.weight, ctx.bias
        dy = dy.contiguous()
        T, K = x.shape
        N = weight.shape[0]
        w = shadow_of(weight)
        if ctx.gelu:
            # dz = dy * gelu'(z): fold into a copy through the dgrad-free path (elementwise in torch
            # would re-read; use the GEMM epilogue only when an input gradient is needed)
            zf = z.float()
            z2 = zf * zf
            t = torch.tanh(zf * (0.7978845608028654 + 0.035677408136300125 * z2))      # tanh-form GELU, as the epilogue
            dyz = (dy.float() * (0.5 * (1 + t) + 0.5 * zf * (1 - t * t) * (0.7978845608028654 + 0.10703222440890037 * z2))).to(torch.bfloat16)
            dy = dyz
        if weight.requires_grad:
            accumulate_wgrad(weight, dy, x)
        if bias is not None and bias.requires_grad:
            accumulate_bgrad(bias, dy)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(T, K, dtype=torch.bfloat16, device=x.device)
            F.gemm(dy, w, dx, T, K, N, b_mn=True)
        notify_grads_written(weight, bias)
        return dx, None, None, None


def _tma_ok(rows: int, feats_in: int, feats_out: int) -> bool:
    return feats_in % 8 == 0 and feats_out % 8 == 0 and rows >= 1


class Linear(nn.Linear):
    """``nn.Linear`` whose CUDA path is the tcgen05 GEMM (bf16 activations, fp32 accumulate) with the
    bias (+ optional GELU) fused in the epilogue. On CPU — or for shapes TMA cannot address (feature
    counts not a multiple of 8) — it is a plain ``torch.nn.functional.linear``."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, activation: Optional[str] = None):
        super().__init__(in_features, out_features, bias=bias)
        if activation not in (None, "gelu"):
            raise ValueError("activation must be None or 'gelu'")
        self.activation = activation

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not (_use_native(x) and _tma_ok(x.numel() // x.shape[-1], self.in_features, self.out_features)):
            y = nn.functional.linear(x, self.weight.to(x.dtype), None if self.bias is None else self.bias.to(x.dtype))
            return nn.functional.gelu(y, approximate="tanh") if self.activation == "gelu" else y
        lead = x.shape[:-1]
        x2 = x.reshape(-1, self.in_features)
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        x2 = x2.contiguous()
        y = _LinearFn.apply(x2, self.weight, self.bias, self.activation == "gelu")
        return y.reshape(*lead, self.out_features)


# ------------------------------------------------------------------------------------------------
# LayerNorm
# ------------------------------------------------------------------------------------------------
class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        y, mean, rstd = F.layernorm_fwd(x, gamma, beta, eps)
        ctx.save_for_backward(x, mean, rstd)
        ctx.gamma, ctx.beta = gamma, beta
        note_param_use(ctx, gamma, beta)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, mean, rstd = ctx.saved_tensors
        dg = grad_buffer(ctx.gamma) if ctx.gamma.requires_grad else None
        db = grad_buffer(ctx.beta) if ctx.beta.requires_grad else None
        dx, _ = F.layernorm_bwd(dy.contiguous(), x, ctx.gamma, mean, rstd, dg, db)
        notify_grads_written(ctx.gamma, ctx.beta)
        return dx, None, None, None


class LayerNorm(nn.LayerNorm):
    """``nn.LayerNorm`` over the last dimension; CUDA path = one-warp-per-row kernels (register-resident
    rows for hidden sizes 256/512/768/1024, a strided variant for any other multiple of 8), gamma/beta gradients
    accumulated with fp32 atomics."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        H = x.shape[-1]
        if not (_use_native(x) and H % 8 == 0 and H <= 16384 and self.elementwise_affine):
            return super().forward(x)
        lead = x.shape[:-1]
        x2 = x.reshape(-1, H).to(torch.bfloat16).contiguous()
        return _LayerNormFn.apply(x2, self.weight, self.bias, self.eps).reshape(*lead, H)


# ------------------------------------------------------------------------------------------------
# Cross entropy
# ------------------------------------------------------------------------------------------------
class _SoftmaxCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, labels):
        lf = logits.float().contiguous()
        loss, dlogits = F.softmax_ce(lf, labels.contiguous(), 1.0 / lf.shape[0])
        ctx.save_for_backward(dlogits)
        ctx.in_dtype = logits.dtype
        return loss.mean()

    @staticmethod
    def backward(ctx, dloss):
        (dlogits,) = ctx.saved_tensors
        return (dlogits * dloss).to(ctx.in_dtype), None


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """Mean softmax cross-entropy; CUDA path computes loss and dlogits in one fused pass."""
    if _use_native(logits) and logits.dim() == 2:
        return _SoftmaxCEFn.apply(logits, labels)
    return nn.functional.cross_entropy(logits.float(), labels)


# ------------------------------------------------------------------------------------------------
# Self-attention (packed QKV)
# ------------------------------------------------------------------------------------------------
class _AttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, mask, B, S, H, p, seed):
        out, lse = F.attention_fwd(qkv, mask, B, S, H, p, seed)
        if S > F.ATTENTION_SINGLE_TILE:
            ctx.save_for_backward(qkv, lse, out)      # the tiled backward needs delta = rowsum(dO * O)
        else:
            ctx.save_for_backward(qkv, lse)
        ctx.mask, ctx.dims, ctx.p, ctx.seed = mask, (B, S, H), p, seed
        return out

    @staticmethod
    def backward(ctx, dctx):
        qkv, lse = ctx.saved_tensors[:2]
        out = ctx.saved_tensors[2] if len(ctx.saved_tensors) > 2 else None
        B, S, H = ctx.dims
        dqkv = F.attention_bwd(qkv, dctx.contiguous(), ctx.mask, lse, B, S, H, ctx.p, ctx.seed, out)
        return dqkv, None, None, None, None, None, None


def fused_self_attention(qkv: torch.Tensor, mask: Optional[torch.Tensor], B: int, S: int, H: int,
                         dropout_p: float = 0.0, seed: int = 0) -> torch.Tensor:
    """softmax(QK^T/8 + mask) V for packed ``qkv`` [B*S, 3*H*64] bf16 → [B*S, H*64] (one CTA per head up to
    128 tokens, the tiled online-softmax kernel beyond)."""
    return _AttentionFn.apply(qkv, mask, B, S, H, dropout_p, seed)


def reference_self_attention(qkv: torch.Tensor, mask: Optional[torch.Tensor], B: int, S: int, H: int) -> torch.Tensor:
    """Plain-PyTorch fp32 statement of the same math (CPU path and numerics oracle)."""
    hid = H * 64
    q, k, v = [t.float().view(B, S, H, 64).transpose(1, 2) for t in qkv.split(hid, dim=-1)]
    sc = q @ k.transpose(-1, -2) / math.sqrt(64.0)
    if mask is not None:
        sc = sc + mask[:, None, None, :].float()
    return (torch.softmax(sc, dim=-1) @ v).transpose(1, 2).reshape(B * S, hid).to(qkv.dtype)
