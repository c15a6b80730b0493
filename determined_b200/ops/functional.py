"""Tensor-level wrappers over the sm_100a kernel library.

Each function takes ``torch`` CUDA tensors, launches on the CURRENT stream (so it composes with
side streams and CUDA-graph capture), and returns/fills torch tensors. No function here
synchronises the device or touches host memory.

Operand conventions for :func:`gemm` mirror the C ABI: an operand is "K-major" when the contraction
dimension is contiguous, "MN-major" when the M/N dimension is. With those two flags forward
(``x @ W.T``), dgrad (``dy @ W``) and wgrad (``dy.T @ x``) need no transposed copies.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Tuple

import torch

from . import _native

EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_DGELU, EPI_BIAS_RESID = 0, 1, 2, 3, 4
OUT_BF16, OUT_F32, OUT_F32_ATOMIC_ADD = 0, 1, 2
OPT_ADAM, OPT_ADAMW, OPT_SGD, OPT_RMSPROP, OPT_ADADELTA = 0, 1, 2, 3, 4
RED_AVG, RED_SUM, RED_MAX, RED_MIN = 0, 1, 2, 3


def _p(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def native_available() -> bool:
    return torch.cuda.is_available() and _native.available()


def set_sm_limit(n: int) -> None:
    """Cap the SM count the persistent kernels size their grids from (0 = whole device); host-side, affects launches
    (and graph captures) made afterwards."""
    _native.check(_native.load().det_set_sm_limit(int(n)))


def device_sm_count(device=None) -> int:
    return torch.cuda.get_device_properties(device if device is not None else torch.cuda.current_device()).multi_processor_count


# ------------------------------------------------------------------------------------------------
# Per-step dropout seed (device resident so CUDA-graph replays draw fresh masks)
# ------------------------------------------------------------------------------------------------
_seed_state = {}


def step_seed_tensor(device=None) -> torch.Tensor:
    dev = torch.device(device if device is not None else torch.cuda.current_device())
    if dev.type == "cuda" and dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    t = _seed_state.get(dev)
    if t is None:
        t = torch.zeros(1, dtype=torch.int32, device=dev)
        _seed_state[dev] = t
    return t


def advance_step_seed(device=None) -> None:
    """Bump the device-side dropout seed (captured as one tiny kernel inside a step graph)."""
    step_seed_tensor(device).add_(1)


def set_step_seed(value: int, device=None) -> None:
    step_seed_tensor(device).fill_(int(value) & 0x7FFFFFFF)


# ------------------------------------------------------------------------------------------------
# GEMM
# ------------------------------------------------------------------------------------------------
def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, *, a_mn: bool = False,
         b_mn: bool = False, out_mode: int = OUT_BF16, epi: int = EPI_NONE, alpha: float = 1.0,
         bias: Optional[torch.Tensor] = None, aux_in: Optional[torch.Tensor] = None,
         aux_out: Optional[torch.Tensor] = None, dropout_p: float = 0.0, seed: int = 0,
         splitk: int = 0, force_bn: int = 0) -> torch.Tensor:
    """``out[M,N] = epilogue(a[M,K] @ b[N,K].T)`` on tcgen05 tensor cores (bf16 in, fp32 accumulate).

    ``a``: [M,K] (K-major) or, with ``a_mn``, stored [K,M].  ``b``: [N,K] or, with ``b_mn``, stored [K,N].
    """
    aux = aux_in if aux_in is not None else aux_out
    _native.call(
        "det_gemm_bf16", _p(a), _p(b), _p(out), M, N, K, int(a_mn), int(b_mn), a.stride(0), b.stride(0),
        out.stride(0), out_mode, epi, float(alpha), _p(bias), _p(aux_in), _p(aux_out),
        aux.stride(0) if aux is not None else 0, float(dropout_p), int(seed) & 0xFFFFFFFF,
        _p(step_seed_tensor(a.device)) if dropout_p > 0 else None, splitk, force_bn, _stream())
    return out


def tma_compatible(t: torch.Tensor) -> bool:
    return (t.dtype == torch.bfloat16 and t.dim() == 2 and t.stride(1) == 1 and t.stride(0) % 8 == 0
            and t.data_ptr() % 16 == 0)


def gemm_small(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int,
               sa: Tuple[int, int], sb: Tuple[int, int], sc: Tuple[int, int], *, bias: Optional[torch.Tensor] = None,
               act_tanh: bool = False, alpha: float = 1.0, accumulate: bool = False) -> torch.Tensor:
    """Strided CUDA-core GEMM for tiny heads: ``out[m,n] (+)= act(alpha * sum_k a[m,k] b[n,k] + bias[n])``.
    ``sa=(stride_m, stride_k)``, ``sb=(stride_n, stride_k)``, ``sc=(stride_m, stride_n)`` in elements."""
    def f32(t):
        if t.dtype == torch.float32:
            return 1
        if t.dtype == torch.bfloat16:
            return 0
        raise TypeError(f"gemm_small supports bf16/fp32, got {t.dtype}")
    _native.call("det_gemm_small", _p(a), _p(b), _p(out), M, N, K, sa[0], sa[1], sb[0], sb[1], sc[0], sc[1],
                 f32(a), f32(b), f32(out), _p(bias), int(act_tanh), float(alpha), int(accumulate), _stream())
    return out


# ------------------------------------------------------------------------------------------------
# Attention / LayerNorm / misc
# ------------------------------------------------------------------------------------------------
ATTENTION_SINGLE_TILE = 128   # sequences up to this length take the one-CTA-per-head kernel, longer ones the tiled one


def attention_fwd(qkv: torch.Tensor, mask: Optional[torch.Tensor], B: int, S: int, H: int, dropout_p: float,
                  seed: int) -> Tuple[torch.Tensor, torch.Tensor]:
    ctx = torch.empty(B * S, H * 64, dtype=torch.bfloat16, device=qkv.device)
    lse = torch.empty(B, H, S, dtype=torch.float32, device=qkv.device)
    name = "det_attention_fwd" if S <= ATTENTION_SINGLE_TILE else "det_attention_flash_fwd"
    _native.call(name, _p(qkv), _p(mask), _p(ctx), _p(lse), B, S, H, float(dropout_p),
                 int(seed) & 0xFFFFFFFF, _p(step_seed_tensor(qkv.device)) if dropout_p > 0 else None, _stream())
    return ctx, lse


def attention_bwd(qkv: torch.Tensor, dctx: torch.Tensor, mask: Optional[torch.Tensor], lse: torch.Tensor, B: int,
                  S: int, H: int, dropout_p: float, seed: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out`` is the forward output; the tiled kernel (S > 128) needs it for delta = rowsum(dO * O)."""
    dqkv = torch.empty_like(qkv)
    seed_dev = _p(step_seed_tensor(qkv.device)) if dropout_p > 0 else None
    if S <= ATTENTION_SINGLE_TILE:
        _native.call("det_attention_bwd", _p(qkv), _p(dctx), _p(mask), _p(lse), _p(dqkv), B, S, H, float(dropout_p),
                     int(seed) & 0xFFFFFFFF, seed_dev, _stream())
        return dqkv
    if out is None:
        raise ValueError("attention_bwd for sequences longer than 128 needs the forward output")
    dq_accum = torch.empty(B * S, H * 64, dtype=torch.float32, device=qkv.device)
    delta = torch.empty(B, H, S, dtype=torch.float32, device=qkv.device)
    _native.call("det_attention_flash_bwd", _p(qkv), _p(dctx), _p(out), _p(mask), _p(lse), _p(dqkv), _p(dq_accum), _p(delta),
                 B, S, H, float(dropout_p), int(seed) & 0xFFFFFFFF, seed_dev, _stream())
    return dqkv


def layernorm_fwd(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float, dropout_p: float = 0.0,
                  seed: int = 0) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    T, H = x.shape
    y = torch.empty_like(x)
    mean = torch.empty(T, dtype=torch.float32, device=x.device)
    rstd = torch.empty(T, dtype=torch.float32, device=x.device)
    _native.call("det_layernorm_fwd", _p(x), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), T, H, float(eps),
                 float(dropout_p), int(seed) & 0xFFFFFFFF,
                 _p(step_seed_tensor(x.device)) if dropout_p > 0 else None, _stream())
    return y, mean, rstd


def layernorm_bwd(dy: torch.Tensor, x: torch.Tensor, gamma: torch.Tensor, mean: torch.Tensor, rstd: torch.Tensor,
                  dgamma: Optional[torch.Tensor], dbeta: Optional[torch.Tensor], *, want_masked: bool = False,
                  in_dropout_p: float = 0.0, in_seed: int = 0, out_dropout_p: float = 0.0,
                  out_seed: int = 0, dxsum: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Returns (dx, dx_masked). dgamma/dbeta (fp32) are ACCUMULATED into; so is ``dxsum`` (fp32 [H]) with the column sums of
    dx_masked (of dx when no masked copy is produced) — the bias gradient of the GEMM that produced the LayerNorm input."""
    T, H = x.shape
    dx = torch.empty_like(x)
    dxm = torch.empty_like(x) if want_masked else None
    any_drop = in_dropout_p > 0 or out_dropout_p > 0
    _native.call("det_layernorm_bwd", _p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dx), _p(dxm), _p(dgamma),
                 _p(dbeta), _p(dxsum), T, H, float(in_dropout_p), int(in_seed) & 0xFFFFFFFF, float(out_dropout_p),
                 int(out_seed) & 0xFFFFFFFF, _p(step_seed_tensor(x.device)) if any_drop else None, _stream())
    return dx, dxm


def colsum_accum(x: torch.Tensor, out_f32: torch.Tensor) -> None:
    """out[n] += sum_t x[t, n] (bf16 in, fp32 accumulate) — bias gradients."""
    T, N = x.shape
    _native.call("det_colsum_bf16", _p(x), _p(out_f32), T, N, x.stride(0), _stream())


def embed_gather(ids: torch.Tensor, type_ids: Optional[torch.Tensor], word: torch.Tensor, pos: torch.Tensor,
                 typ: torch.Tensor, S: int) -> torch.Tensor:
    T = ids.numel()
    H = word.shape[1]
    out = torch.empty(T, H, dtype=torch.bfloat16, device=word.device)
    _native.call("det_embed_gather", _p(ids), _p(type_ids), _p(word), _p(pos), _p(typ), _p(out), T, S, H, _stream())
    return out


def embed_scatter(ids: torch.Tensor, type_ids: Optional[torch.Tensor], de: torch.Tensor, dword: torch.Tensor,
                  dpos: torch.Tensor, dtyp: torch.Tensor, S: int) -> None:
    T, H = de.shape
    if T % S == 0 and (S * H) % 8 == 0 and de.is_contiguous():
        # every position occurs once per sequence: its gradient is a sum over the batch — a column sum of de viewed as
        # [B, S*H] — instead of B-way contended atomics on the same S*H addresses
        _native.call("det_embed_scatter", _p(ids), _p(type_ids), _p(de), _p(dword), None, None, T, S, H, _stream())
        _native.call("det_colsum_bf16", _p(de), _p(dpos), T // S, S * H, S * H, _stream())
    else:
        _native.call("det_embed_scatter", _p(ids), _p(type_ids), _p(de), _p(dword), _p(dpos), None, T, S, H, _stream())
    # token-type rows are few: reduce instead of hammering two addresses with atomics
    if type_ids is None:
        _native.call("det_colsum_bf16", _p(de), _p(dtyp[0]), T, H, de.stride(0), _stream())
    else:
        for t in range(dtyp.shape[0]):
            _native.call("det_colsum_select_bf16", _p(de), _p(type_ids), t, _p(dtyp[t]), T, H, de.stride(0), _stream())


def softmax_ce(logits_f32: torch.Tensor, labels: torch.Tensor, gscale: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-row loss and dlogits = (softmax - onehot) * gscale in one pass."""
    B, C = logits_f32.shape
    loss = torch.empty(B, dtype=torch.float32, device=logits_f32.device)
    dlogits = torch.empty_like(logits_f32)
    _native.call("det_softmax_ce", _p(logits_f32), _p(labels), _p(loss), _p(dlogits), B, C, float(gscale), _stream())
    return loss, dlogits


def cast_to_bf16(src_f32: torch.Tensor, dst_bf16: torch.Tensor) -> None:
    _native.call("det_cast_f32_bf16", _p(src_f32), _p(dst_bf16), src_f32.numel(), _stream())


# ------------------------------------------------------------------------------------------------
# Optimizer / gradient bucket kernels
# ------------------------------------------------------------------------------------------------
def fused_optimizer_step(variant: int, param: torch.Tensor, grad: torch.Tensor, state1: torch.Tensor,
                         state2: Optional[torch.Tensor], shadow: Optional[torch.Tensor], hyper_dev: torch.Tensor,
                         grad_scale_dev: Optional[torch.Tensor], found_inf_dev: Optional[torch.Tensor],
                         zero_grad: bool) -> None:
    _native.call("det_fused_optimizer_step", variant, _p(param), _p(grad), _p(state1), _p(state2), _p(shadow),
                 param.numel(), _p(hyper_dev), _p(grad_scale_dev), _p(found_inf_dev), int(zero_grad), _stream())


def grad_sumsq_finite(grad: torch.Tensor, grad_scale_dev: Optional[torch.Tensor], sumsq_dev: torch.Tensor,
                      found_inf_dev: Optional[torch.Tensor]) -> None:
    _native.call("det_grad_sumsq_finite", _p(grad), grad.numel(), _p(grad_scale_dev), _p(sumsq_dev),
                 _p(found_inf_dev), _stream())


def clip_coef(sumsq_dev: torch.Tensor, max_norm: float, base_scale_dev: Optional[torch.Tensor],
              out2_dev: torch.Tensor) -> None:
    _native.call("det_clip_coef", _p(sumsq_dev), float(max_norm), _p(base_scale_dev), _p(out2_dev), _stream())


def loss_scale_update(state3_dev: torch.Tensor, found_inf_dev: torch.Tensor, growth: float, backoff: float,
                      growth_interval: int, divisor: float) -> None:
    _native.call("det_loss_scale_update", _p(state3_dev), _p(found_inf_dev), float(growth), float(backoff),
                 int(growth_interval), float(divisor), _stream())


# ------------------------------------------------------------------------------------------------
# Convolutional stack (NHWC bf16 activations; csrc/kernels/conv_sm100.cu + conv_ops.cu)
# ------------------------------------------------------------------------------------------------
CONV_FPROP, CONV_DGRAD, CONV_WGRAD = 0, 1, 2
ACT_NONE, ACT_RELU, ACT_LEAKY, ACT_TANH, ACT_SIGMOID = 0, 1, 2, 3, 4


def conv_out_size(size: int, k: int, stride: int, pad: int) -> int:
    return (size + 2 * pad - k) // stride + 1


def conv2d_implicit(mode: int, src: torch.Tensor, other: torch.Tensor, out: torch.Tensor, N: int, H: int, W: int,
                    Cin: int, Cout: int, R: int, S: int, pad_h: int, pad_w: int, *, bias: Optional[torch.Tensor] = None,
                    relu: bool = False, splitk: int = 0, force_bn: int = 0) -> torch.Tensor:
    """Unit-stride implicit-GEMM convolution on tcgen05 (no im2col buffer). ``mode``: CONV_FPROP (src=x,
    other=weight [Cout,R,S,Cin] → y), CONV_DGRAD (src=dy, other=weight → dx), CONV_WGRAD (src=dy, other=x →
    fp32 dW accumulated)."""
    _native.call("det_conv2d_implicit", mode, _p(src), _p(other), _p(out), N, H, W, Cin, Cout, R, S, pad_h, pad_w,
                 _p(bias), int(relu), splitk, force_bn, _stream())
    return out


def im2col(x: torch.Tensor, N: int, H: int, W: int, C: int, R: int, S: int, stride: int, pad_h: int, pad_w: int,
           Kp: int) -> torch.Tensor:
    Ho, Wo = conv_out_size(H, R, stride, pad_h), conv_out_size(W, S, stride, pad_w)
    col = torch.empty(N * Ho * Wo, Kp, dtype=torch.bfloat16, device=x.device)
    _native.call("det_im2col", _p(x), _p(col), N, H, W, C, R, S, stride, pad_h, pad_w, Kp, _stream())
    return col


def col2im(col: torch.Tensor, N: int, H: int, W: int, C: int, R: int, S: int, stride: int, pad_h: int, pad_w: int,
           Kp: int) -> torch.Tensor:
    dx = torch.empty(N, H, W, C, dtype=torch.bfloat16, device=col.device)
    _native.call("det_col2im", _p(col), _p(dx), N, H, W, C, R, S, stride, pad_h, pad_w, Kp, _stream())
    return dx


def bn_stats(x: torch.Tensor, C: int, running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor],
             eps: float, momentum: float) -> Tuple[torch.Tensor, torch.Tensor]:
    M = x.numel() // C
    scratch = torch.empty(2 * C, dtype=torch.float32, device=x.device)
    mean = torch.empty(C, dtype=torch.float32, device=x.device)
    rstd = torch.empty(C, dtype=torch.float32, device=x.device)
    _native.call("det_bn_stats", _p(x), _p(scratch), _p(mean), _p(rstd), _p(running_mean), _p(running_var), M, C,
                 float(eps), float(momentum), _stream())
    return mean, rstd


def bn_apply(x: torch.Tensor, resid: Optional[torch.Tensor], mean: torch.Tensor, rstd: torch.Tensor,
             gamma: Optional[torch.Tensor], beta: Optional[torch.Tensor], C: int, act: int = ACT_NONE,
             slope: float = 0.0) -> torch.Tensor:
    y = torch.empty_like(x)
    _native.call("det_bn_apply", _p(x), _p(resid), _p(y), _p(mean), _p(rstd), _p(gamma), _p(beta), x.numel() // C, C,
                 act, float(slope), _stream())
    return y


def bn_bwd(dy: torch.Tensor, x: torch.Tensor, a: Optional[torch.Tensor], mean: torch.Tensor, rstd: torch.Tensor,
           gamma: Optional[torch.Tensor], dgamma: Optional[torch.Tensor], dbeta: Optional[torch.Tensor], C: int,
           act: int = ACT_NONE, slope: float = 0.0, want_dresid: bool = False):
    dx = torch.empty_like(x)
    dresid = torch.empty_like(x) if want_dresid else None
    scratch = torch.empty(2 * C, dtype=torch.float32, device=x.device)
    _native.call("det_bn_bwd", _p(dy), _p(x), _p(a), _p(mean), _p(rstd), _p(gamma), _p(scratch), _p(dx), _p(dresid),
                 _p(dgamma), _p(dbeta), x.numel() // C, C, act, float(slope), _stream())
    return dx, dresid


def dwconv_out_size(n: int, k: int, stride: int, pad: int, dil: int) -> int:
    return (n + 2 * pad - dil * (k - 1) - 1) // stride + 1


def dwconv2d_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], R: int, S: int, stride: int, pad: int, dil: int,
                 relu: bool = False) -> torch.Tensor:
    """Depthwise convolution: x [N,H,W,C] bf16, w fp32 [R,S,C] → y [N,Ho,Wo,C] bf16."""
    N, H, W, C = x.shape
    y = torch.empty(N, dwconv_out_size(H, R, stride, pad, dil), dwconv_out_size(W, S, stride, pad, dil), C, dtype=torch.bfloat16, device=x.device)
    _native.call("det_dwconv2d_fwd", _p(x), _p(w), _p(bias), _p(y), N, H, W, C, R, S, stride, pad, dil, int(relu), _stream())
    return y


def dwconv2d_dgrad(dy: torch.Tensor, w: torch.Tensor, in_shape, R: int, S: int, stride: int, pad: int, dil: int) -> torch.Tensor:
    N, H, W, C = in_shape
    dx = torch.empty(N, H, W, C, dtype=torch.bfloat16, device=dy.device)
    _native.call("det_dwconv2d_dgrad", _p(dy), _p(w), _p(dx), N, H, W, C, R, S, stride, pad, dil, _stream())
    return dx


def dwconv2d_wgrad(x: torch.Tensor, dy: torch.Tensor, dw: torch.Tensor, R: int, S: int, stride: int, pad: int, dil: int) -> None:
    """dw (fp32 [R,S,C]) += the weight gradient."""
    N, H, W, C = x.shape
    _native.call("det_dwconv2d_wgrad", _p(x), _p(dy), _p(dw), N, H, W, C, R, S, stride, pad, dil, _stream())


def avgpool2d_fwd(x: torch.Tensor, k: int, stride: int, pad: int, count_include_pad: bool) -> torch.Tensor:
    N, H, W, C = x.shape
    y = torch.empty(N, conv_out_size(H, k, stride, pad), conv_out_size(W, k, stride, pad), C, dtype=torch.bfloat16, device=x.device)
    _native.call("det_avgpool2d_fwd", _p(x), _p(y), N, H, W, C, k, stride, pad, int(count_include_pad), _stream())
    return y


def avgpool2d_bwd(dy: torch.Tensor, in_shape, k: int, stride: int, pad: int, count_include_pad: bool) -> torch.Tensor:
    N, H, W, C = in_shape
    dx = torch.empty(N, H, W, C, dtype=torch.bfloat16, device=dy.device)
    _native.call("det_avgpool2d_bwd", _p(dy), _p(dx), N, H, W, C, k, stride, pad, int(count_include_pad), _stream())
    return dx


def maxpool_fwd(x: torch.Tensor, k: int, stride: int, pad: int) -> Tuple[torch.Tensor, torch.Tensor]:
    N, H, W, C = x.shape
    Ho, Wo = conv_out_size(H, k, stride, pad), conv_out_size(W, k, stride, pad)
    y = torch.empty(N, Ho, Wo, C, dtype=torch.bfloat16, device=x.device)
    idx = torch.empty(N, Ho, Wo, C, dtype=torch.uint8, device=x.device)
    _native.call("det_maxpool_fwd", _p(x), _p(y), _p(idx), N, H, W, C, k, stride, pad, _stream())
    return y, idx


def maxpool_bwd(dy: torch.Tensor, idx: torch.Tensor, in_shape, k: int, stride: int, pad: int) -> torch.Tensor:
    N, H, W, C = in_shape
    dx = torch.empty(N, H, W, C, dtype=torch.bfloat16, device=dy.device)
    _native.call("det_maxpool_bwd", _p(dy), _p(idx), _p(dx), N, H, W, C, k, stride, pad, _stream())
    return dx


def avgpool_fwd(x: torch.Tensor) -> torch.Tensor:
    N, H, W, C = x.shape
    y = torch.empty(N, C, dtype=torch.bfloat16, device=x.device)
    _native.call("det_avgpool_fwd", _p(x), _p(y), N, H * W, C, _stream())
    return y


def avgpool_bwd(dy: torch.Tensor, in_shape) -> torch.Tensor:
    N, H, W, C = in_shape
    dx = torch.empty(N, H, W, C, dtype=torch.bfloat16, device=dy.device)
    _native.call("det_avgpool_bwd", _p(dy), _p(dx), N, H * W, C, _stream())
    return dx


def act_fwd(x: torch.Tensor, act: int, slope: float = 0.0) -> torch.Tensor:
    y = torch.empty_like(x)
    _native.call("det_act_fwd", _p(x), _p(y), x.numel(), act, float(slope), _stream())
    return y


def act_bwd(dy: torch.Tensor, y: torch.Tensor, act: int, slope: float = 0.0) -> torch.Tensor:
    dx = torch.empty_like(y)
    _native.call("det_act_bwd", _p(dy), _p(y), _p(dx), y.numel(), act, float(slope), _stream())
    return dx


def dropout(x: torch.Tensor, p: float, seed: int) -> torch.Tensor:
    """Stateless-hash dropout keyed by (seed, device step seed, element index); calling it again with the same seed
    on the gradient IS the backward pass."""
    y = torch.empty_like(x)
    _native.call("det_dropout", _p(x), _p(y), x.numel(), float(p), int(seed) & 0xFFFFFFFF,
                 _p(step_seed_tensor(x.device)), _stream())
    return y


def nchw_to_nhwc_bf16(x: torch.Tensor) -> torch.Tensor:
    """fp32 NCHW image batch → bf16 NHWC (the layout every convolutional kernel here works in)."""
    N, C, H, W = x.shape
    out = torch.empty(N, H, W, C, dtype=torch.bfloat16, device=x.device)
    _native.call("det_nchw_to_nhwc_bf16", _p(x), _p(out), N, C, H * W, _stream())
    return out


def u8_nhwc_to_bf16(x: torch.Tensor, mean, std) -> torch.Tensor:
    """uint8 NHWC image bytes → normalised bf16 NHWC: ``(x / 255 - mean[c]) / std[c]`` in one pass."""
    C = x.shape[-1]
    out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
    m = (ctypes.c_float * C)(*[float(v) for v in mean])
    s = (ctypes.c_float * C)(*[float(v) for v in std])
    _native.call("det_u8_to_bf16_norm", _p(x), _p(out), x.numel(), C, m, s, _stream())
    return out


def gemm_stats(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, colstats: torch.Tensor, *,
               bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out = a @ b.T (+ bias)`` and ``colstats[0:N] += column sums``, ``colstats[N:2N] += column sums of squares`` of the
    bf16 output, from the GEMM epilogue (``colstats`` fp32, cleared by the caller)."""
    _native.call("det_gemm_bf16_stats", _p(a), _p(b), _p(out), M, N, K, a.stride(0), b.stride(0), out.stride(0),
                 EPI_BIAS if bias is not None else EPI_NONE, _p(bias), _p(colstats), _stream())
    return out


def conv2d_fprop_stats(x: torch.Tensor, w: torch.Tensor, y: torch.Tensor, N: int, H: int, W: int, Cin: int, Cout: int, R: int,
                       S: int, pad_h: int, pad_w: int, colstats: torch.Tensor, *, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Implicit-GEMM forward convolution that also accumulates per-channel sum / sum of squares of its output."""
    _native.call("det_conv2d_fprop_stats", _p(x), _p(w), _p(y), N, H, W, Cin, Cout, R, S, pad_h, pad_w, _p(bias), _p(colstats),
                 _stream())
    return y


def bn_finalize(sums: torch.Tensor, M: int, C: int, running_mean: Optional[torch.Tensor], running_var: Optional[torch.Tensor],
                eps: float, momentum: float) -> Tuple[torch.Tensor, torch.Tensor]:
    mean = torch.empty(C, dtype=torch.float32, device=sums.device)
    rstd = torch.empty(C, dtype=torch.float32, device=sums.device)
    _native.call("det_bn_finalize", _p(sums), _p(mean), _p(rstd), _p(running_mean), _p(running_var), M, C, float(eps), float(momentum),
                 _stream())
    return mean, rstd


# ------------------------------------------------------------------------------------------------
# Recurrent-cell gate (csrc/kernels/rnn_ops.cu)
# ------------------------------------------------------------------------------------------------
GATE_ACTS = {"tanh": 0, "relu": 1, "sigmoid": 2, "identity": 3}


def _ptr_array(tensors):
    arr = (ctypes.c_void_p * len(tensors))(*[None if t is None else t.data_ptr() for t in tensors])
    return arr, ctypes.cast(arr, ctypes.c_void_p)


def highway_gate_fwd(ch: torch.Tensor, sp: torch.Tensor, mask: Optional[torch.Tensor], acts: Tuple[int, ...]):
    """``s_j = sp + sigmoid(c_j) * (f_j(h_j) - sp)`` for the ``k = len(acts)`` nodes whose ``[c_j | h_j]`` blocks are the
    columns of ``ch`` [B, k*2*nh]; returns ``(states, masked)`` — ``masked[j] = states[j] * mask`` when a mask is given."""
    k = len(acts)
    B, nh = sp.shape
    assert ch.shape == (B, 2 * nh * k) and ch.is_contiguous() and sp.is_contiguous() and ch.dtype == sp.dtype == torch.bfloat16
    s = [torch.empty_like(sp) for _ in range(k)]
    sm = [torch.empty_like(sp) for _ in range(k)] if mask is not None else []
    s_keep, s_arr = _ptr_array(s)
    m_keep, m_arr = _ptr_array(sm) if sm else (None, None)
    acts_arr = (ctypes.c_int * k)(*acts)
    _native.call("det_highway_gate_fwd", _p(ch), _p(sp), _p(mask), s_arr, m_arr, ctypes.cast(acts_arr, ctypes.c_void_p), k, B, nh, _stream())
    del s_keep, m_keep
    return s, sm


def highway_gate_bwd(ch: torch.Tensor, sp: torch.Tensor, mask: Optional[torch.Tensor], ds, dsm, acts: Tuple[int, ...]):
    """Gradients of :func:`highway_gate_fwd`: ``ds[j]`` / ``dsm[j]`` (either may be ``None``) are the gradients of node j's
    state and of its masked copy; returns ``(d_ch [B, k*2*nh], d_sp [B, nh])``."""
    k = len(acts)
    B, nh = sp.shape
    dch, dsp = torch.empty_like(ch), torch.empty_like(sp)
    a_keep, a_arr = _ptr_array(ds)
    b_keep, b_arr = _ptr_array(dsm) if mask is not None else (None, None)
    acts_arr = (ctypes.c_int * k)(*acts)
    _native.call("det_highway_gate_bwd", _p(ch), _p(sp), _p(mask), a_arr, b_arr, _p(dch), _p(dsp),
                 ctypes.cast(acts_arr, ctypes.c_void_p), k, B, nh, _stream())
    del a_keep, b_keep
    return dch, dsp
