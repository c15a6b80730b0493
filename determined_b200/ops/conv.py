"""Convolutional ``nn.Module`` drop-ins on the sm_100a kernels (Conv2d, ConvTranspose2d, BatchNorm2d with fused
residual-add/activation, pooling, activations, dropout).

Tensor convention
-----------------
Modules take and return tensors with the usual LOGICAL shape ``[N, C, H, W]``. On CUDA the data behind them is bf16 in
channels-last order — ``x.permute(0, 2, 3, 1)`` is the contiguous NHWC array every kernel works on — so user model
code (``flatten``, ``view``, torch functions) keeps its stock-PyTorch meaning. On CPU each module is the plain
``torch.nn.functional`` call in fp32, which is also the numerics oracle in the tests.

Parameter layout
----------------
Convolution weights are held in the order the tensor cores contract over: ``Conv2d.weight`` is ``[Cout, R, S, Cin]``
and ``ConvTranspose2d.weight`` is ``[Cin, R, S, Cout]`` (fp32 master; bf16 shadow for the kernels, see
:mod:`determined_b200.ops.modules`). ``state_dict()`` / ``load_state_dict()`` translate to and from torch's
``[Cout, Cin, R, S]`` / ``[Cin, Cout, R, S]`` so checkpoints interchange with ``torch.nn`` models.

Kernel selection per convolution
--------------------------------
* 1x1, stride 1, no padding: the plain tcgen05 GEMM on the NHWC array viewed as ``[pixels, Cin]``.
* stride 1, channel counts multiples of 8: implicit GEMM (``conv_sm100.cu``) — TMA fetches shifted 4-D boxes, the
  padding halo is TMA's out-of-bounds zero fill, no im2col buffer exists.
* everything else (strided, RGB/greyscale first layers, transposed): gather (im2col) / gather-form col2im kernels
  around the plain GEMM.

The reference reaches all of this through ``torch.nn`` → cuDNN in user model code, e.g.
``examples/tutorials/mnist_pytorch/model_def.py:34-53``, ``examples/computer_vision/cifar10_pytorch/model_def.py:49-74``.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn as nn

from . import functional as F
from .modules import accumulate_bgrad, grad_buffer, note_param_use, notify_grads_written, shadow_of

_ACT_CODES = {None: F.ACT_NONE, "relu": F.ACT_RELU, "leaky_relu": F.ACT_LEAKY, "tanh": F.ACT_TANH, "sigmoid": F.ACT_SIGMOID}


def _pair(v) -> Tuple[int, int]:
    return (v, v) if isinstance(v, int) else (int(v[0]), int(v[1]))


def _round8(n: int) -> int:
    return (n + 7) // 8 * 8


def to_nhwc(x: torch.Tensor) -> torch.Tensor:
    """Logical NCHW tensor → contiguous bf16 NHWC array (a view when the data already is channels-last bf16)."""
    if x.dtype == torch.float32 and x.is_contiguous() and not x.requires_grad:
        return F.nchw_to_nhwc_bf16(x)
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    xp = x.permute(0, 2, 3, 1)
    return xp if xp.is_contiguous() else xp.contiguous()


def from_nhwc(y: torch.Tensor) -> torch.Tensor:
    """NHWC array → logical NCHW tensor (no copy)."""
    return y.permute(0, 3, 1, 2)


def _pad_cols(t: torch.Tensor, cols: int) -> torch.Tensor:
    """[rows, c] → [rows, cols] zero padded (small first/last layers only)."""
    if t.shape[1] == cols:
        return t
    out = torch.zeros(t.shape[0], cols, dtype=t.dtype, device=t.device)
    out[:, :t.shape[1]] = t
    return out


# ------------------------------------------------------------------------------------------------
# Conv2d
# ------------------------------------------------------------------------------------------------
class _Conv2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, pad, relu, want_stats=False, passthrough=False):
        # passthrough: also hand x back as an output. A residual block reads its input twice (first convolution + skip
        # connection); routing the skip through this node lets backward ADD the skip gradient in the data-gradient GEMM's
        # epilogue instead of leaving a separate elementwise add to autograd.
        N, H, W, Cin = x.shape
        Cout, R, S, _ = weight.shape
        ph, pw = pad
        Ho, Wo = F.conv_out_size(H, R, stride, ph), F.conv_out_size(W, S, stride, pw)
        M = N * Ho * Wo
        w = shadow_of(weight)
        aligned = Cin % 8 == 0 and Cout % 8 == 0
        # per-channel sum / sum of squares of the output, accumulated by the producing kernel's epilogue (BatchNorm
        # statistics without a pass over the output); only where the epilogue writes the final values
        sums = torch.zeros(2 * Cout, dtype=torch.float32, device=x.device) if (want_stats and aligned and not relu) else None
        if aligned and stride == 1 and R == 1 and S == 1 and ph == 0 and pw == 0:
            kind = "gemm"
            y = torch.empty(N, Ho, Wo, Cout, dtype=torch.bfloat16, device=x.device)
            if sums is not None:
                F.gemm_stats(x.view(M, Cin), w.view(Cout, Cin), y.view(M, Cout), M, Cout, Cin, sums, bias=bias)
            else:
                F.gemm(x.view(M, Cin), w.view(Cout, Cin), y.view(M, Cout), M, Cout, Cin,
                       epi=F.EPI_BIAS if bias is not None else F.EPI_NONE, bias=bias)
            if relu:
                y = F.act_fwd(y, F.ACT_RELU)
        elif aligned and stride == 1:
            kind = "implicit"
            y = torch.empty(N, Ho, Wo, Cout, dtype=torch.bfloat16, device=x.device)
            if sums is not None:
                F.conv2d_fprop_stats(x, w, y, N, H, W, Cin, Cout, R, S, ph, pw, sums, bias=bias)
            else:
                F.conv2d_implicit(F.CONV_FPROP, x, w, y, N, H, W, Cin, Cout, R, S, ph, pw, bias=bias, relu=relu)
        else:
            kind = "explicit"
            K = R * S * Cin
            Kp, Cp = _round8(K), _round8(Cout)
            col = F.im2col(x, N, H, W, Cin, R, S, stride, ph, pw, Kp)
            wk = _pad_cols(w.view(Cout, K), Kp)
            yp = torch.empty(M, Cp, dtype=torch.bfloat16, device=x.device)
            if sums is not None:     # aligned ⇒ Cp == Cout
                F.gemm_stats(col, wk, yp, M, Cout, Kp, sums, bias=bias)
            else:
                F.gemm(col, wk, yp, M, Cout, Kp, epi=F.EPI_BIAS if bias is not None else F.EPI_NONE, bias=bias)
            y = (yp if Cp == Cout else yp[:, :Cout].contiguous()).view(N, Ho, Wo, Cout)
            if relu:
                y = F.act_fwd(y, F.ACT_RELU)
            # first layer of a network (no input gradient wanted): its few-channel gather is the expensive kind, so the
            # column matrix is kept for the weight gradient instead of being rebuilt in backward
            ctx.col = col if not ctx.needs_input_grad[0] else None
        ctx.save_for_backward(x, y if relu else None)
        ctx.weight, ctx.bias = weight, bias
        ctx.cfg = (kind, stride, ph, pw, relu)
        ctx.outs = (bool(want_stats), bool(passthrough))
        note_param_use(ctx, weight, bias)
        outs = [y]
        if want_stats:
            if sums is None:   # this path's epilogue cannot produce them: signal the caller to run the reduction kernel
                sums = torch.empty(0, dtype=torch.float32, device=x.device)
            ctx.mark_non_differentiable(sums)
            outs.append(sums)
        if passthrough:
            outs.append(x.view_as(x))
        return outs[0] if len(outs) == 1 else tuple(outs)

    @staticmethod
    def backward(ctx, dy, *rest):
        want_stats, passthrough = ctx.outs
        dskip = rest[int(want_stats)] if passthrough else None      # gradient that arrived through the passthrough output
        x, y = ctx.saved_tensors
        weight, bias = ctx.weight, ctx.bias
        kind, stride, ph, pw, relu = ctx.cfg
        N, H, W, Cin = x.shape
        Cout, R, S, _ = weight.shape
        Ho, Wo = F.conv_out_size(H, R, stride, ph), F.conv_out_size(W, S, stride, pw)
        M = N * Ho * Wo
        dy = dy.contiguous()
        if relu:
            dy = F.act_bwd(dy, y, F.ACT_RELU)
        w = shadow_of(weight)
        if bias is not None and bias.requires_grad:
            if Cout % 8 == 0:
                accumulate_bgrad(bias, dy.view(M, Cout))
            else:
                grad_buffer(bias).add_(dy.view(M, Cout).float().sum(0))
        dx = None
        need_dx = ctx.needs_input_grad[0]
        if kind == "gemm":
            if weight.requires_grad:
                F.gemm(dy.view(M, Cout), x.view(M, Cin), grad_buffer(weight).view(Cout, Cin), Cout, Cin, M,
                       a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)
            if need_dx:
                dx = torch.empty_like(x)
                if dskip is not None:      # dx = dy W + dskip in one kernel
                    F.gemm(dy.view(M, Cout), w.view(Cout, Cin), dx.view(M, Cin), M, Cin, Cout, b_mn=True, epi=F.EPI_BIAS_RESID,
                           aux_in=dskip.contiguous().view(M, Cin))
                    dskip = None
                else:
                    F.gemm(dy.view(M, Cout), w.view(Cout, Cin), dx.view(M, Cin), M, Cin, Cout, b_mn=True)
        elif kind == "implicit":
            if weight.requires_grad:
                F.conv2d_implicit(F.CONV_WGRAD, dy, x, grad_buffer(weight), N, H, W, Cin, Cout, R, S, ph, pw)
            if need_dx:
                dx = torch.empty_like(x)
                F.conv2d_implicit(F.CONV_DGRAD, dy, w, dx, N, H, W, Cin, Cout, R, S, ph, pw)
        else:
            K = R * S * Cin
            Kp, Cp = _round8(K), _round8(Cout)
            dyp = _pad_cols(dy.view(M, Cout), Cp)
            if weight.requires_grad:
                col = getattr(ctx, "col", None)
                if col is None:   # recomputed: cheaper than keeping it alive through the whole forward/backward
                    col = F.im2col(x, N, H, W, Cin, R, S, stride, ph, pw, Kp)
                g = grad_buffer(weight).view(Cout, K)
                if Kp == K and Cp == Cout:
                    F.gemm(dyp, col, g, Cout, K, M, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)
                else:
                    tmp = torch.zeros(Cp, Kp, dtype=torch.float32, device=x.device)
                    F.gemm(dyp, col, tmp, Cp, Kp, M, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)
                    g.add_(tmp[:Cout, :K])
            if need_dx:
                wk = _pad_cols(w.view(Cout, K), Kp)
                if Cp != Cout:
                    wk = torch.cat([wk, torch.zeros(Cp - Cout, Kp, dtype=wk.dtype, device=wk.device)], 0)
                dcol = torch.empty(M, Kp, dtype=torch.bfloat16, device=x.device)
                F.gemm(dyp, wk, dcol, M, Kp, Cp, b_mn=True)
                dx = F.col2im(dcol, N, H, W, Cin, R, S, stride, ph, pw, Kp)
        if dskip is not None:
            dx = dskip if dx is None else dx + dskip
        notify_grads_written(weight, bias)
        return dx, None, None, None, None, None, None, None


class Conv2d(nn.Module):
    """2-D convolution (square dilation-1, groups-1) with optional fused bias + ReLU. Weight is ``[Cout, R, S, Cin]``."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size, stride=1, padding=0, bias: bool = True,
                 activation: Optional[str] = None) -> None:
        super().__init__()
        if activation not in (None, "relu"):
            raise ValueError("Conv2d fuses only ReLU; apply other activations as modules")
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size, self.padding = _pair(kernel_size), _pair(padding)
        st = _pair(stride)
        if st[0] != st[1]:
            raise ValueError("Conv2d needs equal strides in both dimensions")
        self.stride = st[0]
        self.activation = activation
        R, S = self.kernel_size
        ref = nn.Conv2d(in_channels, out_channels, (R, S), bias=bias)  # torch's default initialisation
        self.weight = nn.Parameter(ref.weight.detach().permute(0, 2, 3, 1).contiguous())
        self.bias = nn.Parameter(ref.bias.detach().clone()) if bias else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda:
            y = nn.functional.conv2d(x.float(), self.weight.permute(0, 3, 1, 2), self.bias, self.stride, self.padding)
            return torch.relu(y) if self.activation == "relu" else y
        y = _Conv2dFn.apply(to_nhwc(x), self.weight, self.bias, self.stride, self.padding, self.activation == "relu", False)
        return from_nhwc(y)

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        nn.Module._save_to_state_dict(self, destination, prefix, keep_vars)
        k = prefix + "weight"
        if k in destination:
            destination[k] = destination[k].detach().permute(0, 3, 1, 2).contiguous()

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        k = prefix + "weight"
        if k in state_dict and state_dict[k].dim() == 4:
            state_dict = dict(state_dict)
            state_dict[k] = state_dict[k].permute(0, 2, 3, 1).contiguous()
        nn.Module._load_from_state_dict(self, state_dict, prefix, *args, **kwargs)

    def extra_repr(self) -> str:
        return (f"{self.in_channels}, {self.out_channels}, kernel_size={self.kernel_size}, stride={self.stride}, "
                f"padding={self.padding}, activation={self.activation}")


# ------------------------------------------------------------------------------------------------
# Depthwise convolution (groups == channels): stride, padding, dilation
# ------------------------------------------------------------------------------------------------
class _DepthwiseFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, pad, dil, relu):
        R, S, _ = weight.shape
        y = F.dwconv2d_fwd(x, weight, bias, R, S, stride, pad, dil, relu)
        ctx.save_for_backward(x, y if relu else None)
        ctx.weight, ctx.bias, ctx.cfg = weight, bias, (R, S, stride, pad, dil, relu)
        note_param_use(ctx, weight, bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y = ctx.saved_tensors
        R, S, stride, pad, dil, relu = ctx.cfg
        dy = dy.contiguous()
        if relu:
            dy = F.act_bwd(dy, y, F.ACT_RELU)
        F.dwconv2d_wgrad(x, dy, grad_buffer(ctx.weight), R, S, stride, pad, dil)
        if ctx.bias is not None:
            accumulate_bgrad(ctx.bias, dy.view(-1, dy.shape[-1]))
        dx = F.dwconv2d_dgrad(dy, ctx.weight, tuple(x.shape), R, S, stride, pad, dil) if ctx.needs_input_grad[0] else None
        notify_grads_written(ctx.weight, ctx.bias)
        return dx, None, None, None, None, None, None


class DepthwiseConv2d(nn.Module):
    """``nn.Conv2d(C, C, k, groups=C)`` with stride / padding / dilation — the building block of separable and dilated
    convolutions (DARTS ``SepConv`` / ``DilConv``, MobileNet blocks). No channel reduction means no tensor-core work: the
    kernels (``dwconv_ops.cu``) are bandwidth-shaped. Weight is ``[R, S, C]`` fp32 (tap-major, read directly by the
    kernels); ``state_dict`` speaks torch's ``[C, 1, R, S]``."""

    def __init__(self, channels: int, kernel_size, stride: int = 1, padding: int = 0, dilation: int = 1, bias: bool = False,
                 activation: Optional[str] = None) -> None:
        super().__init__()
        if activation not in (None, "relu"):
            raise ValueError("DepthwiseConv2d fuses only ReLU")
        self.channels, self.kernel_size = channels, _pair(kernel_size)
        self.stride, self.padding, self.dilation, self.activation = int(stride), int(padding), int(dilation), activation
        ref = nn.Conv2d(channels, channels, self.kernel_size, groups=channels, bias=bias)
        self.weight = nn.Parameter(ref.weight.detach()[:, 0].permute(1, 2, 0).contiguous())
        self.bias = nn.Parameter(ref.bias.detach().clone()) if bias else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda or self.channels % 8 != 0:
            w = self.weight.permute(2, 0, 1).unsqueeze(1)
            y = nn.functional.conv2d(x.float() if not x.is_cuda else x, w.to(x.dtype if x.is_cuda else torch.float32),
                                     None if self.bias is None else self.bias.to(x.dtype if x.is_cuda else torch.float32),
                                     self.stride, self.padding, self.dilation, groups=self.channels)
            return torch.relu(y) if self.activation == "relu" else y
        y = _DepthwiseFn.apply(to_nhwc(x), self.weight, self.bias, self.stride, self.padding, self.dilation, self.activation == "relu")
        return from_nhwc(y)

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        nn.Module._save_to_state_dict(self, destination, prefix, keep_vars)
        k = prefix + "weight"
        if k in destination:
            destination[k] = destination[k].detach().permute(2, 0, 1).unsqueeze(1).contiguous()

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        k = prefix + "weight"
        if k in state_dict and state_dict[k].dim() == 4:
            state_dict = dict(state_dict)
            state_dict[k] = state_dict[k][:, 0].permute(1, 2, 0).contiguous()
        nn.Module._load_from_state_dict(self, state_dict, prefix, *args, **kwargs)

    def extra_repr(self) -> str:
        return (f"{self.channels}, kernel_size={self.kernel_size}, stride={self.stride}, padding={self.padding}, "
                f"dilation={self.dilation}, activation={self.activation}")


# ------------------------------------------------------------------------------------------------
# ConvTranspose2d  (= the data-gradient of a convolution)
# ------------------------------------------------------------------------------------------------
class _ConvT2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, pad):
        N, Hx, Wx, Cin = x.shape
        _, R, S, Cout = weight.shape
        Hy, Wy = (Hx - 1) * stride - 2 * pad + R, (Wx - 1) * stride - 2 * pad + S
        M = N * Hx * Wx
        K = R * S * Cout
        Kp, Cp = _round8(K), _round8(Cin)
        w = shadow_of(weight).view(Cin, K)
        wk = _pad_cols(w, Kp)
        if Cp != Cin:
            wk = torch.cat([wk, torch.zeros(Cp - Cin, Kp, dtype=wk.dtype, device=wk.device)], 0)
        xp = _pad_cols(x.reshape(M, Cin), Cp)
        col = torch.empty(M, Kp, dtype=torch.bfloat16, device=x.device)
        F.gemm(xp, wk, col, M, Kp, Cp, b_mn=True)                       # col[pixel_x, (r,s,co)] = x · W
        y = F.col2im(col, N, Hy, Wy, Cout, R, S, stride, pad, pad, Kp)   # overlap-add into the output image
        if bias is not None:
            y = y + bias.to(torch.bfloat16)
        ctx.save_for_backward(x)
        ctx.weight, ctx.bias = weight, bias
        ctx.cfg = (stride, pad, Hy, Wy)
        note_param_use(ctx, weight, bias)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        weight, bias = ctx.weight, ctx.bias
        stride, pad, Hy, Wy = ctx.cfg
        N, Hx, Wx, Cin = x.shape
        _, R, S, Cout = weight.shape
        M = N * Hx * Wx
        K = R * S * Cout
        Kp, Cp = _round8(K), _round8(Cin)
        dy = dy.contiguous()
        if bias is not None and bias.requires_grad:
            grad_buffer(bias).add_(dy.reshape(-1, Cout).float().sum(0))
        col = F.im2col(dy, N, Hy, Wy, Cout, R, S, stride, pad, pad, Kp)   # [pixel_x, (r,s,co)] of dy
        if weight.requires_grad:
            xp = _pad_cols(x.reshape(M, Cin), Cp)
            g = grad_buffer(weight).view(Cin, K)
            if Kp == K and Cp == Cin:
                F.gemm(xp, col, g, Cin, K, M, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)
            else:
                tmp = torch.zeros(Cp, Kp, dtype=torch.float32, device=x.device)
                F.gemm(xp, col, tmp, Cp, Kp, M, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)
                g.add_(tmp[:Cin, :K])
        dx = None
        if ctx.needs_input_grad[0]:
            wk = _pad_cols(shadow_of(weight).view(Cin, K), Kp)
            if Cp != Cin:
                wk = torch.cat([wk, torch.zeros(Cp - Cin, Kp, dtype=wk.dtype, device=wk.device)], 0)
            dxp = torch.empty(M, Cp, dtype=torch.bfloat16, device=x.device)
            F.gemm(col, wk, dxp, M, Cin, Kp)
            dx = (dxp if Cp == Cin else dxp[:, :Cin].contiguous()).view(N, Hx, Wx, Cin)
        notify_grads_written(weight, bias)
        return dx, None, None, None, None


class ConvTranspose2d(nn.Module):
    """Transposed convolution (output_padding 0, dilation 1, groups 1). Weight is ``[Cin, R, S, Cout]``."""

    def __init__(self, in_channels: int, out_channels: int, kernel_size, stride=1, padding=0, bias: bool = True) -> None:
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.kernel_size = _pair(kernel_size)
        self.stride, self.padding = int(stride), int(padding)
        ref = nn.ConvTranspose2d(in_channels, out_channels, self.kernel_size, bias=bias)
        self.weight = nn.Parameter(ref.weight.detach().permute(0, 2, 3, 1).contiguous())
        self.bias = nn.Parameter(ref.bias.detach().clone()) if bias else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda:
            return nn.functional.conv_transpose2d(x.float(), self.weight.permute(0, 3, 1, 2), self.bias, self.stride, self.padding)
        return from_nhwc(_ConvT2dFn.apply(to_nhwc(x), self.weight, self.bias, self.stride, self.padding))

    _save_to_state_dict = Conv2d._save_to_state_dict
    _load_from_state_dict = Conv2d._load_from_state_dict


# ------------------------------------------------------------------------------------------------
# BatchNorm2d with fused residual add + activation
# ------------------------------------------------------------------------------------------------
class _BatchNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, resid, gamma, beta, mean, rstd, act, slope):
        C = x.shape[-1]
        y = F.bn_apply(x, resid, mean, rstd, gamma, beta, C, act, slope)
        ctx.save_for_backward(x, y if act != F.ACT_NONE else None, mean, rstd)
        ctx.gamma, ctx.beta = gamma, beta
        ctx.cfg = (act, slope, resid is not None)
        note_param_use(ctx, gamma, beta)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, mean, rstd = ctx.saved_tensors
        act, slope, has_resid = ctx.cfg
        gamma, beta = ctx.gamma, ctx.beta
        dg = grad_buffer(gamma) if gamma is not None and gamma.requires_grad else None
        db = grad_buffer(beta) if beta is not None and beta.requires_grad else None
        dx, dres = F.bn_bwd(dy.contiguous(), x, y, mean, rstd, gamma, dg, db, x.shape[-1], act, slope,
                            want_dresid=has_resid and ctx.needs_input_grad[1])
        if has_resid and ctx.needs_input_grad[1] and act == F.ACT_NONE:
            dres = dy
        notify_grads_written(gamma, beta)
        return dx, dres, None, None, None, None, None, None


class _FrozenBatchNormFn(torch.autograd.Function):
    """Eval-mode BatchNorm (running statistics are constants): dx = gamma * rstd * dy through the activation."""

    @staticmethod
    def forward(ctx, x, resid, gamma, beta, mean, rstd, act, slope):
        y = F.bn_apply(x, resid, mean, rstd, gamma, beta, x.shape[-1], act, slope)
        ctx.save_for_backward(y if act != F.ACT_NONE else None, rstd, x, mean)
        ctx.gamma, ctx.beta, ctx.cfg = gamma, beta, (act, slope, resid is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        y, rstd, x, mean = ctx.saved_tensors
        act, slope, has_resid = ctx.cfg
        dy = dy.contiguous()
        if act != F.ACT_NONE:
            dy = F.act_bwd(dy, y, act, slope)
        C = dy.shape[-1]
        if ctx.beta is not None and ctx.beta.requires_grad:      # rare path (fine-tuning with frozen statistics)
            grad_buffer(ctx.beta).add_(dy.reshape(-1, C).float().sum(0))
        if ctx.gamma is not None and ctx.gamma.requires_grad:
            xhat = (x.reshape(-1, C).float() - mean) * rstd
            grad_buffer(ctx.gamma).add_((dy.reshape(-1, C).float() * xhat).sum(0))
        scale = rstd if ctx.gamma is None else rstd * ctx.gamma.detach()
        dx = (dy.float() * scale).to(torch.bfloat16)
        return dx, (dy if has_resid else None), None, None, None, None, None, None


class BatchNorm2d(nn.BatchNorm2d):
    """``nn.BatchNorm2d`` with an optional fused epilogue: ``y = act(bn(x) + residual)``.

    Training mode computes batch statistics in one column-reduction pass and applies them in a second; the backward is
    a reduction pass plus one pass that writes both ``dx`` and the residual-branch gradient. Statistics are per
    process, exactly like the reference under Horovod (no cross-GPU BatchNorm there either)."""

    def __init__(self, num_features: int, eps: float = 1e-5, momentum: float = 0.1, affine: bool = True,
                 track_running_stats: bool = True, activation: Optional[str] = None, negative_slope: float = 0.01) -> None:
        super().__init__(num_features, eps=eps, momentum=momentum, affine=affine, track_running_stats=track_running_stats)
        if activation not in (None, "relu", "leaky_relu"):
            raise ValueError("BatchNorm2d fuses relu or leaky_relu")
        self.activation, self.negative_slope = activation, float(negative_slope)

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        C = self.num_features
        if not x.is_cuda or C % 8 != 0 or (C > 2048 and C % 2048 != 0):
            y = super().forward(x.float() if not x.is_cuda else x)
            if residual is not None:
                y = y + residual
            if self.activation == "relu":
                y = torch.relu(y)
            elif self.activation == "leaky_relu":
                y = nn.functional.leaky_relu(y, self.negative_slope)
            return y
        xn = to_nhwc(x)
        rn = to_nhwc(residual) if residual is not None else None
        act = _ACT_CODES[self.activation]
        use_batch = self.training or not self.track_running_stats
        if use_batch:
            mom = 0.0 if self.momentum is None else self.momentum
            track = self.training and self.track_running_stats
            if track and self.num_batches_tracked is not None:
                self.num_batches_tracked.add_(1)
            mean, rstd = F.bn_stats(xn, C, self.running_mean if track else None, self.running_var if track else None,
                                    self.eps, mom)
            y = _BatchNormFn.apply(xn, rn, self.weight, self.bias, mean, rstd, act, self.negative_slope)
        else:
            rstd = torch.rsqrt(self.running_var + self.eps)
            y = _FrozenBatchNormFn.apply(xn, rn, self.weight, self.bias, self.running_mean, rstd, act, self.negative_slope)
        return from_nhwc(y)


def conv_bn(conv: "Conv2d", bn: "BatchNorm2d", x: torch.Tensor, residual: Optional[torch.Tensor] = None, passthrough: bool = False):
    """``bn(conv(x), residual)`` with the BatchNorm batch statistics taken from the convolution's own epilogue: the
    kernel that writes the convolution output also accumulates per-channel sum and sum of squares, so training-mode
    BatchNorm needs no statistics pass over that tensor. Falls back to the two separate modules whenever that is not
    possible (CPU, eval mode, fused ReLU in the convolution, unaligned channel counts)."""
    C = bn.num_features
    use_batch = bn.training or not bn.track_running_stats
    if not (x.is_cuda and use_batch and conv.activation is None and C % 8 == 0 and (C <= 2048 or C % 2048 == 0)):
        out = bn(conv(x), residual)
        return (out, x) if passthrough else out
    res = _Conv2dFn.apply(to_nhwc(x), conv.weight, conv.bias, conv.stride, conv.padding, False, True, passthrough)
    y, sums = res[0], res[1]
    rn = to_nhwc(residual) if residual is not None else None
    track = bn.training and bn.track_running_stats
    if track and bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)
    mom = 0.0 if bn.momentum is None else bn.momentum
    rm, rv = (bn.running_mean, bn.running_var) if track else (None, None)
    if sums.numel() == 2 * C:
        mean, rstd = F.bn_finalize(sums, y.numel() // C, C, rm, rv, bn.eps, mom)
    else:
        mean, rstd = F.bn_stats(y, C, rm, rv, bn.eps, mom)
    out = _BatchNormFn.apply(y, rn, bn.weight, bn.bias, mean, rstd, _ACT_CODES[bn.activation], bn.negative_slope)
    return (from_nhwc(out), from_nhwc(res[2])) if passthrough else from_nhwc(out)


# ------------------------------------------------------------------------------------------------
# Pooling, activations, dropout
# ------------------------------------------------------------------------------------------------
class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad):
        y, idx = F.maxpool_fwd(x, k, stride, pad)
        ctx.save_for_backward(idx)
        ctx.cfg = (tuple(x.shape), k, stride, pad)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        shape, k, stride, pad = ctx.cfg
        return F.maxpool_bwd(dy.contiguous(), idx, shape, k, stride, pad), None, None, None


class MaxPool2d(nn.Module):
    def __init__(self, kernel_size: int, stride: Optional[int] = None, padding: int = 0) -> None:
        super().__init__()
        self.kernel_size, self.stride, self.padding = int(kernel_size), int(stride or kernel_size), int(padding)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda or x.shape[1] % 8 != 0:
            return nn.functional.max_pool2d(x, self.kernel_size, self.stride, self.padding)
        return from_nhwc(_MaxPoolFn.apply(to_nhwc(x), self.kernel_size, self.stride, self.padding))


class _AvgPool2dFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad, include_pad):
        ctx.cfg = (tuple(x.shape), k, stride, pad, include_pad)
        return F.avgpool2d_fwd(x, k, stride, pad, include_pad)

    @staticmethod
    def backward(ctx, dy):
        shape, k, stride, pad, include_pad = ctx.cfg
        return F.avgpool2d_bwd(dy.contiguous(), shape, k, stride, pad, include_pad), None, None, None, None


class AvgPool2d(nn.Module):
    """``nn.AvgPool2d`` (square window) on the NHWC kernels."""

    def __init__(self, kernel_size: int, stride: Optional[int] = None, padding: int = 0, count_include_pad: bool = True) -> None:
        super().__init__()
        self.kernel_size, self.stride, self.padding = int(kernel_size), int(stride or kernel_size), int(padding)
        self.count_include_pad = bool(count_include_pad)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda or x.shape[1] % 8 != 0:
            return nn.functional.avg_pool2d(x, self.kernel_size, self.stride, self.padding, count_include_pad=self.count_include_pad)
        return from_nhwc(_AvgPool2dFn.apply(to_nhwc(x), self.kernel_size, self.stride, self.padding, self.count_include_pad))


class _AvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.shape = tuple(x.shape)
        return F.avgpool_fwd(x)

    @staticmethod
    def backward(ctx, dy):
        return F.avgpool_bwd(dy.contiguous(), ctx.shape)


class GlobalAvgPool2d(nn.Module):
    """``AdaptiveAvgPool2d(1)`` + ``flatten(1)``: [N, C, H, W] → [N, C]."""

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda or x.shape[1] % 8 != 0:
            return x.float().mean(dim=(2, 3)) if not x.is_cuda else x.mean(dim=(2, 3))
        return _AvgPoolFn.apply(to_nhwc(x))


class _ActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, act, slope):
        y = F.act_fwd(x, act, slope)
        ctx.save_for_backward(y)
        ctx.cfg = (act, slope)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return F.act_bwd(dy.contiguous(), y, *ctx.cfg), None, None


class Activation(nn.Module):
    """relu / leaky_relu / tanh / sigmoid as one elementwise kernel each way (backward reads the saved output)."""

    def __init__(self, kind: str, negative_slope: float = 0.01) -> None:
        super().__init__()
        if kind not in ("relu", "leaky_relu", "tanh", "sigmoid"):
            raise ValueError(f"unknown activation {kind!r}")
        self.kind, self.negative_slope = kind, float(negative_slope)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda or x.dtype != torch.bfloat16:
            fn = {"relu": torch.relu, "tanh": torch.tanh, "sigmoid": torch.sigmoid,
                  "leaky_relu": lambda t: nn.functional.leaky_relu(t, self.negative_slope)}[self.kind]
            return fn(x)
        if x.dim() == 4:   # keep the channels-last data in place
            return from_nhwc(_ActFn.apply(to_nhwc(x), _ACT_CODES[self.kind], self.negative_slope))
        return _ActFn.apply(x.contiguous(), _ACT_CODES[self.kind], self.negative_slope)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed):
        ctx.cfg = (p, seed)
        return F.dropout(x, p, seed)

    @staticmethod
    def backward(ctx, dy):
        return F.dropout(dy.contiguous(), *ctx.cfg), None, None


_dropout_sites = 0


class Dropout(nn.Module):
    """Dropout whose mask is a stateless hash of (site seed, per-step device seed, element index): nothing is stored
    for backward and CUDA-graph replays draw fresh masks (advance with ``functional.advance_step_seed``)."""

    def __init__(self, p: float = 0.5) -> None:
        super().__init__()
        global _dropout_sites
        _dropout_sites += 1
        self.p, self.seed = float(p), 0x5EED0000 + 7919 * _dropout_sites

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not self.training or self.p == 0.0:
            return x
        if not x.is_cuda or x.dtype != torch.bfloat16:
            return nn.functional.dropout(x, self.p, True)
        if x.dim() == 4:
            return from_nhwc(_DropoutFn.apply(to_nhwc(x), self.p, self.seed))
        return _DropoutFn.apply(x.contiguous(), self.p, self.seed)
