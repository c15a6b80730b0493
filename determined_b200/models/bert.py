"""BERT-base encoder + sequence-classification head (GLUE shape: seq 128) on the sm_100a kernels.

Architecture and default sizes follow ``bert-base-uncased`` as fine-tuned by the reference's
``examples/nlp/bert_glue_pytorch/model_def.py:27-174`` (HF ``BertForSequenceClassification``): post-LN
encoder, erf-GELU, learned position/type embeddings, tanh pooler, dropout 0.1.

On CUDA each encoder layer runs as ONE autograd node (:class:`_BertLayerFn`) that launches the
hand-written kernels directly:

    fwd   QKV GEMM(+bias) → fused attention → out-proj GEMM(+bias+dropout+residual) → LN →
          FFN1 GEMM(+bias+GELU, saves pre-activation) → FFN2 GEMM(+bias+dropout+residual) → LN
    bwd   LN-bwd (emits both the residual gradient and the dropout-masked gradient) → wgrad GEMMs
          accumulating into the flat fp32 gradient bucket → dgrad GEMMs with dGELU / residual-add
          epilogues → fused attention backward

so no elementwise, transpose, head-split or mask kernels are ever launched. On CPU (tests, local
mode) the same parameters run through a plain fp32 PyTorch statement of the same math, which is
also the numerics oracle for the CUDA path.
"""
from __future__ import annotations

import dataclasses
import math
import os
from typing import Optional, Tuple

import torch
import torch.nn as nn

from ..ops import functional as F
from ..ops.modules import (accumulate_bgrad, accumulate_wgrad, grad_buffer, note_param_use, notify_grads_written, side_grads,
                           shadow_of)


@dataclasses.dataclass
class BertConfig:
    vocab_size: int = 30522
    hidden_size: int = 768
    num_hidden_layers: int = 12
    num_attention_heads: int = 12
    intermediate_size: int = 3072
    max_position_embeddings: int = 512
    type_vocab_size: int = 2
    hidden_dropout_prob: float = 0.1
    attention_probs_dropout_prob: float = 0.1
    layer_norm_eps: float = 1e-12
    initializer_range: float = 0.02
    num_labels: int = 2


# The LayerNorm backward kernel can also emit the bias gradient of the GEMM in front of it (column sums of its output).
# Measured on B200 it LOSES to the separate column-sum kernel — 25.61 vs 24.71 ms per BERT-base step — because the sums go
# through shared-memory atomics inside the row loop; it stays opt-in (DET_LN_FUSED_BIAS=1) for shapes where launches dominate.
_FUSED_BIAS = os.environ.get("DET_LN_FUSED_BIAS", "0") == "1"


def _gelu(x: torch.Tensor) -> torch.Tensor:
    # tanh form, as in the original BERT code and in the GEMM epilogue (csrc/kernels/gemm_common.cuh)
    return nn.functional.gelu(x, approximate="tanh")


# ------------------------------------------------------------------------------------------------
# Encoder layer
# ------------------------------------------------------------------------------------------------
class _BertLayerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, mask, layer: "BertLayer", B: int, S: int, training: bool, seed_base: Optional[int] = None):
        c = layer.cfg
        T, Hd, I, NH = x.shape[0], c.hidden_size, c.intermediate_size, c.num_attention_heads
        p_h = c.hidden_dropout_prob if training else 0.0
        p_a = c.attention_probs_dropout_prob if training else 0.0
        sd = layer.seed_base if seed_base is None else seed_base   # a layer shared across depths draws a mask per depth
        dev = x.device
        bf = torch.bfloat16
        qkv = torch.empty(T, 3 * Hd, dtype=bf, device=dev)
        F.gemm(x, shadow_of(layer.qkv_weight), qkv, T, 3 * Hd, Hd, epi=F.EPI_BIAS, bias=layer.qkv_bias)
        att, lse = F.attention_fwd(qkv, mask, B, S, NH, p_a, sd + 0)
        h1 = torch.empty(T, Hd, dtype=bf, device=dev)
        F.gemm(att, shadow_of(layer.out_weight), h1, T, Hd, Hd, epi=F.EPI_BIAS_RESID, bias=layer.out_bias, aux_in=x,
               dropout_p=p_h, seed=sd + 1)
        x1, mean1, rstd1 = F.layernorm_fwd(h1, layer.ln1_weight, layer.ln1_bias, c.layer_norm_eps)
        z = torch.empty(T, I, dtype=bf, device=dev)
        a = torch.empty(T, I, dtype=bf, device=dev)
        F.gemm(x1, shadow_of(layer.ffn1_weight), a, T, I, Hd, epi=F.EPI_BIAS_GELU, bias=layer.ffn1_bias, aux_out=z)
        h2 = torch.empty(T, Hd, dtype=bf, device=dev)
        F.gemm(a, shadow_of(layer.ffn2_weight), h2, T, Hd, I, epi=F.EPI_BIAS_RESID, bias=layer.ffn2_bias, aux_in=x1,
               dropout_p=p_h, seed=sd + 2)
        x2, mean2, rstd2 = F.layernorm_fwd(h2, layer.ln2_weight, layer.ln2_bias, c.layer_norm_eps)
        ctx.save_for_backward(x, qkv, lse, att, h1, mean1, rstd1, x1, z, a, h2, mean2, rstd2)
        ctx.layer, ctx.mask, ctx.dims, ctx.drop, ctx.sd = layer, mask, (B, S), (p_h, p_a), sd
        note_param_use(ctx, *layer.parameters())
        return x2

    @staticmethod
    def backward(ctx, dx2):
        x, qkv, lse, att, h1, mean1, rstd1, x1, z, a, h2, mean2, rstd2 = ctx.saved_tensors
        L: BertLayer = ctx.layer
        c = L.cfg
        B, S = ctx.dims
        p_h, p_a = ctx.drop
        sd = ctx.sd
        T, Hd, I, NH = x.shape[0], c.hidden_size, c.intermediate_size, c.num_attention_heads
        dev, bf = x.device, torch.bfloat16
        dx2 = dx2.contiguous()
        with side_grads(dev):    # weight / bias gradients run beside the data-gradient chain, joined on the way out
            out = _BertLayerFn._backward_body(ctx, dx2, (x, qkv, lse, att, h1, mean1, rstd1, x1, z, a, h2, mean2, rstd2))
        notify_grads_written(*L.parameters())   # this layer's slice of the gradient bucket is final: its all-reduce may start
        return out

    @staticmethod
    def _backward_body(ctx, dx2, saved):
        x, qkv, lse, att, h1, mean1, rstd1, x1, z, a, h2, mean2, rstd2 = saved
        L: BertLayer = ctx.layer
        c = L.cfg
        B, S = ctx.dims
        p_h, p_a = ctx.drop
        sd = ctx.sd
        T, Hd, I, NH = x.shape[0], c.hidden_size, c.intermediate_size, c.num_attention_heads
        dev, bf = x.device, torch.bfloat16
        masked = p_h > 0.0
        # LN2 backward: dh2 (residual branch) and dh2m (dropout-masked: gradient of FFN2's pre-dropout output)
        dh2, dh2m = F.layernorm_bwd(dx2, h2, L.ln2_weight, mean2, rstd2, grad_buffer(L.ln2_weight),
                                    grad_buffer(L.ln2_bias), want_masked=masked, out_dropout_p=p_h, out_seed=sd + 2,
                                    dxsum=grad_buffer(L.ffn2_bias) if _FUSED_BIAS else None)
        if not masked:
            dh2m = dh2
        if not _FUSED_BIAS:
            accumulate_bgrad(L.ffn2_bias, dh2m)
        accumulate_wgrad(L.ffn2_weight, dh2m, a)
        dz = torch.empty(T, I, dtype=bf, device=dev)
        F.gemm(dh2m, shadow_of(L.ffn2_weight), dz, T, I, Hd, b_mn=True, epi=F.EPI_DGELU, aux_in=z)
        accumulate_wgrad(L.ffn1_weight, dz, x1)
        accumulate_bgrad(L.ffn1_bias, dz)
        dx1 = torch.empty(T, Hd, dtype=bf, device=dev)
        F.gemm(dz, shadow_of(L.ffn1_weight), dx1, T, Hd, I, b_mn=True, epi=F.EPI_BIAS_RESID, aux_in=dh2)
        dh1, dh1m = F.layernorm_bwd(dx1, h1, L.ln1_weight, mean1, rstd1, grad_buffer(L.ln1_weight),
                                    grad_buffer(L.ln1_bias), want_masked=masked, out_dropout_p=p_h, out_seed=sd + 1,
                                    dxsum=grad_buffer(L.out_bias) if _FUSED_BIAS else None)
        if not masked:
            dh1m = dh1
        if not _FUSED_BIAS:
            accumulate_bgrad(L.out_bias, dh1m)
        accumulate_wgrad(L.out_weight, dh1m, att)
        datt = torch.empty(T, Hd, dtype=bf, device=dev)
        F.gemm(dh1m, shadow_of(L.out_weight), datt, T, Hd, Hd, b_mn=True)
        dqkv = F.attention_bwd(qkv, datt, ctx.mask, lse, B, S, NH, p_a, sd + 0, att)
        accumulate_wgrad(L.qkv_weight, dqkv, x)
        accumulate_bgrad(L.qkv_bias, dqkv)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(T, Hd, dtype=bf, device=dev)
            F.gemm(dqkv, shadow_of(L.qkv_weight), dx, T, Hd, 3 * Hd, b_mn=True, epi=F.EPI_BIAS_RESID, aux_in=dh1)
        return dx, None, None, None, None, None, None


class BertLayer(nn.Module):
    def __init__(self, cfg: BertConfig, index: int) -> None:
        super().__init__()
        self.cfg = cfg
        self.seed_base = 0x1000 + 16 * index
        Hd, I = cfg.hidden_size, cfg.intermediate_size
        std = cfg.initializer_range
        self.qkv_weight = nn.Parameter(torch.randn(3 * Hd, Hd) * std)
        self.qkv_bias = nn.Parameter(torch.zeros(3 * Hd))
        self.out_weight = nn.Parameter(torch.randn(Hd, Hd) * std)
        self.out_bias = nn.Parameter(torch.zeros(Hd))
        self.ln1_weight = nn.Parameter(torch.ones(Hd))
        self.ln1_bias = nn.Parameter(torch.zeros(Hd))
        self.ffn1_weight = nn.Parameter(torch.randn(I, Hd) * std)
        self.ffn1_bias = nn.Parameter(torch.zeros(I))
        self.ffn2_weight = nn.Parameter(torch.randn(Hd, I) * std)
        self.ffn2_bias = nn.Parameter(torch.zeros(Hd))
        self.ln2_weight = nn.Parameter(torch.ones(Hd))
        self.ln2_bias = nn.Parameter(torch.zeros(Hd))

    def forward(self, x: torch.Tensor, mask: Optional[torch.Tensor], B: int, S: int, seed_base: Optional[int] = None) -> torch.Tensor:
        if x.is_cuda:
            return _BertLayerFn.apply(x, mask, self, B, S, self.training, seed_base)
        return self._reference_forward(x, mask, B, S)

    def _reference_forward(self, x, mask, B, S):
        c = self.cfg
        Hd, NH = c.hidden_size, c.num_attention_heads
        drop = nn.functional.dropout
        xf = x.float()
        qkv = xf @ self.qkv_weight.t() + self.qkv_bias
        q, k, v = [t.view(B, S, NH, 64).transpose(1, 2) for t in qkv.split(Hd, dim=-1)]
        sc = q @ k.transpose(-1, -2) / math.sqrt(64.0)
        if mask is not None:
            sc = sc + mask[:, None, None, :].float()
        p = drop(torch.softmax(sc, dim=-1), c.attention_probs_dropout_prob, self.training)
        att = (p @ v).transpose(1, 2).reshape(B * S, Hd)
        h1 = drop(att @ self.out_weight.t() + self.out_bias, c.hidden_dropout_prob, self.training) + xf
        x1 = nn.functional.layer_norm(h1, (Hd,), self.ln1_weight, self.ln1_bias, c.layer_norm_eps)
        a = _gelu(x1 @ self.ffn1_weight.t() + self.ffn1_bias)
        h2 = drop(a @ self.ffn2_weight.t() + self.ffn2_bias, c.hidden_dropout_prob, self.training) + x1
        return nn.functional.layer_norm(h2, (Hd,), self.ln2_weight, self.ln2_bias, c.layer_norm_eps).to(x.dtype)


# ------------------------------------------------------------------------------------------------
# Embeddings
# ------------------------------------------------------------------------------------------------
class _BertEmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, tts, emb: "BertEmbeddings", S: int, training: bool, anchor):
        c = emb.cfg
        p = c.hidden_dropout_prob if training else 0.0
        e = F.embed_gather(ids, tts, emb.word_embeddings, emb.position_embeddings, emb.token_type_embeddings, S)
        x0, mean, rstd = F.layernorm_fwd(e, emb.ln_weight, emb.ln_bias, c.layer_norm_eps, p, emb.seed)
        ctx.save_for_backward(ids, tts, e, mean, rstd)
        ctx.emb, ctx.S, ctx.p = emb, S, p
        note_param_use(ctx, *emb.parameters())
        return x0

    @staticmethod
    def backward(ctx, dx0):
        ids, tts, e, mean, rstd = ctx.saved_tensors
        emb: BertEmbeddings = ctx.emb
        de, _ = F.layernorm_bwd(dx0.contiguous(), e, emb.ln_weight, mean, rstd, grad_buffer(emb.ln_weight),
                                grad_buffer(emb.ln_bias), in_dropout_p=ctx.p, in_seed=emb.seed)
        F.embed_scatter(ids, tts, de, grad_buffer(emb.word_embeddings), grad_buffer(emb.position_embeddings),
                        grad_buffer(emb.token_type_embeddings), ctx.S)
        notify_grads_written(*emb.parameters())
        return None, None, None, None, None, None


class BertEmbeddings(nn.Module):
    def __init__(self, cfg: BertConfig) -> None:
        super().__init__()
        self.cfg = cfg
        self.seed = 0x0777
        std = cfg.initializer_range
        self.word_embeddings = nn.Parameter(torch.randn(cfg.vocab_size, cfg.hidden_size) * std)
        self.position_embeddings = nn.Parameter(torch.randn(cfg.max_position_embeddings, cfg.hidden_size) * std)
        self.token_type_embeddings = nn.Parameter(torch.randn(cfg.type_vocab_size, cfg.hidden_size) * std)
        self.ln_weight = nn.Parameter(torch.ones(cfg.hidden_size))
        self.ln_bias = nn.Parameter(torch.zeros(cfg.hidden_size))
        # autograd anchor: gives the embedding node a differentiable input so its backward runs
        self._anchor = nn.Parameter(torch.zeros(1), requires_grad=True)

    def forward(self, ids: torch.Tensor, tts: Optional[torch.Tensor], S: int) -> torch.Tensor:
        ids = ids.reshape(-1).to(torch.int64).contiguous()
        tts = None if tts is None else tts.reshape(-1).to(torch.int64).contiguous()
        if ids.is_cuda:
            return _BertEmbedFn.apply(ids, tts, self, S, self.training, self._anchor)
        pos = torch.arange(ids.numel(), device=ids.device) % S
        e = self.word_embeddings[ids] + self.position_embeddings[pos]
        e = e + self.token_type_embeddings[tts if tts is not None else torch.zeros_like(ids)]
        x = nn.functional.layer_norm(e, (self.cfg.hidden_size,), self.ln_weight, self.ln_bias, self.cfg.layer_norm_eps)
        return nn.functional.dropout(x, self.cfg.hidden_dropout_prob, self.training)


# ------------------------------------------------------------------------------------------------
# Pooler + classifier + loss (tiny: CUDA-core strided GEMM kernel)
# ------------------------------------------------------------------------------------------------
class _BertHeadFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, head: "BertClassificationHead", B: int, S: int, training: bool):
        c = head.cfg
        Hd, C = c.hidden_size, c.num_labels
        dev = x.device
        x_cls = x.view(B, S, Hd)[:, 0]  # strided [B, Hd] view (row pitch S*Hd): TMA reads it in place
        pre = torch.empty(B, Hd, dtype=torch.bfloat16, device=dev)
        F.gemm(x_cls, shadow_of(head.pooler_weight), pre, B, Hd, Hd, epi=F.EPI_BIAS, bias=head.pooler_bias)
        pooled = torch.tanh(pre.float())
        if training and c.hidden_dropout_prob > 0:
            keep = (torch.rand(B, Hd, device=dev) >= c.hidden_dropout_prob).float() / (1.0 - c.hidden_dropout_prob)
            pd = pooled * keep
        else:
            keep, pd = None, pooled
        logits = torch.empty(B, C, dtype=torch.float32, device=dev)
        F.gemm_small(pd, head.classifier_weight, logits, B, C, Hd, (Hd, 1), (Hd, 1), (C, 1), bias=head.classifier_bias)
        ctx.save_for_backward(x, pooled, keep, pd)
        ctx.head, ctx.dims = head, (B, S)
        note_param_use(ctx, *head.parameters())
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        x, pooled, keep, pd = ctx.saved_tensors
        head: BertClassificationHead = ctx.head
        B, S = ctx.dims
        c = head.cfg
        Hd, C = c.hidden_size, c.num_labels
        dlogits = dlogits.float().contiguous()
        F.gemm_small(dlogits, pd, grad_buffer(head.classifier_weight), C, Hd, B, (1, C), (1, Hd), (Hd, 1), accumulate=True)
        grad_buffer(head.classifier_bias).add_(dlogits.sum(0))
        dpd = torch.empty(B, Hd, dtype=torch.float32, device=x.device)
        F.gemm_small(dlogits, head.classifier_weight, dpd, B, Hd, C, (C, 1), (1, Hd), (Hd, 1))
        if keep is not None:
            dpd = dpd * keep
        dpre32 = dpd * (1.0 - pooled * pooled)
        dpre = dpre32.to(torch.bfloat16)
        x_cls = x.view(B, S, Hd)[:, 0]
        F.gemm(dpre, x_cls, grad_buffer(head.pooler_weight), Hd, Hd, B, a_mn=True, b_mn=True,
               out_mode=F.OUT_F32_ATOMIC_ADD)
        grad_buffer(head.pooler_bias).add_(dpre32.sum(0))
        dcls = torch.empty(B, Hd, dtype=torch.bfloat16, device=x.device)
        F.gemm(dpre, shadow_of(head.pooler_weight), dcls, B, Hd, Hd, b_mn=True)
        dx = torch.zeros_like(x)
        dx.view(B, S, Hd)[:, 0] = dcls
        notify_grads_written(*head.parameters())
        return dx, None, None, None, None


class BertClassificationHead(nn.Module):
    def __init__(self, cfg: BertConfig) -> None:
        super().__init__()
        self.cfg = cfg
        std = cfg.initializer_range
        self.pooler_weight = nn.Parameter(torch.randn(cfg.hidden_size, cfg.hidden_size) * std)
        self.pooler_bias = nn.Parameter(torch.zeros(cfg.hidden_size))
        self.classifier_weight = nn.Parameter(torch.randn(cfg.num_labels, cfg.hidden_size) * std)
        self.classifier_bias = nn.Parameter(torch.zeros(cfg.num_labels))

    def forward(self, x: torch.Tensor, B: int, S: int) -> torch.Tensor:
        if x.is_cuda:
            return _BertHeadFn.apply(x, self, B, S, self.training)
        cls = x.view(B, S, -1)[:, 0].float()
        pooled = torch.tanh(cls @ self.pooler_weight.t() + self.pooler_bias)
        pooled = nn.functional.dropout(pooled, self.cfg.hidden_dropout_prob, self.training)
        return pooled @ self.classifier_weight.t() + self.classifier_bias


class BertForSequenceClassification(nn.Module):
    """``forward(input_ids, attention_mask=None, token_type_ids=None, labels=None)`` →
    ``(loss, logits)`` when labels are given, else ``(logits,)`` — the HF calling convention the
    reference's trial uses (``model_def.py:141-151``)."""

    def __init__(self, cfg: Optional[BertConfig] = None) -> None:
        super().__init__()
        self.cfg = cfg or BertConfig()
        if self.cfg.hidden_size != self.cfg.num_attention_heads * 64:
            raise ValueError("the fused attention kernel is built for head_dim 64")
        self.embeddings = BertEmbeddings(self.cfg)
        self.layers = nn.ModuleList([BertLayer(self.cfg, i) for i in range(self.cfg.num_hidden_layers)])
        self.head = BertClassificationHead(self.cfg)

    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                token_type_ids: Optional[torch.Tensor] = None, labels: Optional[torch.Tensor] = None) -> Tuple:
        B, S = input_ids.shape
        mask = None
        if attention_mask is not None:
            mask = ((1.0 - attention_mask.to(torch.float32)) * -10000.0).contiguous()
        x = self.embeddings(input_ids, token_type_ids, S)
        for layer in self.layers:
            x = layer(x, mask, B, S)
        logits = self.head(x, B, S)
        if labels is None:
            return (logits,)
        from ..ops.modules import cross_entropy

        return cross_entropy(logits, labels.reshape(-1)), logits

    def flops_per_sample(self, seq_len: int = 128) -> float:
        c = self.cfg
        per_tok = 2 * (4 * c.hidden_size * c.hidden_size + 2 * c.hidden_size * c.intermediate_size)
        attn = 4 * seq_len * c.hidden_size
        return 3.0 * c.num_hidden_layers * seq_len * (per_tok + attn)


# ------------------------------------------------------------------------------------------------
# Extractive question answering (SQuAD shape: seq 384) — BERT and ALBERT
# ------------------------------------------------------------------------------------------------
class _QAHeadFn(torch.autograd.Function):
    """start/end logits = x W^T + b with W [2, hidden]: two output columns, so the CUDA-core strided GEMM."""

    @staticmethod
    def forward(ctx, x, head: "BertQAHead"):
        T, Hd = x.shape
        logits = torch.empty(T, 2, dtype=torch.float32, device=x.device)
        F.gemm_small(x, head.weight, logits, T, 2, Hd, (Hd, 1), (Hd, 1), (2, 1), bias=head.bias)
        ctx.save_for_backward(x)
        ctx.head = head
        note_param_use(ctx, head.weight, head.bias)
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        (x,) = ctx.saved_tensors
        head: BertQAHead = ctx.head
        T, Hd = x.shape
        dlogits = dlogits.float().contiguous()
        F.gemm_small(dlogits, x, grad_buffer(head.weight), 2, Hd, T, (1, 2), (1, Hd), (Hd, 1), accumulate=True)
        grad_buffer(head.bias).add_(dlogits.sum(0))
        dx = torch.empty(T, Hd, dtype=torch.bfloat16, device=x.device)
        F.gemm_small(dlogits, head.weight, dx, T, Hd, 2, (2, 1), (1, Hd), (Hd, 1))
        notify_grads_written(head.weight, head.bias)
        return dx, None


class BertQAHead(nn.Module):
    def __init__(self, hidden_size: int, std: float) -> None:
        super().__init__()
        self.weight = nn.Parameter(torch.randn(2, hidden_size) * std)
        self.bias = nn.Parameter(torch.zeros(2))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if x.is_cuda:
            return _QAHeadFn.apply(x, self)
        return x.float() @ self.weight.t() + self.bias


def _qa_outputs(logits: torch.Tensor, B: int, S: int, start_positions, end_positions) -> Tuple:
    """[B*S, 2] logits → HF's (loss?, start_logits, end_logits); loss = (CE(start) + CE(end)) / 2, positions clamped
    into the sequence like HF's ``ignored_index`` handling of out-of-span answers."""
    lg = logits.view(B, S, 2)
    start_logits, end_logits = lg[..., 0], lg[..., 1]
    if start_positions is None or end_positions is None:
        return start_logits, end_logits
    from ..ops.modules import cross_entropy

    both = lg.transpose(1, 2).reshape(2 * B, S)                       # row 2b = start logits of b, row 2b+1 = end logits
    labels = torch.stack([start_positions.reshape(-1), end_positions.reshape(-1)], dim=1).reshape(-1).clamp(0, S - 1)
    return cross_entropy(both, labels), start_logits, end_logits


class BertForQuestionAnswering(nn.Module):
    """HF ``BertForQuestionAnswering`` calling convention (reference: ``examples/nlp/bert_squad_pytorch/model_def.py``):
    ``forward(input_ids, attention_mask, token_type_ids, start_positions=None, end_positions=None)``. Sequences longer
    than 128 tokens run the tiled attention kernels."""

    def __init__(self, cfg: Optional[BertConfig] = None) -> None:
        super().__init__()
        self.cfg = cfg or BertConfig()
        if self.cfg.hidden_size != self.cfg.num_attention_heads * 64:
            raise ValueError("the fused attention kernels are built for head_dim 64")
        self.embeddings = BertEmbeddings(self.cfg)
        self.layers = nn.ModuleList([BertLayer(self.cfg, i) for i in range(self.cfg.num_hidden_layers)])
        self.qa_outputs = BertQAHead(self.cfg.hidden_size, self.cfg.initializer_range)

    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                token_type_ids: Optional[torch.Tensor] = None, start_positions: Optional[torch.Tensor] = None,
                end_positions: Optional[torch.Tensor] = None) -> Tuple:
        B, S = input_ids.shape
        mask = None
        if attention_mask is not None:
            mask = ((1.0 - attention_mask.to(torch.float32)) * -10000.0).contiguous()
        x = self.embeddings(input_ids, token_type_ids, S)
        for layer in self.layers:
            x = layer(x, mask, B, S)
        return _qa_outputs(self.qa_outputs(x), B, S, start_positions, end_positions)


@dataclasses.dataclass
class AlbertConfig(BertConfig):
    """ALBERT: factorised embeddings (``embedding_size`` → ``hidden_size`` projection) and ONE encoder layer whose
    parameters are shared by every depth (reference: ``examples/nlp/albert_squad_pytorch``, HF ``AlbertModel``).
    Defaults are albert-base-v2; xxlarge is hidden 4096 / 64 heads / intermediate 16384."""
    vocab_size: int = 30000
    embedding_size: int = 128
    hidden_dropout_prob: float = 0.0
    attention_probs_dropout_prob: float = 0.0


class AlbertForQuestionAnswering(nn.Module):
    """Shared-layer encoder + span head. The layer is the BERT layer of this module (post-LN, erf-GELU; HF ALBERT uses
    the tanh GELU approximation, which differs by < 1e-3 per activation). Gradients of the shared parameters accumulate
    across depths in the flat fp32 bucket exactly like ``aggregation_frequency`` micro-batches do."""

    def __init__(self, cfg: Optional[AlbertConfig] = None) -> None:
        super().__init__()
        from ..ops.modules import Linear

        self.cfg = cfg or AlbertConfig()
        c = self.cfg
        if c.hidden_size != c.num_attention_heads * 64:
            raise ValueError("the fused attention kernels are built for head_dim 64")
        self.embeddings = BertEmbeddings(dataclasses.replace(c, hidden_size=c.embedding_size))
        self.embedding_hidden_mapping_in = Linear(c.embedding_size, c.hidden_size)
        self.shared_layer = BertLayer(c, 0)
        self.qa_outputs = BertQAHead(c.hidden_size, c.initializer_range)

    def forward(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
                token_type_ids: Optional[torch.Tensor] = None, start_positions: Optional[torch.Tensor] = None,
                end_positions: Optional[torch.Tensor] = None) -> Tuple:
        B, S = input_ids.shape
        mask = None
        if attention_mask is not None:
            mask = ((1.0 - attention_mask.to(torch.float32)) * -10000.0).contiguous()
        x = self.embedding_hidden_mapping_in(self.embeddings(input_ids, token_type_ids, S))
        if x.is_cuda and x.dtype != torch.bfloat16:
            x = x.to(torch.bfloat16)
        for depth in range(self.cfg.num_hidden_layers):
            x = self.shared_layer(x, mask, B, S, seed_base=0x1000 + 16 * depth)
        return _qa_outputs(self.qa_outputs(x), B, S, start_positions, end_positions)


# ------------------------------------------------------------------------------------------------
# Hugging Face checkpoint interchange
# ------------------------------------------------------------------------------------------------
_HF_LAYER = {
    "attention.output.dense.weight": "out_weight", "attention.output.dense.bias": "out_bias",
    "attention.output.LayerNorm.weight": "ln1_weight", "attention.output.LayerNorm.bias": "ln1_bias",
    "intermediate.dense.weight": "ffn1_weight", "intermediate.dense.bias": "ffn1_bias",
    "output.dense.weight": "ffn2_weight", "output.dense.bias": "ffn2_bias",
    "output.LayerNorm.weight": "ln2_weight", "output.LayerNorm.bias": "ln2_bias",
}
_HF_TOP = {
    "bert.embeddings.word_embeddings.weight": "embeddings.word_embeddings",
    "bert.embeddings.position_embeddings.weight": "embeddings.position_embeddings",
    "bert.embeddings.token_type_embeddings.weight": "embeddings.token_type_embeddings",
    "bert.embeddings.LayerNorm.weight": "embeddings.ln_weight", "bert.embeddings.LayerNorm.bias": "embeddings.ln_bias",
    "bert.pooler.dense.weight": "head.pooler_weight", "bert.pooler.dense.bias": "head.pooler_bias",
    "classifier.weight": "head.classifier_weight", "classifier.bias": "head.classifier_bias",
    "qa_outputs.weight": "qa_outputs.weight", "qa_outputs.bias": "qa_outputs.bias",
}


def convert_hf_state_dict(hf: dict) -> dict:
    """``transformers`` BERT ``state_dict`` (``BertForSequenceClassification`` / ``BertForQuestionAnswering`` / bare
    ``BertModel``) → the parameter names of this module: per-layer query / key / value projections are concatenated into
    the packed ``qkv_weight`` the fused kernels read, everything else is renamed. The reference's trials start from
    ``from_pretrained`` weights (``examples/nlp/bert_glue_pytorch/model_def.py:60-68``); this is how such a checkpoint
    gets into the sm_100a model."""
    hf = {(k if k.startswith(("bert.", "classifier.", "qa_outputs.")) else "bert." + k): v for k, v in hf.items()}
    out = {}
    for k, name in _HF_TOP.items():
        if k in hf:
            out[name] = hf[k]
    n = 0
    while f"bert.encoder.layer.{n}.attention.self.query.weight" in hf:
        pre = f"bert.encoder.layer.{n}."
        out[f"layers.{n}.qkv_weight"] = torch.cat([hf[pre + f"attention.self.{t}.weight"] for t in ("query", "key", "value")], dim=0)
        out[f"layers.{n}.qkv_bias"] = torch.cat([hf[pre + f"attention.self.{t}.bias"] for t in ("query", "key", "value")], dim=0)
        for src, dst in _HF_LAYER.items():
            out[f"layers.{n}.{dst}"] = hf[pre + src]
        n += 1
    return out


def to_hf_state_dict(model: nn.Module) -> dict:
    """The inverse of :func:`convert_hf_state_dict`: a ``state_dict`` ``transformers`` can load."""
    sd = model.state_dict()
    out = {}
    for k, name in _HF_TOP.items():
        if name in sd:
            out[k] = sd[name]
    n = 0
    while f"layers.{n}.qkv_weight" in sd:
        pre = f"bert.encoder.layer.{n}."
        for t, w, b in zip(("query", "key", "value"), sd[f"layers.{n}.qkv_weight"].chunk(3, dim=0), sd[f"layers.{n}.qkv_bias"].chunk(3, dim=0)):
            out[pre + f"attention.self.{t}.weight"], out[pre + f"attention.self.{t}.bias"] = w.clone(), b.clone()
        for src, dst in _HF_LAYER.items():
            out[pre + src] = sd[f"layers.{n}.{dst}"]
        n += 1
    return out


def load_hf_state_dict(model: nn.Module, hf: dict) -> None:
    """Load a ``transformers`` BERT checkpoint into ``model`` (heads the checkpoint does not carry keep their init)."""
    own = set(model.state_dict())
    mine = {k: v for k, v in convert_hf_state_dict(hf).items() if k in own or not k.startswith(("head.", "qa_outputs."))}
    missing, unexpected = model.load_state_dict(mine, strict=False)
    bad = [k for k in missing if not (k.endswith("_anchor") or k.startswith(("head.", "qa_outputs.")))]
    if bad or unexpected:
        raise KeyError(f"checkpoint does not match the model: missing {bad}, unexpected {list(unexpected)}")


def config_from_hf(hf_config) -> BertConfig:
    """``transformers.BertConfig`` → :class:`BertConfig`."""
    g = lambda k, d: getattr(hf_config, k, d)  # noqa: E731
    return BertConfig(vocab_size=g("vocab_size", 30522), hidden_size=g("hidden_size", 768), num_hidden_layers=g("num_hidden_layers", 12),
                      num_attention_heads=g("num_attention_heads", 12), intermediate_size=g("intermediate_size", 3072),
                      max_position_embeddings=g("max_position_embeddings", 512), type_vocab_size=g("type_vocab_size", 2),
                      hidden_dropout_prob=g("hidden_dropout_prob", 0.1), attention_probs_dropout_prob=g("attention_probs_dropout_prob", 0.1),
                      layer_norm_eps=g("layer_norm_eps", 1e-12), initializer_range=g("initializer_range", 0.02), num_labels=g("num_labels", 2))
