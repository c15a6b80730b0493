"""DETR — end-to-end object detection with a transformer over ResNet features, authored on this repo's modules.

Counterpart of the reference's ``examples/computer_vision/detr_coco_pytorch`` (``model.py:1-282`` builds the model from the
external ``facebookresearch/detr`` checkout its startup hook clones; ``model_def.py:36-219`` is the trial). That checkout
cannot be fetched here, so the network is stated from its published description: ResNet-50 trunk with frozen batch-norm
statistics, a 1x1 projection to ``hidden_dim``, sine position embeddings added to queries and keys in every attention
layer, a post-norm encoder / decoder, ``num_queries`` learned object queries, a class head with a "no object" class and a
3-layer box head; Hungarian matching between predictions and targets, then cross entropy + L1 + generalised IoU on the
matched pairs, repeated on every decoder layer's output when ``aux_loss`` is on.

Where the work runs on a B200: the trunk is ``models/resnet.py`` (implicit-GEMM tcgen05 convolutions, fused batch-norm),
every projection / feed-forward layer is ``ops.modules.Linear`` (tcgen05 GEMM, bias in the epilogue), LayerNorm is the
in-tree kernel, and self-attention runs through ``fused_self_attention`` when the head width is 64 (``nheads =
hidden_dim / 64``: 4 heads at the default width 256). The decoder's cross-attention has different query and key lengths,
which the packed-QKV kernels do not cover; it — and any head width other than 64 — goes through torch's
``scaled_dot_product_attention``.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Tuple

import torch
from torch import nn

from ..ops.modules import LayerNorm, Linear, fused_self_attention
from .resnet import resnet18, resnet50


# ---- boxes ------------------------------------------------------------------------------------------------------------------
def box_cxcywh_to_xyxy(b: torch.Tensor) -> torch.Tensor:
    cx, cy, w, h = b.unbind(-1)
    return torch.stack([cx - 0.5 * w, cy - 0.5 * h, cx + 0.5 * w, cy + 0.5 * h], dim=-1)


def box_xyxy_to_cxcywh(b: torch.Tensor) -> torch.Tensor:
    x0, y0, x1, y1 = b.unbind(-1)
    return torch.stack([(x0 + x1) / 2, (y0 + y1) / 2, x1 - x0, y1 - y0], dim=-1)


def box_area(b: torch.Tensor) -> torch.Tensor:
    return (b[..., 2] - b[..., 0]).clamp(min=0) * (b[..., 3] - b[..., 1]).clamp(min=0)


def box_iou(a: torch.Tensor, b: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Pairwise IoU and union of ``a`` [n, 4] against ``b`` [m, 4], both xyxy."""
    lt = torch.max(a[:, None, :2], b[None, :, :2])
    rb = torch.min(a[:, None, 2:], b[None, :, 2:])
    inter = (rb - lt).clamp(min=0).prod(-1)
    union = box_area(a)[:, None] + box_area(b)[None, :] - inter
    return inter / union.clamp(min=1e-9), union


def generalized_box_iou(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Pairwise GIoU = IoU - (hull - union) / hull."""
    iou, union = box_iou(a, b)
    lt = torch.min(a[:, None, :2], b[None, :, :2])
    rb = torch.max(a[:, None, 2:], b[None, :, 2:])
    hull = (rb - lt).clamp(min=0).prod(-1)
    return iou - (hull - union) / hull.clamp(min=1e-9)


# ---- position embedding -------------------------------------------------------------------------------------------------------
def sine_position_embedding(pad_mask: torch.Tensor, num_pos_feats: int, temperature: float = 10000.0) -> torch.Tensor:
    """``pad_mask`` [B, H, W] (True on padding) → [B, H*W, 2 * num_pos_feats]: interleaved sin / cos of the normalised
    row and column index of every valid cell, as in "Attention is all you need" extended to two axes."""
    valid = (~pad_mask).float()
    y = valid.cumsum(1)
    x = valid.cumsum(2)
    eps, scale = 1e-6, 2 * math.pi
    y = y / (y[:, -1:, :] + eps) * scale
    x = x / (x[:, :, -1:] + eps) * scale
    i = torch.arange(num_pos_feats, dtype=torch.float32, device=pad_mask.device)
    dim_t = temperature ** (2 * torch.div(i, 2, rounding_mode="floor") / num_pos_feats)
    px, py = x[..., None] / dim_t, y[..., None] / dim_t
    px = torch.stack([px[..., 0::2].sin(), px[..., 1::2].cos()], dim=-1).flatten(-2)
    py = torch.stack([py[..., 0::2].sin(), py[..., 1::2].cos()], dim=-1).flatten(-2)
    return torch.cat([py, px], dim=-1).flatten(1, 2)


# ---- transformer --------------------------------------------------------------------------------------------------------------
class Attention(nn.Module):
    """Multi-head attention with separate inputs for queries, keys and values (DETR adds position embeddings to the first
    two only). Tokens are ``[B, S, hidden]``; ``key_bias`` is an additive ``[B, S_k]`` float mask (-1e4 on padding)."""

    def __init__(self, hidden: int, nheads: int, dropout: float, seed: int) -> None:
        super().__init__()
        if hidden % nheads:
            raise ValueError("hidden_dim must be a multiple of nheads")
        self.hidden, self.nheads, self.dropout, self.seed = hidden, nheads, dropout, seed
        self.q_proj, self.k_proj = Linear(hidden, hidden), Linear(hidden, hidden)
        self.v_proj, self.out_proj = Linear(hidden, hidden), Linear(hidden, hidden)
        for lin in (self.q_proj, self.k_proj, self.v_proj, self.out_proj):
            nn.init.xavier_uniform_(lin.weight)
            nn.init.zeros_(lin.bias)

    def forward(self, q_in: torch.Tensor, k_in: torch.Tensor, v_in: torch.Tensor, key_bias: Optional[torch.Tensor] = None,
                self_attention: bool = False) -> torch.Tensor:
        B, Sq, _ = q_in.shape
        Sk = k_in.shape[1]
        q, k, v = self.q_proj(q_in), self.k_proj(k_in), self.v_proj(v_in)
        p = self.dropout if self.training else 0.0
        hd = self.hidden // self.nheads
        if self_attention and q.is_cuda and hd == 64 and q.dtype == torch.bfloat16:
            qkv = torch.cat([q, k, v], dim=-1).reshape(B * Sq, 3 * self.hidden)
            o = fused_self_attention(qkv, key_bias, B, Sq, self.nheads, p, self.seed).reshape(B, Sq, self.hidden)
        else:
            q, k, v = (t.reshape(B, s, self.nheads, hd).transpose(1, 2) for t, s in ((q, Sq), (k, Sk), (v, Sk)))
            bias = None if key_bias is None else key_bias[:, None, None, :].to(q.dtype)
            o = nn.functional.scaled_dot_product_attention(q, k, v, attn_mask=bias, dropout_p=p)
            o = o.transpose(1, 2).reshape(B, Sq, self.hidden)
        return self.out_proj(o)


class _FFN(nn.Module):
    def __init__(self, hidden: int, dim_feedforward: int, dropout: float) -> None:
        super().__init__()
        self.linear1, self.linear2 = Linear(hidden, dim_feedforward), Linear(dim_feedforward, hidden)
        self.drop = nn.Dropout(dropout)
        for lin in (self.linear1, self.linear2):
            nn.init.xavier_uniform_(lin.weight)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.linear2(self.drop(torch.relu(self.linear1(x))))


class EncoderLayer(nn.Module):
    def __init__(self, hidden: int, nheads: int, dim_feedforward: int, dropout: float, pre_norm: bool, seed: int) -> None:
        super().__init__()
        self.self_attn = Attention(hidden, nheads, dropout, seed)
        self.ffn = _FFN(hidden, dim_feedforward, dropout)
        self.norm1, self.norm2 = LayerNorm(hidden), LayerNorm(hidden)
        self.drop1, self.drop2 = nn.Dropout(dropout), nn.Dropout(dropout)
        self.pre_norm = pre_norm

    def forward(self, src: torch.Tensor, pos: torch.Tensor, key_bias: Optional[torch.Tensor]) -> torch.Tensor:
        if self.pre_norm:
            x = self.norm1(src)
            src = src + self.drop1(self.self_attn(x + pos, x + pos, x, key_bias, self_attention=True))
            return src + self.drop2(self.ffn(self.norm2(src)))
        qk = src + pos
        src = self.norm1(src + self.drop1(self.self_attn(qk, qk, src, key_bias, self_attention=True)))
        return self.norm2(src + self.drop2(self.ffn(src)))


class DecoderLayer(nn.Module):
    def __init__(self, hidden: int, nheads: int, dim_feedforward: int, dropout: float, pre_norm: bool, seed: int) -> None:
        super().__init__()
        self.self_attn = Attention(hidden, nheads, dropout, seed)
        self.cross_attn = Attention(hidden, nheads, dropout, seed + 1)
        self.ffn = _FFN(hidden, dim_feedforward, dropout)
        self.norm1, self.norm2, self.norm3 = LayerNorm(hidden), LayerNorm(hidden), LayerNorm(hidden)
        self.drop1, self.drop2, self.drop3 = nn.Dropout(dropout), nn.Dropout(dropout), nn.Dropout(dropout)
        self.pre_norm = pre_norm

    def forward(self, tgt: torch.Tensor, memory: torch.Tensor, query_pos: torch.Tensor, pos: torch.Tensor,
                memory_key_bias: Optional[torch.Tensor]) -> torch.Tensor:
        if self.pre_norm:
            x = self.norm1(tgt)
            tgt = tgt + self.drop1(self.self_attn(x + query_pos, x + query_pos, x, None, self_attention=True))
            x = self.norm2(tgt)
            tgt = tgt + self.drop2(self.cross_attn(x + query_pos, memory + pos, memory, memory_key_bias))
            return tgt + self.drop3(self.ffn(self.norm3(tgt)))
        qk = tgt + query_pos
        tgt = self.norm1(tgt + self.drop1(self.self_attn(qk, qk, tgt, None, self_attention=True)))
        tgt = self.norm2(tgt + self.drop2(self.cross_attn(tgt + query_pos, memory + pos, memory, memory_key_bias)))
        return self.norm3(tgt + self.drop3(self.ffn(tgt)))


class Transformer(nn.Module):
    """Encoder over the image tokens, decoder over the object queries; returns every decoder layer's (normalised)
    output ``[L, B, Q, hidden]`` so that auxiliary losses can be attached."""

    def __init__(self, hidden: int = 256, nheads: int = 4, enc_layers: int = 6, dec_layers: int = 6,
                 dim_feedforward: int = 2048, dropout: float = 0.1, pre_norm: bool = False) -> None:
        super().__init__()
        self.encoder = nn.ModuleList(EncoderLayer(hidden, nheads, dim_feedforward, dropout, pre_norm, 0x2000 + 4 * i)
                                     for i in range(enc_layers))
        self.decoder = nn.ModuleList(DecoderLayer(hidden, nheads, dim_feedforward, dropout, pre_norm, 0x3000 + 4 * i)
                                     for i in range(dec_layers))
        self.encoder_norm = LayerNorm(hidden) if pre_norm else None
        self.decoder_norm = LayerNorm(hidden)

    def forward(self, src: torch.Tensor, key_bias: Optional[torch.Tensor], query_embed: torch.Tensor,
                pos: torch.Tensor) -> torch.Tensor:
        B = src.shape[0]
        pos = pos.to(src.dtype)
        for layer in self.encoder:
            src = layer(src, pos, key_bias)
        if self.encoder_norm is not None:
            src = self.encoder_norm(src)
        query_pos = query_embed.to(src.dtype)[None].expand(B, -1, -1)
        tgt = torch.zeros_like(query_pos)
        outs = []
        for layer in self.decoder:
            tgt = layer(tgt, src, query_pos, pos, key_bias)
            outs.append(self.decoder_norm(tgt))
        return torch.stack(outs)


# ---- network ------------------------------------------------------------------------------------------------------------------
class Backbone(nn.Module):
    """ResNet trunk up to ``layer4`` (stride 32). Batch-norm layers always use their running statistics and the stem plus
    ``layer1`` are frozen, as in the published training recipe; the remaining layers train at ``lr_backbone``."""

    def __init__(self, name: str = "resnet50") -> None:
        super().__init__()
        net = {"resnet50": resnet50, "resnet18": resnet18}[name](num_classes=8)
        del net.fc, net.avgpool
        self.body = net
        self.num_channels = 2048 if name == "resnet50" else 512
        for n, p in net.named_parameters():
            if not n.startswith(("layer2", "layer3", "layer4")) or ".bn" in n or "downsample.1" in n:
                p.requires_grad_(False)

    def train(self, mode: bool = True) -> "Backbone":
        super().train(mode)
        for m in self.body.modules():
            if m.__class__.__name__.startswith("BatchNorm"):
                m.eval()
        return self

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        from ..ops.conv import conv_bn

        b = self.body
        with torch.no_grad():
            x = b.layer1(b.maxpool(conv_bn(b.conv1, b.bn1, x)))
        return b.layer4(b.layer3(b.layer2(x)))


class MLP(nn.Module):
    def __init__(self, dims: List[int]) -> None:
        super().__init__()
        self.layers = nn.ModuleList(Linear(a, b) for a, b in zip(dims[:-1], dims[1:]))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        for i, layer in enumerate(self.layers):
            x = layer(x)
            if i + 1 < len(self.layers):
                x = torch.relu(x)
        return x


class DETR(nn.Module):
    """``forward(images [B, 3, H, W], pad_mask [B, H, W] bool)`` → ``{"pred_logits": [B, Q, classes + 1], "pred_boxes":
    [B, Q, 4] (cxcywh in [0, 1]), "aux_outputs": [...]}``."""

    def __init__(self, num_classes: int, num_queries: int = 100, hidden_dim: int = 256, nheads: int = 4, enc_layers: int = 6,
                 dec_layers: int = 6, dim_feedforward: int = 2048, dropout: float = 0.1, pre_norm: bool = False,
                 backbone: str = "resnet50", aux_loss: bool = True) -> None:
        super().__init__()
        self.backbone = Backbone(backbone)
        self.input_proj = Linear(self.backbone.num_channels, hidden_dim)      # the 1x1 convolution, on tokens
        self.transformer = Transformer(hidden_dim, nheads, enc_layers, dec_layers, dim_feedforward, dropout, pre_norm)
        self.query_embed = nn.Embedding(num_queries, hidden_dim)
        # the class head is padded to a multiple of 8 outputs so that it stays on the tensor-core GEMM; the padding
        # columns are sliced off before the loss
        self.num_logits = num_classes + 1
        self.class_embed = Linear(hidden_dim, (self.num_logits + 7) // 8 * 8)
        self.bbox_embed = MLP([hidden_dim, hidden_dim, hidden_dim, 8])
        self.hidden_dim, self.aux_loss = hidden_dim, aux_loss

    def forward(self, images: torch.Tensor, pad_mask: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        feats = self.backbone(images)                                         # [B, C, h, w]
        B, C, h, w = feats.shape
        if pad_mask is None:
            pad_mask = torch.zeros(B, images.shape[2], images.shape[3], dtype=torch.bool, device=images.device)
        m = nn.functional.interpolate(pad_mask[:, None].float(), size=(h, w)).to(torch.bool)[:, 0]
        pos = sine_position_embedding(m, self.hidden_dim // 2)
        key_bias = m.flatten(1).float() * -1e4
        tokens = self.input_proj(feats.permute(0, 2, 3, 1).reshape(B, h * w, C))
        hs = self.transformer(tokens, key_bias, self.query_embed.weight, pos)   # [L, B, Q, hidden]
        logits = self.class_embed(hs)[..., :self.num_logits].float()
        boxes = self.bbox_embed(hs)[..., :4].float().sigmoid()
        out = {"pred_logits": logits[-1], "pred_boxes": boxes[-1]}
        if self.aux_loss:
            out["aux_outputs"] = [{"pred_logits": a, "pred_boxes": b} for a, b in zip(logits[:-1], boxes[:-1])]
        return out


# ---- matching and losses ------------------------------------------------------------------------------------------------------
class HungarianMatcher(nn.Module):
    """One-to-one assignment of predictions to targets that minimises ``cost_class * (-p[class]) + cost_bbox * L1 +
    cost_giou * (-GIoU)``, solved per image with the Jonker-Volgenant solver in SciPy."""

    def __init__(self, cost_class: float = 1.0, cost_bbox: float = 5.0, cost_giou: float = 2.0) -> None:
        super().__init__()
        self.cost_class, self.cost_bbox, self.cost_giou = cost_class, cost_bbox, cost_giou

    @torch.no_grad()
    def forward(self, outputs: Dict[str, torch.Tensor], targets: List[Dict[str, torch.Tensor]]):
        from scipy.optimize import linear_sum_assignment

        B, Q, _ = outputs["pred_logits"].shape
        prob = outputs["pred_logits"].flatten(0, 1).softmax(-1)
        box = outputs["pred_boxes"].flatten(0, 1)
        ids = torch.cat([t["labels"] for t in targets])
        tbox = torch.cat([t["boxes"] for t in targets])
        if tbox.numel() == 0:
            empty = torch.empty(0, dtype=torch.int64)
            return [(empty, empty) for _ in targets]
        cost = (self.cost_bbox * torch.cdist(box, tbox, p=1) - self.cost_class * prob[:, ids]
                - self.cost_giou * generalized_box_iou(box_cxcywh_to_xyxy(box), box_cxcywh_to_xyxy(tbox)))
        cost = cost.view(B, Q, -1).cpu()                                      # ONE device → host copy for the batch
        sizes = [len(t["boxes"]) for t in targets]
        out = []
        for i, c in enumerate(cost.split(sizes, -1)):
            r, col = linear_sum_assignment(c[i].numpy())
            out.append((torch.as_tensor(r, dtype=torch.int64), torch.as_tensor(col, dtype=torch.int64)))
        return out


class SetCriterion(nn.Module):
    """Losses on the matched pairs: ``loss_ce`` (the "no object" class down-weighted by ``eos_coef``), ``loss_bbox``
    (L1 / number of boxes), ``loss_giou``, plus the logging-only ``class_error`` and ``cardinality_error``; every decoder
    layer's output gets its own matching and the same losses with an ``_{i}`` suffix."""

    def __init__(self, num_classes: int, matcher: HungarianMatcher, weight_dict: Dict[str, float], eos_coef: float = 0.1,
                 world_size: int = 1) -> None:
        super().__init__()
        self.num_classes, self.matcher, self.weight_dict, self.world_size = num_classes, matcher, weight_dict, world_size
        w = torch.ones(num_classes + 1)
        w[-1] = eos_coef
        self.register_buffer("empty_weight", w)

    def _losses(self, outputs, targets, indices, num_boxes: float, log: bool) -> Dict[str, torch.Tensor]:
        logits, boxes = outputs["pred_logits"], outputs["pred_boxes"]
        dev = logits.device
        bi = torch.cat([torch.full_like(src, i) for i, (src, _) in enumerate(indices)]).to(dev)
        si = torch.cat([src for src, _ in indices]).to(dev)
        tcls = torch.cat([t["labels"][j.to(dev)] for t, (_, j) in zip(targets, indices)])
        tbox = torch.cat([t["boxes"][j.to(dev)] for t, (_, j) in zip(targets, indices)])
        target = torch.full(logits.shape[:2], self.num_classes, dtype=torch.int64, device=dev)
        target[bi, si] = tcls
        out = {"loss_ce": nn.functional.cross_entropy(logits.transpose(1, 2), target, self.empty_weight)}
        src = boxes[bi, si]
        out["loss_bbox"] = (src - tbox).abs().sum() / num_boxes
        giou = generalized_box_iou(box_cxcywh_to_xyxy(src), box_cxcywh_to_xyxy(tbox)).diag() if len(src) else src.sum()
        out["loss_giou"] = (1 - giou).sum() / num_boxes if len(src) else giou
        if log:
            with torch.no_grad():
                hit = (logits[bi, si].argmax(-1) == tcls).float().mean() if len(src) else torch.ones((), device=dev)
                out["class_error"] = 100 - 100 * hit
                n_pred = (logits.argmax(-1) != self.num_classes).sum(1).float()
                n_true = torch.tensor([len(t["labels"]) for t in targets], dtype=torch.float32, device=dev)
                out["cardinality_error"] = (n_pred - n_true).abs().mean()
        return out

    def forward(self, outputs: Dict[str, torch.Tensor], targets: List[Dict[str, torch.Tensor]]) -> Dict[str, torch.Tensor]:
        main = {k: v for k, v in outputs.items() if k != "aux_outputs"}
        num_boxes = max(float(sum(len(t["labels"]) for t in targets)), 1.0)   # per-rank count (no cross-rank average)
        losses = self._losses(main, targets, self.matcher(main, targets), num_boxes, log=True)
        for i, aux in enumerate(outputs.get("aux_outputs", [])):
            for k, v in self._losses(aux, targets, self.matcher(aux, targets), num_boxes, log=False).items():
                losses[f"{k}_{i}"] = v
        return losses


def build_weight_dict(bbox_loss_coef: float, giou_loss_coef: float, dec_layers: int, aux_loss: bool) -> Dict[str, float]:
    wd = {"loss_ce": 1.0, "loss_bbox": float(bbox_loss_coef), "loss_giou": float(giou_loss_coef)}
    if aux_loss:
        for i in range(dec_layers - 1):
            wd.update({f"{k}_{i}": v for k, v in list(wd.items())[:3]})
    return wd


@torch.no_grad()
def post_process(outputs: Dict[str, torch.Tensor], target_sizes: torch.Tensor) -> List[Dict[str, torch.Tensor]]:
    """Per image: the best real class of every query, its probability, and the box in absolute xyxy pixels."""
    prob = outputs["pred_logits"].softmax(-1)
    scores, labels = prob[..., :-1].max(-1)
    h, w = target_sizes.unbind(1)
    scale = torch.stack([w, h, w, h], dim=1).to(outputs["pred_boxes"].dtype)
    boxes = box_cxcywh_to_xyxy(outputs["pred_boxes"]) * scale[:, None, :]
    return [{"scores": s, "labels": l, "boxes": b} for s, l, b in zip(scores, labels, boxes)]


# ---- COCO-style average precision ---------------------------------------------------------------------------------------------
def coco_average_precision(detections: Dict[int, Dict[str, "np.ndarray"]], ground_truth: Dict[int, Dict[str, "np.ndarray"]],
                           max_dets: int = 100) -> Dict[str, float]:
    """Box AP the way the COCO protocol defines it (pycocotools is not available here): per class and IoU threshold in
    0.50:0.05:0.95, detections (best ``max_dets`` per image) are matched greedily in score order to the free ground-truth
    box of highest IoU; AP is the mean of the interpolated precision at 101 recall points; classes without ground truth
    are left out. Area ranges (small < 32², medium < 96², large) ignore ground truth outside the range and unmatched
    detections outside it. Boxes are xyxy in pixels. Returns ``mAP, mAP_50, mAP_75, mAP_small, mAP_medium, mAP_large``."""
    import numpy as np

    thresholds = np.linspace(0.5, 0.95, 10)
    recall_points = np.linspace(0.0, 1.0, 101)
    ranges = {"all": (0.0, 1e10), "small": (0.0, 32.0 ** 2), "medium": (32.0 ** 2, 96.0 ** 2), "large": (96.0 ** 2, 1e10)}
    classes = sorted({int(c) for g in ground_truth.values() for c in g["labels"]})

    def area(b):
        return np.clip(b[:, 2] - b[:, 0], 0, None) * np.clip(b[:, 3] - b[:, 1], 0, None)

    def iou(a, b):
        lt = np.maximum(a[:, None, :2], b[None, :, :2])
        rb = np.minimum(a[:, None, 2:], b[None, :, 2:])
        inter = np.clip(rb - lt, 0, None).prod(-1)
        return inter / np.clip(area(a)[:, None] + area(b)[None, :] - inter, 1e-9, None)

    ap = {k: [] for k in ranges}
    ap50, ap75 = [], []
    for cls in classes:
        for rk, (lo, hi) in ranges.items():
            scores, matched, ignored = [], [], []     # per detection, per threshold
            n_gt = 0
            for img, g in ground_truth.items():
                gb = np.asarray(g["boxes"], dtype=np.float64).reshape(-1, 4)[np.asarray(g["labels"]) == cls]
                g_ign = ~((area(gb) >= lo) & (area(gb) < hi)) if len(gb) else np.zeros(0, bool)
                n_gt += int((~g_ign).sum())
                d = detections.get(img)
                if d is None:
                    continue
                keep = np.asarray(d["labels"]) == cls
                db = np.asarray(d["boxes"], dtype=np.float64).reshape(-1, 4)[keep]
                ds = np.asarray(d["scores"], dtype=np.float64)[keep]
                order = np.argsort(-ds, kind="mergesort")[:max_dets]
                db, ds = db[order], ds[order]
                if not len(db):
                    continue
                gorder = np.argsort(g_ign, kind="mergesort")          # real ground truth first, ignored last
                gb, g_ign = gb[gorder], g_ign[gorder]
                ious = iou(db, gb) if len(gb) else np.zeros((len(db), 0))
                m = np.zeros((len(thresholds), len(db)), bool)
                ig = np.zeros((len(thresholds), len(db)), bool)
                for ti, thr in enumerate(thresholds):
                    taken = np.zeros(len(gb), bool)
                    for di in range(len(db)):
                        best, best_iou = -1, min(thr, 1 - 1e-10)
                        for gi in range(len(gb)):
                            if taken[gi]:
                                continue
                            if best >= 0 and not g_ign[best] and g_ign[gi]:
                                break                                   # a real match beats any ignored one
                            if ious[di, gi] < best_iou:
                                continue
                            best, best_iou = gi, ious[di, gi]
                        if best >= 0:
                            taken[best] = True
                            m[ti, di], ig[ti, di] = True, g_ign[best]
                d_out = ~((area(db) >= lo) & (area(db) < hi))
                ig |= (~m) & d_out[None, :]
                scores.append(ds)
                matched.append(m)
                ignored.append(ig)
            if n_gt == 0:
                continue
            if scores:
                sc = np.concatenate(scores)
                order = np.argsort(-sc, kind="mergesort")
                m = np.concatenate(matched, 1)[:, order]
                ig = np.concatenate(ignored, 1)[:, order]
            else:
                m = ig = np.zeros((len(thresholds), 0), bool)
            per_thr = []
            for ti in range(len(thresholds)):
                tp = np.cumsum(m[ti] & ~ig[ti])
                fp = np.cumsum(~m[ti] & ~ig[ti])
                rc = tp / n_gt
                pr = tp / np.maximum(tp + fp, 1e-12)
                for i in range(len(pr) - 1, 0, -1):                     # precision envelope
                    pr[i - 1] = max(pr[i - 1], pr[i])
                idx = np.searchsorted(rc, recall_points, side="left")
                q = np.where(idx < len(pr), pr[np.minimum(idx, max(len(pr) - 1, 0))] if len(pr) else 0.0, 0.0)
                per_thr.append(float(np.mean(q)))
            ap[rk].append(float(np.mean(per_thr)))
            if rk == "all":
                ap50.append(per_thr[0])
                ap75.append(per_thr[5])

    def mean(v):
        return float(np.mean(v)) if v else -1.0          # COCO prints -1 when a bucket has no ground truth

    return {"mAP": mean(ap["all"]), "mAP_50": mean(ap50), "mAP_75": mean(ap75), "mAP_small": mean(ap["small"]),
            "mAP_medium": mean(ap["medium"]), "mAP_large": mean(ap["large"])}
