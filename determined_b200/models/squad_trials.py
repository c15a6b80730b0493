"""Extractive question answering (SQuAD shape, seq 384) as ``PyTorchTrial``s: BERT and ALBERT.

Counterparts of the reference's ``examples/nlp/bert_squad_pytorch/model_def.py`` (HF ``BertForQuestionAnswering``,
AdamW + linear warmup stepped every batch, ``max_seq_length: 384``) and ``examples/nlp/albert_squad_pytorch/model_def.py``
(``albert-xxlarge-v2``, ``aggregation_frequency: 24``). There is no network here, so the records are synthetic: random
token ids with a question segment, a context segment and an answer span inside the context, in the tensor layout
``squad_convert_examples_to_features`` produces (``input_ids, attention_mask, token_type_ids, start_positions,
end_positions``). Validation reports the span metrics of ``squad_evaluate`` (exact match and token-overlap F1 of the
arg-max span) computed on the device.

Hyperparameters: ``global_batch_size``, ``learning_rate``, ``adam_epsilon``, ``weight_decay``, ``num_warmup_steps``,
``num_training_steps``, ``max_seq_length``, ``max_query_length``, ``max_answer_length``, ``max_grad_norm``; optional
model-size overrides ``num_hidden_layers``, ``hidden_size``, ``num_attention_heads``, ``intermediate_size``,
``embedding_size`` (ALBERT), ``vocab_size``, ``dropout``; ``dataset_size``, ``validation_size``, ``use_cuda_graphs``.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from .. import pytorch as dp
from .bert import AlbertConfig, AlbertForQuestionAnswering, BertConfig, BertForQuestionAnswering
from .bert_glue_trial import linear_schedule_with_warmup


class SyntheticSquadDataset(torch.utils.data.Dataset):
    """[CLS] question [SEP] context [SEP] padding, answer span inside the context."""

    def __init__(self, size: int, seq_len: int, vocab_size: int, max_query_length: int = 64, max_answer_length: int = 30,
                 seed: int = 0) -> None:
        g = torch.Generator().manual_seed(seed)
        pos = torch.arange(seq_len)[None, :]
        qlen = torch.randint(8, max(9, min(max_query_length, seq_len // 3)) + 1, (size,), generator=g)
        total = torch.randint(seq_len // 2, seq_len + 1, (size,), generator=g)
        total = torch.maximum(total, qlen + 8)
        self.attention_mask = (pos < total[:, None]).to(torch.int64)
        self.token_type_ids = ((pos > qlen[:, None]) & (pos < total[:, None])).to(torch.int64)
        self.input_ids = torch.randint(1000, vocab_size, (size, seq_len), generator=g, dtype=torch.int64) * self.attention_mask
        ctx_lo = qlen + 2
        ctx_len = (total - 1 - ctx_lo).clamp_min(1)
        self.start_positions = ctx_lo + (torch.rand(size, generator=g) * ctx_len).long().clamp_max(seq_len - 1)
        alen = torch.randint(1, max_answer_length + 1, (size,), generator=g)
        self.end_positions = torch.minimum(self.start_positions + alen - 1, total - 2).clamp_min(0)
        self.end_positions = torch.maximum(self.end_positions, self.start_positions)
        # make the task learnable: the answer span is marked by a reserved token id range
        span = (pos >= self.start_positions[:, None]) & (pos <= self.end_positions[:, None])
        self.input_ids = torch.where(span, self.input_ids % 900 + 50, self.input_ids)

    def __len__(self) -> int:
        return self.input_ids.shape[0]

    def __getitem__(self, i: int):
        return (self.input_ids[i], self.attention_mask[i], self.token_type_ids[i], self.start_positions[i],
                self.end_positions[i])


def span_metrics(start_logits: torch.Tensor, end_logits: torch.Tensor, start: torch.Tensor, end: torch.Tensor) -> Dict[str, torch.Tensor]:
    """Exact match and token-overlap F1 of the arg-max span, the two numbers ``squad_evaluate`` reports."""
    ps, pe = start_logits.argmax(-1), end_logits.argmax(-1)
    pe = torch.maximum(pe, ps)
    inter = (torch.minimum(pe, end) - torch.maximum(ps, start) + 1).clamp_min(0).float()
    plen, glen = (pe - ps + 1).float(), (end - start + 1).float()
    prec, rec = inter / plen, inter / glen
    f1 = torch.where(inter > 0, 2 * prec * rec / (prec + rec).clamp_min(1e-9), torch.zeros_like(prec))
    return {"exact": ((ps == start) & (pe == end)).float().mean(), "f1": f1.mean()}


class _SquadTrialBase(dp.PyTorchTrial):
    def _build_model(self, hp: Dict[str, Any]) -> torch.nn.Module:
        raise NotImplementedError

    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.seq_len = int(hp.get("max_seq_length", 384))
        self.model = context.wrap_model(self._build_model(hp))
        self.cfg = self.model.cfg
        no_decay = ("bias", "ln1_", "ln2_", "ln_weight", "ln_bias")
        named = list(self.model.named_parameters())
        decay = [p for n, p in named if not any(t in n for t in no_decay)]
        nodecay = [p for n, p in named if any(t in n for t in no_decay)]
        opt = torch.optim.AdamW([{"params": decay, "weight_decay": float(hp.get("weight_decay", 0.0))},
                                 {"params": nodecay, "weight_decay": 0.0}],
                                lr=float(hp.get("learning_rate", 3e-5)), eps=float(hp.get("adam_epsilon", 1e-8)))
        self.optimizer = context.wrap_optimizer(opt)
        self.lr_scheduler = context.wrap_lr_scheduler(
            linear_schedule_with_warmup(self.optimizer, int(hp.get("num_warmup_steps", 0)),
                                        int(hp.get("num_training_steps", 15000))),
            dp.LRScheduler.StepMode.STEP_EVERY_BATCH)
        g = float(hp.get("max_grad_norm", 1.0))
        self.clip = dp.ClipGradsL2Norm(g) if g > 0 else None
        if hp.get("use_cuda_graphs", False):
            context.experimental.enable_cuda_graphs()

    def _dataset(self, size: int, seed: int) -> SyntheticSquadDataset:
        hp = self.context.get_hparams()
        return SyntheticSquadDataset(size, self.seq_len, self.cfg.vocab_size, int(hp.get("max_query_length", 64)),
                                     int(hp.get("max_answer_length", 30)), seed)

    def build_training_data_loader(self) -> dp.DataLoader:
        size = int(self.context.get_hparams().get("dataset_size", 4096))
        return dp.DataLoader(self._dataset(size, 1), batch_size=self.context.get_per_slot_batch_size(), shuffle=True,
                             drop_last=True, pin_memory=torch.cuda.is_available())

    def build_validation_data_loader(self) -> dp.DataLoader:
        size = int(self.context.get_hparams().get("validation_size", 512))
        return dp.DataLoader(self._dataset(size, 2), batch_size=self.context.get_per_slot_batch_size(), drop_last=True,
                             pin_memory=torch.cuda.is_available())

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        input_ids, attention_mask, token_type_ids, start, end = batch
        loss = self.model(input_ids, attention_mask, token_type_ids, start, end)[0]
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self.clip)
        return {"loss": loss.detach()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        input_ids, attention_mask, token_type_ids, start, end = batch
        loss, sl, el = self.model(input_ids, attention_mask, token_type_ids, start, end)
        return {"validation_loss": loss, **span_metrics(sl.float(), el.float(), start, end)}


def _size_overrides(hp: Dict[str, Any], keys) -> Dict[str, Any]:
    out = {k: int(hp[k]) for k in keys if k in hp}
    if "dropout" in hp:
        out["hidden_dropout_prob"] = out["attention_probs_dropout_prob"] = float(hp["dropout"])
    return out


class BertSquadTrial(_SquadTrialBase):
    def _build_model(self, hp: Dict[str, Any]) -> torch.nn.Module:
        keys = ("num_hidden_layers", "hidden_size", "num_attention_heads", "intermediate_size", "vocab_size")
        return BertForQuestionAnswering(BertConfig(**_size_overrides(hp, keys)))


class AlbertSquadTrial(_SquadTrialBase):
    def _build_model(self, hp: Dict[str, Any]) -> torch.nn.Module:
        keys = ("num_hidden_layers", "hidden_size", "num_attention_heads", "intermediate_size", "vocab_size", "embedding_size")
        return AlbertForQuestionAnswering(AlbertConfig(**_size_overrides(hp, keys)))
