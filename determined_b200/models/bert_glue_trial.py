"""BERT-base GLUE-shape fine-tuning as a ``PyTorchTrial`` (the flagship benchmark trial).

Counterpart of the reference's ``examples/nlp/bert_glue_pytorch/model_def.py:27-174`` (HF
``BertForSequenceClassification`` on MRPC, seq 128, AdamW + linear warmup stepped every batch). There is no
network in this environment, so the data set is synthetic: random token ids / labels of the GLUE shape,
generated once on the host and served through the ordinary ``det.pytorch.DataLoader`` path (pinned host
memory → device copy every batch).

Hyperparameters: ``global_batch_size``, ``learning_rate``, ``adam_epsilon``, ``weight_decay``,
``num_warmup_steps``, ``num_training_steps``, ``max_seq_length``, optional ``num_hidden_layers``,
``max_grad_norm``, ``dataset_size``, ``use_cuda_graphs``.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

from .. import pytorch as dp
from .bert import BertConfig, BertForSequenceClassification


class SyntheticGlueDataset(torch.utils.data.Dataset):
    """Random (input_ids, attention_mask, token_type_ids, label) records with the MRPC tensor shapes."""

    def __init__(self, size: int, seq_len: int, vocab_size: int, num_labels: int = 2, seed: int = 0) -> None:
        g = torch.Generator().manual_seed(seed)
        self.input_ids = torch.randint(1000, vocab_size, (size, seq_len), generator=g, dtype=torch.int64)
        lengths = torch.randint(seq_len // 2, seq_len + 1, (size,), generator=g)
        pos = torch.arange(seq_len)[None, :]
        self.attention_mask = (pos < lengths[:, None]).to(torch.int64)
        self.input_ids = self.input_ids * self.attention_mask
        first = torch.randint(seq_len // 4, seq_len // 2, (size,), generator=g)
        self.token_type_ids = ((pos >= first[:, None]) & (pos < lengths[:, None])).to(torch.int64)
        self.labels = torch.randint(0, num_labels, (size,), generator=g, dtype=torch.int64)

    def __len__(self) -> int:
        return self.input_ids.shape[0]

    def __getitem__(self, i: int):
        return self.input_ids[i], self.attention_mask[i], self.token_type_ids[i], self.labels[i]


def linear_schedule_with_warmup(optimizer, num_warmup_steps: int, num_training_steps: int):
    def lr_lambda(step: int) -> float:
        if step < num_warmup_steps:
            return float(step) / float(max(1, num_warmup_steps))
        return max(0.0, float(num_training_steps - step) / float(max(1, num_training_steps - num_warmup_steps)))

    return torch.optim.lr_scheduler.LambdaLR(optimizer, lr_lambda)


class BertGlueTrial(dp.PyTorchTrial):
    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        self.context = context
        hp = context.get_hparams()
        self.seq_len = int(hp.get("max_seq_length", 128))
        cfg = BertConfig(num_hidden_layers=int(hp.get("num_hidden_layers", 12)),
                         vocab_size=int(hp.get("vocab_size", 30522)),
                         hidden_dropout_prob=float(hp.get("dropout", 0.1)),
                         attention_probs_dropout_prob=float(hp.get("dropout", 0.1)))
        self.cfg = cfg
        self.model = context.wrap_model(BertForSequenceClassification(cfg))
        no_decay = ("bias", "ln1_", "ln2_", "ln_weight", "ln_bias")
        decay = [p for n, p in self.model.named_parameters() if not any(t in n for t in no_decay)]
        nodecay = [p for n, p in self.model.named_parameters() if any(t in n for t in no_decay)]
        opt = torch.optim.AdamW([{"params": decay, "weight_decay": float(hp.get("weight_decay", 0.0))},
                                 {"params": nodecay, "weight_decay": 0.0}],
                                lr=float(hp.get("learning_rate", 2e-5)), eps=float(hp.get("adam_epsilon", 1e-8)))
        self.optimizer = context.wrap_optimizer(opt)
        self.lr_scheduler = context.wrap_lr_scheduler(
            linear_schedule_with_warmup(self.optimizer, int(hp.get("num_warmup_steps", 0)),
                                        int(hp.get("num_training_steps", 459))),
            dp.LRScheduler.StepMode.STEP_EVERY_BATCH)
        self.max_grad_norm = float(hp.get("max_grad_norm", 1.0))
        self.clip = dp.ClipGradsL2Norm(self.max_grad_norm) if self.max_grad_norm > 0 else None
        if hp.get("overlap_grad_allreduce", False):
            context.experimental.enable_grad_overlap()   # opt-in: measured slower than the step-time all-reduce, see DESIGN.md
        if hp.get("use_cuda_graphs", False):
            context.experimental.enable_cuda_graphs()

    def _dataset(self, size: int, seed: int) -> SyntheticGlueDataset:
        return SyntheticGlueDataset(size, self.seq_len, self.cfg.vocab_size, self.cfg.num_labels, seed)

    def build_training_data_loader(self) -> dp.DataLoader:
        size = int(self.context.get_hparams().get("dataset_size", 3668))  # MRPC train size
        return dp.DataLoader(self._dataset(size, 1), batch_size=self.context.get_per_slot_batch_size(), shuffle=True,
                             drop_last=True, pin_memory=torch.cuda.is_available())

    def build_validation_data_loader(self) -> dp.DataLoader:
        size = int(self.context.get_hparams().get("validation_size", 408))
        return dp.DataLoader(self._dataset(size, 2), batch_size=self.context.get_per_slot_batch_size(),
                             drop_last=True, pin_memory=torch.cuda.is_available())

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Dict[str, torch.Tensor]:
        input_ids, attention_mask, token_type_ids, labels = batch
        loss, _ = self.model(input_ids, attention_mask, token_type_ids, labels)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, clip_grads=self.clip)
        return {"loss": loss.detach()}

    def evaluate_batch(self, batch: Any) -> Dict[str, Any]:
        input_ids, attention_mask, token_type_ids, labels = batch
        loss, logits = self.model(input_ids, attention_mask, token_type_ids, labels)
        acc = (logits.argmax(dim=-1) == labels).float().mean()
        return {"validation_loss": loss, "acc": acc}
