"""Reference arm worker: runs the UNMODIFIED reference harness (``baseline/_ref``: determined 0.13.10.dev0) through
its own public API — a ``determined.pytorch.PyTorchTrial`` driven by the reference's ``PyTorchTrialController`` over
RUN_STEP workloads — on the same model/config as ``bench.py``'s own arm.

Executed as a subprocess by ``baseline/reference_arm.py`` with ``PYTHONPATH=baseline/shims:baseline/_ref`` and a
working directory outside the repository, so nothing of determined_b200 is importable here. The model is the
stock HF ``BertForSequenceClassification`` (random-init ``bert-base`` config; the reference example uses the same
class), stock ``torch.optim.AdamW``, ``torch.nn.utils.clip_grad_norm_``; GEMMs are cuBLAS, attention is torch
SDPA. apex is not installable offline, so mixed precision is ``torch.autocast(bfloat16)`` inside ``train_batch``
(the reference's ``configure_apex_amp`` path is unavailable); ``aggregation_frequency=4`` uses the reference's own
``step_optimizer`` logic.
"""
import det_ref_compat  # noqa: F401  (must precede the reference imports)

import argparse
import json
import sys
import time

import torch

import determined as det
from determined import _execution, load, workload
from determined.pytorch import DataLoader, LRScheduler, PyTorchTrial, PyTorchTrialContext

PER_SLOT_BATCH, SEQ_LEN, AGG_FREQ = 48, 128, 4


class SyntheticGlue(torch.utils.data.Dataset):
    def __init__(self, size, seq_len, vocab, seed):
        g = torch.Generator().manual_seed(seed)
        self.ids = torch.randint(1000, vocab, (size, seq_len), generator=g, dtype=torch.int64)
        lengths = torch.randint(seq_len // 2, seq_len + 1, (size,), generator=g)
        pos = torch.arange(seq_len)[None, :]
        self.mask = (pos < lengths[:, None]).to(torch.int64)
        self.ids = self.ids * self.mask
        first = torch.randint(seq_len // 4, seq_len // 2, (size,), generator=g)
        self.tt = ((pos >= first[:, None]) & (pos < lengths[:, None])).to(torch.int64)
        self.labels = torch.randint(0, 2, (size,), generator=g, dtype=torch.int64)

    def __len__(self):
        return self.ids.shape[0]

    def __getitem__(self, i):
        return self.ids[i], self.mask[i], self.tt[i], self.labels[i]


class RefBertTrial(PyTorchTrial):
    def __init__(self, context: PyTorchTrialContext) -> None:
        from transformers import BertConfig, BertForSequenceClassification

        self.context = context
        cfg = BertConfig(num_labels=2)  # bert-base-uncased architecture, random init
        self.model = context.wrap_model(BertForSequenceClassification(cfg))
        no_decay = ["bias", "LayerNorm.weight"]
        groups = [
            {"params": [p for n, p in self.model.named_parameters() if not any(nd in n for nd in no_decay)],
             "weight_decay": 0.0},
            {"params": [p for n, p in self.model.named_parameters() if any(nd in n for nd in no_decay)],
             "weight_decay": 0.0},
        ]
        self.optimizer = context.wrap_optimizer(torch.optim.AdamW(groups, lr=2e-5, eps=1e-8))
        sched = torch.optim.lr_scheduler.LambdaLR(self.optimizer, lambda s: max(0.0, (100000 - s) / 100000.0))
        self.lr_scheduler = context.wrap_lr_scheduler(sched, LRScheduler.StepMode.STEP_EVERY_BATCH)

    def build_training_data_loader(self):
        return DataLoader(SyntheticGlue(4096, SEQ_LEN, 30522, 1), batch_size=self.context.get_per_slot_batch_size(),
                          shuffle=True, drop_last=True, pin_memory=True)

    def build_validation_data_loader(self):
        return DataLoader(SyntheticGlue(96, SEQ_LEN, 30522, 2), batch_size=self.context.get_per_slot_batch_size(),
                          drop_last=True, pin_memory=True)

    def train_batch(self, batch, epoch_idx, batch_idx):
        ids, mask, tt, labels = batch
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = self.model(input_ids=ids, attention_mask=mask, token_type_ids=tt, labels=labels)
        loss = out[0]
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer,
                                    clip_grads=lambda params: torch.nn.utils.clip_grad_norm_(params, 1.0))
        return {"loss": loss}

    def evaluate_batch(self, batch):
        ids, mask, tt, labels = batch
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = self.model(input_ids=ids, attention_mask=mask, token_type_ids=tt, labels=labels)
        return {"validation_loss": out[0], "acc": (out[1].argmax(-1) == labels).float().mean()}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    W, K = max(args.warmup, 3), args.steps
    config = {
        "hyperparameters": {"global_batch_size": PER_SLOT_BATCH},
        "searcher": {"name": "single", "metric": "acc", "smaller_is_better": False, "max_length": {"batches": 400}},
        "optimizations": {"aggregation_frequency": AGG_FREQ, "average_aggregated_gradients": True,
                          "average_training_metrics": False, "gradient_compression": False, "mixed_precision": "O0",
                          "tensor_fusion_threshold": 64, "tensor_fusion_cycle_time": 5, "auto_tune_tensor_fusion": False},
        "reproducibility": {"experiment_seed": 1234},
    }
    env, rendezvous_info, hvd_config = _execution._make_local_execution_env(
        True, config, hparams=config["hyperparameters"], limit_gpus=1)
    timing = {}

    def workloads():
        icpt = workload.WorkloadResponseInterceptor()
        total = 0
        for i in range(W):
            yield from icpt.send(workload.train_workload(i + 1, num_batches=AGG_FREQ, total_batches_processed=total), [])
            total += AGG_FREQ
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        losses = []
        for i in range(K):
            yield from icpt.send(workload.train_workload(W + i + 1, num_batches=AGG_FREQ, total_batches_processed=total), [])
            total += AGG_FREQ
            losses.append(icpt.metrics_result()["metrics"]["avg_metrics"]["loss"])
        e1.record()
        torch.cuda.synchronize()
        timing["ms"] = e0.elapsed_time(e1)
        timing["wall_ms"] = (time.perf_counter() - t0) * 1e3
        timing["loss"] = float(losses[-1])
        yield from icpt.send(workload.terminate_workload(), [])

    controller = load.load_controller_from_trial(RefBertTrial, env, workloads(), None, rendezvous_info, hvd_config)
    controller.run()
    samples = PER_SLOT_BATCH * AGG_FREQ * K
    print("REF_RESULT " + json.dumps({"ms": timing["ms"], "wall_ms": timing["wall_ms"], "samples": samples,
                                      "loss": timing["loss"], "det_version": det.__version__,
                                      "torch": torch.__version__}), flush=True)


if __name__ == "__main__":
    main()
