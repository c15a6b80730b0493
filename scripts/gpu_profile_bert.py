"""Per-kernel time breakdown of one BERT-base optimizer step (4 micro-batches of 48x128) using the
torch profiler's CUDA activity records. Diagnostic only — numbers under a profiler are not bench values."""
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.models.bert import BertConfig, BertForSequenceClassification  # noqa: E402
from determined_b200.ops import functional as F  # noqa: E402
from determined_b200.parallel.fused_optimizer import FusedOptimizer  # noqa: E402

cfg = BertConfig()
model = BertForSequenceClassification(cfg).cuda().train()
opt = FusedOptimizer(torch.optim.AdamW(model.parameters(), lr=2e-5))
B, S = int(os.environ.get("B", 48)), 128
ids = torch.randint(0, cfg.vocab_size, (B, S), device="cuda")
am = torch.ones(B, S, dtype=torch.long, device="cuda")
tt = torch.zeros(B, S, dtype=torch.long, device="cuda")
y = torch.randint(0, 2, (B,), device="cuda")


def step():
    for _ in range(4):
        loss, _ = model(ids, am, tt, y)
        loss.backward()
    F.advance_step_seed()
    opt.step(max_grad_norm=1.0, zero_grad=True)


for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step()
    torch.cuda.synchronize()
tab = prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=90)
print(tab)
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/bert_kernel_breakdown.txt", "w").write(tab)
