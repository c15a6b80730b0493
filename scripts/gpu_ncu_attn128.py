"""Tiny driver for ncu: the single-tile attention kernels on the BERT-base GLUE shape (batch 48, seq 128, 12 heads)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402

B, S, H = 48, 128, 12
qkv = (torch.randn(B * S, 3 * H * 64, device="cuda") * 0.5).bfloat16()
dctx = torch.randn(B * S, H * 64, device="cuda").bfloat16()
mask = torch.zeros(B, S, device="cuda")
for _ in range(3):
    ctx, lse = F.attention_fwd(qkv, mask, B, S, H, 0.1, 3)
    F.attention_bwd(qkv, dctx, mask, lse, B, S, H, 0.1, 3)
torch.cuda.synchronize()
print("ok")
