"""Tiny driver for ncu: the tiled attention kernels on the BERT-base SQuAD shape (batch 32, seq 384, 12 heads)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402

B, S, H = 32, 384, 12
qkv = (torch.randn(B * S, 3 * H * 64, device="cuda") * 0.5).bfloat16()
dctx = torch.randn(B * S, H * 64, device="cuda").bfloat16()
for _ in range(3):
    ctx, lse = F.attention_fwd(qkv, None, B, S, H, 0.1, 3)
    F.attention_bwd(qkv, dctx, None, lse, B, S, H, 0.1, 3, ctx)
torch.cuda.synchronize()
print("ok")
