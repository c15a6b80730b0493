"""A/B of the one-shot and the persistent single-tile attention backward (DET_ATTN_ONESHOT=1 selects the former):
numerics against each other and timing on the BERT-base GLUE shape."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402


def timeit(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


out = []
for B, S, H, p in [(48, 128, 12, 0.1), (48, 128, 12, 0.0), (64, 100, 16, 0.1), (13, 128, 12, 0.1)]:
    torch.manual_seed(B)
    qkv = (torch.randn(B * S, 3 * H * 64, device="cuda") * 0.5).bfloat16()
    dctx = torch.randn(B * S, H * 64, device="cuda").bfloat16()
    mask = torch.zeros(B, S, device="cuda")
    mask[:, S - 20:] = -10000.0
    ctx, lse = F.attention_fwd(qkv, mask, B, S, H, p, 3)
    d = F.attention_bwd(qkv, dctx, mask, lse, B, S, H, p, 3)
    t_b = timeit(lambda: F.attention_bwd(qkv, dctx, mask, lse, B, S, H, p, 3))
    t_f = timeit(lambda: F.attention_fwd(qkv, mask, B, S, H, p, 3))
    row = {"B": B, "S": S, "H": H, "dropout": p, "oneshot": os.environ.get("DET_ATTN_ONESHOT", "0"), "fwd_us": round(t_f * 1e3, 2),
           "bwd_us": round(t_b * 1e3, 2), "checksum": float(d.float().abs().sum()), "finite": bool(torch.isfinite(d.float()).all())}
    torch.save(torch.cat([d.float().flatten(), ctx.float().flatten(), lse.flatten()]).cpu(), f"gpurun_out/attn128_{B}_{S}_{H}_{p}_{row['oneshot']}.pt")
    print(json.dumps(row), flush=True)
    out.append(row)
