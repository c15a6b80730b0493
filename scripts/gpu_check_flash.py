"""Time the tiled attention kernels on SQuAD shapes against torch SDPA (library flash attention) on the same tensors,
and a BERT-base SQuAD-shape training step through the trial controller. Writes gpurun_out/flash_timing.json."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402


def timeit(fn, iters=20, warm=5):
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


def main():
    out = []
    for B, S, H in [(32, 384, 12), (8, 384, 64), (8, 1024, 12), (4, 2048, 16), (48, 128, 12)]:
        hid = H * 64
        qkv = (torch.randn(B * S, 3 * hid, device="cuda") * 0.5).bfloat16()
        dctx = torch.randn(B * S, hid, device="cuda").bfloat16()
        ctx, lse = F.attention_fwd(qkv, None, B, S, H, 0.0, 0)
        t_f = timeit(lambda: F.attention_fwd(qkv, None, B, S, H, 0.0, 0))
        t_b = timeit(lambda: F.attention_bwd(qkv, dctx, None, lse, B, S, H, 0.0, 0, ctx))
        q, k, v = [t.view(B, S, H, 64).transpose(1, 2).contiguous().requires_grad_(True) for t in qkv.split(hid, dim=1)]
        t_sf = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v))
        o = torch.nn.functional.scaled_dot_product_attention(q, k, v)
        do = dctx.view(B, S, H, 64).transpose(1, 2).contiguous()
        t_sb = timeit(lambda: torch.autograd.grad(o, (q, k, v), do, retain_graph=True))
        fl = 4.0 * B * H * S * S * 64
        row = {"B": B, "S": S, "H": H, "fwd_ms": round(t_f, 4), "bwd_ms": round(t_b, 4), "fwd_tflops": round(fl / t_f / 1e9, 1),
               "bwd_tflops": round(2.5 * fl / t_b / 1e9, 1), "sdpa_fwd_ms": round(t_sf, 4), "sdpa_bwd_ms": round(t_sb, 4)}
        print(row, flush=True)
        out.append(row)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/flash_timing.json", "w"), indent=1)


if __name__ == "__main__":
    main()
