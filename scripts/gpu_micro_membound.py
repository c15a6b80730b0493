"""Isolated timing (CUDA events) of the memory-bound kernels against the measured HBM copy bandwidth."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402

PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) else 6650.0


def bench(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


out = []
n = 109_000_000 // 4 * 4
p, g, s1, s2 = [torch.randn(n, device="cuda") for _ in range(4)]
s2.abs_()
sh = torch.empty(n, device="cuda", dtype=torch.bfloat16)
hp = torch.tensor([1e-5, 0.9, 0.999, 1e-8, 0.0, 0.1, 0.001, 0, 0, 0, 0, 0], device="cuda")
gs = torch.ones(1, device="cuda")
ms = bench(lambda: F.fused_optimizer_step(F.OPT_ADAMW, p, g, s1, s2, sh, hp, gs, None, True), 10)
out.append({"kernel": "fused_adamw 109M", "ms": round(ms, 4), "bytes": n * 34, "gbs": round(n * 34 / ms / 1e6, 1)})
ms = bench(lambda: F.fused_optimizer_step(F.OPT_ADAMW, p, g, s1, s2, None, hp, gs, None, False), 10)
out.append({"kernel": "fused_adamw no-shadow no-zero", "ms": round(ms, 4), "bytes": n * 28, "gbs": round(n * 28 / ms / 1e6, 1)})
a = torch.randn(n, device="cuda")
b = torch.empty_like(a)
ms = bench(lambda: b.copy_(a), 10)
out.append({"kernel": "torch copy 436MB", "ms": round(ms, 4), "bytes": n * 8, "gbs": round(n * 8 / ms / 1e6, 1)})
T, H = 6144, 768
x = torch.randn(T, H, device="cuda").bfloat16()
gm, bt = torch.ones(H, device="cuda"), torch.zeros(H, device="cuda")
ms = bench(lambda: F.layernorm_fwd(x, gm, bt, 1e-12))
out.append({"kernel": "ln_fwd 6144x768", "ms": round(ms, 4), "bytes": T * H * 4, "gbs": round(T * H * 4 / ms / 1e6, 1)})
y, mean, rstd = F.layernorm_fwd(x, gm, bt, 1e-12)
dy = torch.randn_like(x)
dg, db = torch.zeros(H, device="cuda"), torch.zeros(H, device="cuda")
ms = bench(lambda: F.layernorm_bwd(dy, x, gm, mean, rstd, dg, db, want_masked=True, out_dropout_p=0.1, out_seed=3))
out.append({"kernel": "ln_bwd+mask 6144x768", "ms": round(ms, 4), "bytes": T * H * 8, "gbs": round(T * H * 8 / ms / 1e6, 1)})
z = torch.randn(T, 3072, device="cuda").bfloat16()
acc = torch.zeros(3072, device="cuda")
ms = bench(lambda: F.colsum_accum(z, acc))
out.append({"kernel": "colsum 6144x3072", "ms": round(ms, 4), "bytes": T * 3072 * 2, "gbs": round(T * 3072 * 2 / ms / 1e6, 1)})
B, S, Hh = 48, 128, 12
qkv = (torch.randn(B * S, 3 * 768, device="cuda") * 0.5).bfloat16()
ms = bench(lambda: F.attention_fwd(qkv, None, B, S, Hh, 0.1, 1))
out.append({"kernel": "attn_fwd B48", "ms": round(ms, 4), "bytes": B * S * 768 * 2 * 4, "gbs": round(B * S * 768 * 8 / ms / 1e6, 1), "tflops": round(4 * B * Hh * S * S * 64 / ms / 1e9, 1)})
ctx, lse = F.attention_fwd(qkv, None, B, S, Hh, 0.1, 1)
dctx = torch.randn_like(ctx)
ms = bench(lambda: F.attention_bwd(qkv, dctx, None, lse, B, S, Hh, 0.1, 1))
out.append({"kernel": "attn_bwd B48", "ms": round(ms, 4), "bytes": B * S * 768 * 2 * 7, "gbs": round(B * S * 768 * 14 / ms / 1e6, 1), "tflops": round(10 * B * Hh * S * S * 64 / ms / 1e9, 1)})
for o in out:
    o["frac_of_measured_hbm"] = round(o["gbs"] / PEAK, 3)
    print(json.dumps(o), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/micro_membound.json", "w"), indent=1)
