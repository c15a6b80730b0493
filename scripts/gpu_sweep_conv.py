"""Tile-width / split-K sweep of the implicit-GEMM convolution on the four 3x3 layers of ResNet-50 (batch 128): what the
launcher's automatic choices leave on the table. One JSON line per (layer, mode)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402


def bench(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


N = 128
for H, C in [(56, 64), (28, 128), (14, 256), (7, 512)]:
    x = torch.randn(N, H, H, C, device="cuda").bfloat16()
    w = (torch.randn(C, 3, 3, C, device="cuda") * 0.05).bfloat16()
    y = torch.empty(N, H, H, C, device="cuda", dtype=torch.bfloat16)
    dw = torch.zeros(C, 3, 3, C, device="cuda")
    for mode, name, args in ((F.CONV_FPROP, "fprop", (x, w, y)), (F.CONV_DGRAD, "dgrad", (y, w, x)), (F.CONV_WGRAD, "wgrad", (y, x, dw))):
        row = {"H": H, "C": C, "mode": name, "auto_us": round(bench(lambda: F.conv2d_implicit(mode, *args, N, H, H, C, C, 3, 3, 1, 1)), 2)}
        best = None
        for bn in (64, 128, 256):
            if bn > max(C, 64):
                continue
            for sk in ((0,) if mode != F.CONV_WGRAD else (1, 2, 4, 8, 12, 16, 24, 32, 48, 64)):
                try:
                    us = bench(lambda: F.conv2d_implicit(mode, *args, N, H, H, C, C, 3, 3, 1, 1, splitk=sk, force_bn=bn), iters=10)
                except Exception:      # noqa: BLE001
                    continue
                row[f"bn{bn}_sk{sk}"] = round(us, 2)
                if best is None or us < best[0]:
                    best = (us, bn, sk)
        row["best"] = {"us": round(best[0], 2), "bn": best[1], "splitk": best[2]}
        print(json.dumps(row), flush=True)
