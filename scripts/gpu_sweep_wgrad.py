"""Tile-width / split-K sweep of the weight-gradient GEMMs of one BERT-base layer (dW = dY^T X, K = 6144 tokens, fp32
atomic accumulate): shows what the launcher's automatic choice leaves on the table. Prints one JSON line per shape."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402


def bench(fn, iters=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


K = 6144
for M, N in [(768, 768), (2304, 768), (3072, 768), (768, 3072)]:
    dY = torch.randn(K, M, device="cuda").bfloat16()
    X = torch.randn(K, N, device="cuda").bfloat16()
    G = torch.zeros(M, N, device="cuda")
    row = {"M": M, "N": N, "K": K, "auto_us": round(bench(lambda: F.gemm(dY, X, G, M, N, K, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)), 2)}
    best = None
    for bn in (64, 128, 256):
        for sk in (1, 2, 3, 4, 6, 8, 12, 16):
            try:
                us = bench(lambda: F.gemm(dY, X, G, M, N, K, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD, splitk=sk, force_bn=bn), iters=15)
            except Exception as e:      # noqa: BLE001
                continue
            row[f"bn{bn}_sk{sk}"] = round(us, 2)
            if best is None or us < best[0]:
                best = (us, bn, sk)
    Gb = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    row["cublas_us"] = round(bench(lambda: torch.matmul(dY.t(), X, out=Gb)), 2)
    row["best"] = {"us": round(best[0], 2), "bn": best[1], "splitk": best[2]}
    print(json.dumps(row), flush=True)
