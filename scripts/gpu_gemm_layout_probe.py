"""Which part of the weight-gradient GEMM costs more than cuBLAS?  Times the CTA-pair kernel on one square problem for every
operand layout (K-major / MN-major A and B), then the BERT weight-gradient shapes with bf16 output vs fp32 reduce-add output,
static split-K vs stream-K — each against cuBLAS on the same tensors."""
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
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


dev = "cuda"
rows = []
M = N = K = 4096
for a_mn in (False, True):
    for b_mn in (False, True):
        A = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
        B = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
        C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        us = bench(lambda: F.gemm(A, B, C, M, N, K, a_mn=a_mn, b_mn=b_mn, force_bn=512))
        Am = A.t() if a_mn else A
        Bm = B.t() if b_mn else B
        cu = bench(lambda: torch.matmul(Am, Bm.t(), out=C))
        rows.append({"square": M, "a_mn": a_mn, "b_mn": b_mn, "pair_us": round(us, 1), "pair_tflops": round(2 * M * N * K / us / 1e6, 1),
                     "cublas_us": round(cu, 1)})
        print(json.dumps(rows[-1]), flush=True)

T = 6144
for (Mw, Nw) in ((768, 768), (2304, 768), (3072, 768), (768, 3072)):
    dY = torch.randn(T, Mw, device=dev).bfloat16()
    X = torch.randn(T, Nw, device=dev).bfloat16()
    G = torch.zeros(Mw, Nw, device=dev)
    Gb = torch.empty(Mw, Nw, device=dev, dtype=torch.bfloat16)
    r = {"wgrad": [Mw, Nw, T]}
    r["f32_add_streamk_us"] = round(bench(lambda: F.gemm(dY, X, G, Mw, Nw, T, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD, splitk=-1, force_bn=512)), 1)
    r["f32_add_auto_us"] = round(bench(lambda: F.gemm(dY, X, G, Mw, Nw, T, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)), 1)
    for sk in (1, 2, 3, 4):
        r[f"f32_add_splitk{sk}_us"] = round(bench(lambda: F.gemm(dY, X, G, Mw, Nw, T, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD,
                                                                  splitk=sk, force_bn=512)), 1)
    r["bf16_out_us"] = round(bench(lambda: F.gemm(dY, X, Gb, Mw, Nw, T, a_mn=True, b_mn=True, force_bn=512)), 1)
    r["bf16_out_bn256_us"] = round(bench(lambda: F.gemm(dY, X, Gb, Mw, Nw, T, a_mn=True, b_mn=True, force_bn=256)), 1)
    r["cublas_bf16_us"] = round(bench(lambda: torch.matmul(dY.t(), X, out=Gb)), 1)
    r["ideal_us_at_cublas_8192_rate"] = round(2 * Mw * Nw * T / 1533e6, 1)
    rows.append(r)
    print(json.dumps(r), flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/gemm_layout_probe.json", "w"), indent=1)
