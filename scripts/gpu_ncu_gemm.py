"""Tiny driver for ncu: the BERT FFN1 GEMM (6144 x 3072 x 768) through the CTA-pair kernel — bare, with the
bias + GELU + aux epilogue, and the weight-gradient shape (fp32 reduce-add output) — a few launches each."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402

M, N, K = 6144, 3072, 768
A = torch.randn(M, K, device="cuda").bfloat16()
B = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
bias = torch.randn(N, device="cuda")
C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
Z = torch.empty_like(C)
G = torch.zeros(N, K, device="cuda")
for _ in range(3):
    F.gemm(A, B, C, M, N, K)
    F.gemm(A, B, C, M, N, K, epi=F.EPI_BIAS_GELU, bias=bias, aux_out=Z)
    F.gemm(C, A, G, N, K, M, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)
torch.cuda.synchronize()
print("ok")
