"""Tiny driver for ncu: the kernels of the step tail and the embedding path at BERT-base size — fused optimizer update
(110 M fp32 parameters: clip coefficient, AdamW, zero-grad, bf16 shadow refresh), the gradient sum-of-squares / finiteness
scan, LayerNorm forward/backward, the embedding scatter-add — plus the FFN GEMM with the GELU epilogue, a few launches each."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402
from determined_b200.ops import modules as M  # noqa: E402
from determined_b200.parallel.fused_optimizer import FusedOptimizer  # noqa: E402

torch.manual_seed(0)
dev = torch.device("cuda")
n = 110_000_000
p = torch.nn.Parameter(torch.randn(n, device=dev) * 0.02)
opt = FusedOptimizer(torch.optim.AdamW([p], lr=1e-4, weight_decay=0.01))
assert opt.fused, "fused optimizer path is not active"
for _ in range(3):
    opt.grad_buckets()[0].normal_()
    opt.step(max_grad_norm=1.0, zero_grad=True)
torch.cuda.synchronize()
opt.grad_buckets()[0].normal_()
torch.cuda.synchronize()
torch.cuda.profiler.start()          # ncu --profile-from-start off: only the launches between start() and stop() are captured
opt.step(max_grad_norm=1.0, zero_grad=True)
torch.cuda.synchronize()
torch.cuda.profiler.stop()

ln = M.LayerNorm(768).to(dev)
x = torch.randn(6144, 768, device=dev, dtype=torch.bfloat16, requires_grad=True)
for _ in range(3):
    y = ln(x)
    y.backward(torch.ones_like(y))
torch.cuda.synchronize()
go = torch.ones_like(y)
torch.cuda.profiler.start()
y = ln(x)
y.backward(go)
torch.cuda.synchronize()
torch.cuda.profiler.stop()

Mm, N, K = 6144, 3072, 768
A = torch.randn(Mm, K, device=dev).bfloat16()
B = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
bias = torch.randn(N, device=dev)
C = torch.empty(Mm, N, device=dev, dtype=torch.bfloat16)
Z = torch.empty_like(C)
G = torch.zeros(N, K, device=dev)
for _ in range(3):
    F.gemm(A, B, C, Mm, N, K, epi=F.EPI_BIAS_GELU, bias=bias, aux_out=Z)
    F.gemm(A, B, C, Mm, N, K)
    F.gemm(C, A, G, N, K, Mm, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)
torch.cuda.synchronize()
torch.cuda.profiler.start()
F.gemm(A, B, C, Mm, N, K, epi=F.EPI_BIAS_GELU, bias=bias, aux_out=Z)
F.gemm(A, B, C, Mm, N, K)
F.gemm(C, A, G, N, K, Mm, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
