"""Tiny driver for ncu: implicit-GEMM convolution (fprop / dgrad / wgrad) on a ResNet-50 layer shape, the BatchNorm
kernels on the matching activation, and the CTA-pair GEMM on a large problem."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402

N, H, W, Cin, Cout = 128, 28, 28, 128, 128
x = torch.randn(N, H, W, Cin, device="cuda").bfloat16()
w = (torch.randn(Cout, 3, 3, Cin, device="cuda") * 0.05).bfloat16()
y = torch.empty(N, H, W, Cout, device="cuda", dtype=torch.bfloat16)
dx = torch.empty_like(x)
dw = torch.zeros(Cout, 3, 3, Cin, device="cuda")
gamma, beta = torch.ones(Cout, device="cuda"), torch.zeros(Cout, device="cuda")
dg, db = torch.zeros(Cout, device="cuda"), torch.zeros(Cout, device="cuda")
A = torch.randn(8192, 4096, device="cuda").bfloat16()
B = torch.randn(8192, 4096, device="cuda").bfloat16()
C = torch.empty(8192, 8192, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    F.conv2d_implicit(F.CONV_FPROP, x, w, y, N, H, W, Cin, Cout, 3, 3, 1, 1)
    F.conv2d_implicit(F.CONV_DGRAD, y, w, dx, N, H, W, Cin, Cout, 3, 3, 1, 1)
    F.conv2d_implicit(F.CONV_WGRAD, y, x, dw, N, H, W, Cin, Cout, 3, 3, 1, 1)
    mean, rstd = F.bn_stats(y, Cout, None, None, 1e-5, 0.1)
    a = F.bn_apply(y, x, mean, rstd, gamma, beta, Cout, F.ACT_RELU)
    F.bn_bwd(dx, y, a, mean, rstd, gamma, dg, db, Cout, F.ACT_RELU, 0.0, want_dresid=True)
    F.gemm(A, B, C, 8192, 8192, 4096, force_bn=512)
torch.cuda.synchronize()
print("ok")
