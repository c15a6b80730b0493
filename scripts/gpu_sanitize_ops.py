"""Small-shape pass over the hand-written kernels: every family once, shapes small enough for a memory checker's 10-100x
slow-down to stay in seconds (compute-sanitizer is closed on the GPU pool this was developed on, so the pass has only been
run bare: it exits 0; kernels trap on their own bounded waits). Checks nothing numerically — tests/test_gpu_kernels.py does."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402
from determined_b200.ops import modules as M  # noqa: E402
from determined_b200.parallel.fused_optimizer import FusedOptimizer  # noqa: E402

dev = "cuda"
torch.manual_seed(0)
# GEMM: single-CTA kernel, CTA-pair kernel (bf16 out, GELU+aux epilogue, fp32 reduce-add with stream-K), K tail
for (m, n, k, bn) in ((256, 256, 192, 0), (512, 512, 256, 512), (384, 320, 200, 0)):
    a = torch.randn(m, k, device=dev).bfloat16()
    b = torch.randn(n, k, device=dev).bfloat16()
    c = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
    F.gemm(a, b, c, m, n, k, force_bn=bn)
    if k % 8 == 0 and n % 8 == 0:
        z = torch.empty_like(c)
        F.gemm(a, b, c, m, n, k, epi=F.EPI_BIAS_GELU, bias=torch.randn(n, device=dev), aux_out=z, force_bn=bn)
dy = torch.randn(512, 256, device=dev).bfloat16()
x = torch.randn(512, 512, device=dev).bfloat16()
g = torch.zeros(256, 512, device=dev)
F.gemm(dy, x, g, 256, 512, 512, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD, force_bn=512)
F.gemm(dy, x, g, 256, 512, 512, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD, splitk=-1, force_bn=512)
# attention (single tile) forward / backward, both forward forms
B, S, H = 2, 128, 2
qkv = (torch.randn(B * S, 3 * H * 64, device=dev) * 0.5).bfloat16()
ctx, lse = F.attention_fwd(qkv, None, B, S, H, 0.1, 3)
F.attention_bwd(qkv, torch.randn_like(ctx), None, lse, B, S, H, 0.1, 3)
# tiled attention
B, S, H = 1, 256, 2
qkv = (torch.randn(B * S, 3 * H * 64, device=dev) * 0.5).bfloat16()
ctx, lse = F.attention_fwd(qkv, None, B, S, H, 0.0, 0)
F.attention_bwd(qkv, torch.randn_like(ctx), None, lse, B, S, H, 0.0, 0, out=ctx)
# LayerNorm forward / backward through the module, cross entropy, embedding
ln = M.LayerNorm(768).to(dev)
xx = torch.randn(64, 768, device=dev, dtype=torch.bfloat16, requires_grad=True)
ln(xx).float().sum().backward()
logits = torch.randn(32, 10, device=dev, requires_grad=True)
M.cross_entropy(logits, torch.randint(0, 10, (32,), device=dev)).backward()
# fused optimizer + grad norm
p = torch.nn.Parameter(torch.randn(10000, device=dev))
opt = FusedOptimizer(torch.optim.AdamW([p], lr=1e-3))
opt.grad_buckets()[0].normal_()
opt.step(max_grad_norm=1.0, zero_grad=True)
torch.cuda.synchronize()
print("ok")
