"""Times a handful of GEMM shapes through one kernel choice (default: the CTA-pair kernel) — the A/B harness used with
DET_KERNELS_LIB=<variant .so> when a kernel parameter (stage count, epilogue variant) is being decided."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402

bn = int(os.environ.get("BN", "512"))


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


out = {"lib": os.environ.get("DET_KERNELS_LIB", "default"), "bn": bn}
for (M, N, K) in [(6144, 768, 768), (6144, 2304, 768), (6144, 3072, 768), (6144, 768, 3072), (4096, 4096, 4096)]:
    A = torch.randn(M, K, device="cuda").bfloat16()
    B = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    bias = torch.randn(N, device="cuda")
    C = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    Z = torch.empty_like(C)
    R = torch.randn(M, N, device="cuda").bfloat16()
    row = {"none": bench(lambda: F.gemm(A, B, C, M, N, K, force_bn=bn))}
    if os.environ.get("EPIS", "1") == "1":
        row["gelu_aux"] = bench(lambda: F.gemm(A, B, C, M, N, K, epi=F.EPI_BIAS_GELU, bias=bias, aux_out=Z, force_bn=bn))
        row["resid_drop"] = bench(lambda: F.gemm(A, B, C, M, N, K, epi=F.EPI_BIAS_RESID, bias=bias, aux_in=R, dropout_p=0.1, seed=3, force_bn=bn))
        row["dgelu"] = bench(lambda: F.gemm(A, B, C, M, N, K, epi=F.EPI_DGELU, aux_in=Z, force_bn=bn))
    row["cublas"] = bench(lambda: torch.matmul(A, B.t(), out=C))
    out[f"{M}x{N}x{K}"] = {k: round(v, 1) for k, v in row.items()}
for (M, N, K) in [(768, 768, 6144), (2304, 768, 6144), (3072, 768, 6144), (768, 3072, 6144)]:
    dY = torch.randn(K, M, device="cuda").bfloat16()
    X = torch.randn(K, N, device="cuda").bfloat16()
    G = torch.zeros(M, N, device="cuda", dtype=torch.float32)
    Gb = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    row = {}
    for sk in (0, -1, 1, 2, 4, 8):
        try:
            row[f"splitk{sk}"] = bench(lambda: F.gemm(dY, X, G, M, N, K, a_mn=True, b_mn=True, out_mode=F.OUT_F32_ATOMIC_ADD, force_bn=bn, splitk=sk))
        except Exception as e:  # noqa
            row[f"splitk{sk}"] = -1.0
    row["cublas"] = bench(lambda: torch.matmul(dY.t(), X, out=Gb))
    out[f"wgrad {M}x{N}x{K}"] = {k: round(v, 1) for k, v in row.items()}
print(json.dumps(out), flush=True)
