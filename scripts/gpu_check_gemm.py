"""GPU bring-up check for the tcgen05 GEMM: every operand-major combination, tile width, epilogue and
split-K mode against an fp32 torch reference, then a timing sweep against torch.matmul (cuBLAS).
Prints one line per case and never stops at the first failure, so one gpurun call shows everything."""
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from determined_b200.ops import functional as F  # noqa: E402


def gemm(A, B, C, M, N, K, a_mn, b_mn, out_mode=0, epi=0, alpha=1.0, bias=None, aux_in=None,
         aux_out=None, dropout_p=0.0, seed=0, splitk=0, force_bn=0):
    F.gemm(A, B, C, M, N, K, a_mn=bool(a_mn), b_mn=bool(b_mn), out_mode=out_mode, epi=epi, alpha=alpha, bias=bias,
           aux_in=aux_in, aux_out=aux_out, dropout_p=dropout_p, seed=seed, splitk=splitk, force_bn=force_bn)


results = []


def report(name, err, tol):
    ok = bool(err <= tol) and math.isfinite(err)
    results.append((name, ok))
    print(f"{'PASS' if ok else 'FAIL'} {name}: max_rel_err={err:.3e} tol={tol:.1e}", flush=True)


def rel_err(out, ref):
    return ((out.float() - ref).abs().max() / ref.abs().max().clamp_min(1e-6)).item()


def main():
    torch.manual_seed(0)
    dev = "cuda"
    print(torch.cuda.get_device_name(0), flush=True)
    for (M, N, K) in [(256, 256, 128), (384, 512, 320), (1000, 776, 200)]:
        for a_mn in (0, 1):
            for b_mn in (0, 1):
                for bn in (64, 128, 256, 512):   # 512 = CTA-pair kernel (256 x 256 over two SMs)
                    A = torch.randn(M, K, device=dev).bfloat16()
                    B = torch.randn(N, K, device=dev).bfloat16()
                    ref = A.float() @ B.float().t()
                    As = A.t().contiguous() if a_mn else A
                    Bs = B.t().contiguous() if b_mn else B
                    C = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
                    try:
                        gemm(As, Bs, C, M, N, K, a_mn, b_mn, force_bn=bn)
                        torch.cuda.synchronize()
                        report(f"gemm M{M} N{N} K{K} a_mn{a_mn} b_mn{b_mn} bn{bn}", rel_err(C, ref), 1.5e-2)
                    except Exception as e:  # noqa
                        print("ERROR", M, N, K, a_mn, b_mn, bn, repr(e), flush=True)
                        results.append((f"gemm {M} {N} {K} {a_mn} {b_mn} {bn}", False))
                        if "launch failure" in repr(e) or "illegal" in repr(e):
                            summarize()
                            return
    # epilogues
    M, N, K = 512, 768, 256
    A = torch.randn(M, K, device=dev).bfloat16()
    B = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
    bias = torch.randn(N, device=dev)
    acc = A.float() @ B.float().t()
    C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    gemm(A, B, C, M, N, K, 0, 0, epi=1, bias=bias)
    report("epi bias", rel_err(C, acc + bias), 1.5e-2)
    C.zero_()
    gemm(A, B, C, M, N, K, 0, 0, epi=1, bias=bias, force_bn=512)
    report("epi bias (CTA pair)", rel_err(C, acc + bias), 1.5e-2)
    Zp = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    gemm(A, B, C, M, N, K, 0, 0, epi=2, bias=bias, aux_out=Zp, force_bn=512)
    report("epi bias_gelu C (CTA pair)", rel_err(C, torch.nn.functional.gelu(acc + bias, approximate="tanh")), 1.5e-2)
    report("epi bias_gelu Z (CTA pair)", rel_err(Zp, acc + bias), 1.5e-2)
    Z = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    gemm(A, B, C, M, N, K, 0, 0, epi=2, bias=bias, aux_out=Z)
    report("epi bias_gelu C", rel_err(C, torch.nn.functional.gelu(acc + bias, approximate="tanh")), 1.5e-2)
    report("epi bias_gelu Z", rel_err(Z, acc + bias), 1.5e-2)
    zf = Z.float().requires_grad_(True)
    torch.nn.functional.gelu(zf, approximate="tanh").sum().backward()
    gemm(A, B, C, M, N, K, 0, 0, epi=3, aux_in=Z)
    report("epi dgelu", rel_err(C, acc * zf.grad), 2e-2)
    R = torch.randn(M, N, device=dev).bfloat16()
    gemm(A, B, C, M, N, K, 0, 0, epi=4, bias=bias, aux_in=R)
    report("epi bias_resid", rel_err(C, acc + bias + R.float()), 1.5e-2)
    gemm(A, B, C, M, N, K, 0, 0, epi=4, bias=bias, aux_in=R, dropout_p=0.1, seed=1234)
    d = C.float() - R.float()
    kept = (d.abs() > 1e-6)
    frac = kept.float().mean().item()
    ref_d = (acc + bias) / 0.9
    err = ((d - ref_d).abs() * kept).max().item() / ref_d.abs().max().item()
    print(f"dropout keep fraction {frac:.4f} (expect ~0.9)")
    report("epi bias_resid dropout (kept values)", err, 2e-2)
    results.append(("dropout fraction", abs(frac - 0.9) < 0.01))
    # fp32 out + atomic split-K accumulate
    Cf = torch.empty(M, N, device=dev, dtype=torch.float32)
    gemm(A, B, Cf, M, N, K, 0, 0, out_mode=1)
    report("out f32", rel_err(Cf, acc), 1e-3)
    M2, N2, K2 = 768, 768, 6144
    dY = torch.randn(K2, M2, device=dev).bfloat16()   # [tokens, out]  -> A MN-major
    X = torch.randn(K2, N2, device=dev).bfloat16()    # [tokens, in]   -> B MN-major
    ref = dY.float().t() @ X.float()
    G = torch.ones(M2, N2, device=dev, dtype=torch.float32)
    gemm(dY, X, G, M2, N2, K2, 1, 1, out_mode=2)
    report("wgrad splitk auto (+1 preexisting)", rel_err(G - 1.0, ref), 1e-3)
    G.zero_()
    gemm(dY, X, G, M2, N2, K2, 1, 1, out_mode=2, splitk=1)
    report("wgrad splitk=1 atomic", rel_err(G, ref), 1e-3)
    G.zero_()
    gemm(dY, X, G, M2, N2, K2, 1, 1, out_mode=2, force_bn=512)
    report("wgrad splitk auto (CTA pair)", rel_err(G, ref), 1e-3)
    # a long persistent run: many tiles per cluster, both TMEM buffers and every smem stage wrap several times
    Ml, Nl, Kl = 6144, 3072, 768
    Al = torch.randn(Ml, Kl, device=dev).bfloat16()
    Bl = (torch.randn(Nl, Kl, device=dev) * 0.05).bfloat16()
    Cl = torch.empty(Ml, Nl, device=dev, dtype=torch.bfloat16)
    gemm(Al, Bl, Cl, Ml, Nl, Kl, 0, 0, force_bn=512)
    report("CTA pair 6144x3072x768", rel_err(Cl, Al.float() @ Bl.float().t()), 1.5e-2)

    # timing
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

    timing = []
    # the big square problem runs LAST: it pulls the board into its power cap and everything timed right after it reads
    # 1.5-1.7x slow (seen as "none" = 52 us vs 30 us for the same FFN1 GEMM within one run)
    def fwd_row(M, N, K):
        A = torch.randn(M, K, device=dev).bfloat16()
        B = torch.randn(N, K, device=dev).bfloat16()
        C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        row = {"M": M, "N": N, "K": K}
        for bn in (0, 128, 256, 512):
            ms = bench(lambda: gemm(A, B, C, M, N, K, 0, 0, force_bn=bn))
            row[f"det_bn{bn}_ms"] = round(ms, 4)
            row[f"det_bn{bn}_tflops"] = round(2 * M * N * K / ms / 1e9, 1)
        ms = bench(lambda: torch.matmul(A, B.t(), out=C))
        row["cublas_ms"] = round(ms, 4)
        row["cublas_tflops"] = round(2 * M * N * K / ms / 1e9, 1)
        timing.append(row)
        print(json.dumps(row), flush=True)

    for shape in [(6144, 768, 768), (6144, 2304, 768), (6144, 3072, 768), (6144, 768, 3072)]:
        fwd_row(*shape)
    # what the fused epilogues cost on the BERT shapes (same GEMM, different epilogue)
    for (name, M, N, K) in [("qkv", 6144, 2304, 768), ("out_proj", 6144, 768, 768), ("ffn1", 6144, 3072, 768), ("ffn2", 6144, 768, 3072)]:
        A = torch.randn(M, K, device=dev).bfloat16()
        B = (torch.randn(N, K, device=dev) * 0.05).bfloat16()
        bias = torch.randn(N, device=dev)
        R = torch.randn(M, N, device=dev).bfloat16()
        C = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        Z = torch.empty_like(C)
        row = {"epilogue_cost": name, "M": M, "N": N, "K": K}
        row["none_us"] = round(bench(lambda: gemm(A, B, C, M, N, K, 0, 0)) * 1e3, 1)
        row["bias_us"] = round(bench(lambda: gemm(A, B, C, M, N, K, 0, 0, epi=1, bias=bias)) * 1e3, 1)
        row["bias_gelu_aux_us"] = round(bench(lambda: gemm(A, B, C, M, N, K, 0, 0, epi=2, bias=bias, aux_out=Z)) * 1e3, 1)
        row["bias_resid_us"] = round(bench(lambda: gemm(A, B, C, M, N, K, 0, 0, epi=4, bias=bias, aux_in=R)) * 1e3, 1)
        row["bias_resid_dropout_us"] = round(bench(lambda: gemm(A, B, C, M, N, K, 0, 0, epi=4, bias=bias, aux_in=R, dropout_p=0.1, seed=7)) * 1e3, 1)
        row["dgelu_us"] = round(bench(lambda: gemm(A, B, C, M, N, K, 0, 0, epi=3, aux_in=Z)) * 1e3, 1)
        timing.append(row)
        print(json.dumps(row), flush=True)
    # dgrad timing (B MN-major)
    for (M, N, K) in [(6144, 768, 2304), (6144, 768, 3072), (6144, 3072, 768)]:
        dY = torch.randn(M, K, device=dev).bfloat16()
        Wt = torch.randn(K, N, device=dev).bfloat16()
        dX = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        row = {"dgrad": True, "M": M, "N": N, "K": K}
        for bn in (0, 512):
            ms = bench(lambda: gemm(dY, Wt, dX, M, N, K, 0, 1, force_bn=bn))
            row[f"det_bn{bn}_ms"] = round(ms, 4)
            row[f"det_bn{bn}_tflops"] = round(2 * M * N * K / ms / 1e9, 1)
        ms = bench(lambda: torch.matmul(dY, Wt, out=dX))
        row["cublas_ms"] = round(ms, 4)
        row["cublas_tflops"] = round(2 * M * N * K / ms / 1e9, 1)
        timing.append(row)
        print(json.dumps(row), flush=True)
    # wgrad timing
    for (M, N, K) in [(768, 768, 6144), (3072, 768, 6144), (768, 3072, 6144)]:
        dY = torch.randn(K, M, device=dev).bfloat16()
        X = torch.randn(K, N, device=dev).bfloat16()
        G = torch.zeros(M, N, device=dev, dtype=torch.float32)
        ms = bench(lambda: gemm(dY, X, G, M, N, K, 1, 1, out_mode=2))
        Gb = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        ms2 = bench(lambda: torch.matmul(dY.t(), X, out=Gb))
        msp = bench(lambda: gemm(dY, X, G, M, N, K, 1, 1, out_mode=2, force_bn=512))
        row = {"wgrad": True, "M": M, "N": N, "K": K, "det_ms": round(ms, 4), "det_tflops": round(2 * M * N * K / ms / 1e9, 1),
               "det_pair_ms": round(msp, 4), "det_pair_tflops": round(2 * M * N * K / msp / 1e9, 1),
               "cublas_ms": round(ms2, 4), "cublas_tflops": round(2 * M * N * K / ms2 / 1e9, 1)}
        timing.append(row)
        print(json.dumps(row), flush=True)
    fwd_row(8192, 8192, 8192)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "gemm_timing.json"), "w") as f:
        json.dump(timing, f, indent=1)
    summarize()


def summarize():
    bad = [n for n, ok in results if not ok]
    print(f"SUMMARY: {len(results) - len(bad)}/{len(results)} passed")
    for n in bad:
        print("  FAILED:", n)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
