"""GPU check of the convolutional stack against fp32 torch references (forward, input gradient, weight/bias gradients),
then a timing sweep of ResNet-50 convolution shapes against cuDNN (bf16, channels_last). Never stops at the first
failure; prints a summary and writes gpurun_out/conv_timing.json."""
import json
import math
import os
import sys

import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from determined_b200.ops import conv as C  # noqa: E402
from determined_b200.ops import functional as F  # noqa: E402

results = []
dev = "cuda"


def report(name, err, tol):
    ok = math.isfinite(err) and err <= tol
    results.append((name, ok))
    print(f"{'PASS' if ok else 'FAIL'} {name}: rel_err={err:.3e} tol={tol:.1e}", flush=True)


def rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


def bf(t):
    return t.to(torch.bfloat16).float()


def check_module(name, mine, ref, x, tol=2e-2, extra_inputs=(), gtol=None):
    """mine/ref: modules with identical state_dicts; x fp32 NCHW (values bf16-representable)."""
    gtol = gtol or tol
    try:
        ref = ref.to(dev).float()
        mine = mine.to(dev)
        mine.load_state_dict(ref.state_dict())
        xr = x.clone().requires_grad_(True)
        xm = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        yr = ref(xr, *extra_inputs)
        ym = mine(xm, *[e.to(torch.bfloat16) for e in extra_inputs])
        report(name + " fwd", rel(ym, yr), tol)
        g = bf(torch.randn_like(yr))
        yr.backward(g)
        ym.backward(g.to(torch.bfloat16))
        torch.cuda.synchronize()
        report(name + " dx", rel(xm.grad, xr.grad), gtol)
        sd_r = dict(ref.named_parameters())
        msd = {}
        for k, p in mine.named_parameters():
            gk = p.grad
            if gk is not None and gk.dim() == 4:
                gk = gk.permute(0, 3, 1, 2)
            msd[k] = gk
        for k, p in sd_r.items():
            if p.grad is not None and msd.get(k) is not None:
                report(f"{name} d{k}", rel(msd[k], p.grad), gtol)
    except Exception as e:  # noqa
        print("ERROR", name, repr(e)[:400], flush=True)
        results.append((name, False))
        if "illegal" in repr(e) or "launch failure" in repr(e) or "unspecified" in repr(e):
            summarize()


def conv_case(N, H, W, Cin, Cout, k, stride, pad, bias=True, relu=False):
    torch.manual_seed(N * 7 + H + Cin + Cout + k)
    x = bf(torch.randn(N, Cin, H, W, device=dev))
    mine = C.Conv2d(Cin, Cout, k, stride, pad, bias=bias, activation="relu" if relu else None)
    core = nn.Conv2d(Cin, Cout, k, stride, pad, bias=bias)
    with torch.no_grad():
        core.weight.copy_(bf(core.weight))
    ref = core
    if relu:  # same parameter names as the fused module
        class R(nn.Module):
            def __init__(self):
                super().__init__()
                self.weight, self.bias = core.weight, core.bias

            def forward(self, x):
                return torch.relu(nn.functional.conv2d(x, self.weight, self.bias, stride, pad))
        ref = R()
    check_module(f"conv N{N} {H}x{W} {Cin}->{Cout} k{k} s{stride} p{pad}{' relu' if relu else ''}", mine, ref, x)


def convt_case(N, H, W, Cin, Cout, k, stride, pad):
    torch.manual_seed(N + H + Cin + Cout)
    x = bf(torch.randn(N, Cin, H, W, device=dev))
    mine = C.ConvTranspose2d(Cin, Cout, k, stride, pad)
    ref = nn.ConvTranspose2d(Cin, Cout, k, stride, pad)
    with torch.no_grad():
        ref.weight.copy_(bf(ref.weight))
    check_module(f"convT N{N} {H}x{W} {Cin}->{Cout} k{k} s{stride} p{pad}", mine, ref, x)


class RefBN(nn.Module):
    def __init__(self, Cn, act, slope):
        super().__init__()
        self.bn = nn.BatchNorm2d(Cn)
        self.act, self.slope = act, slope

    def forward(self, x, r=None):
        y = self.bn(x)
        if r is not None:
            y = y + r
        if self.act == "relu":
            y = torch.relu(y)
        elif self.act == "leaky_relu":
            y = nn.functional.leaky_relu(y, self.slope)
        return y

    def state_dict(self, *a, **k):
        return self.bn.state_dict(*a, **k)

    def named_parameters(self, *a, **k):
        return self.bn.named_parameters(*a, **k)


def bn_case(N, Cn, H, W, act, resid, slope=0.2):
    torch.manual_seed(Cn + H)
    x = bf(torch.randn(N, Cn, H, W, device=dev) * 2 + 0.5)
    ref = RefBN(Cn, act, slope)
    with torch.no_grad():
        ref.bn.weight.copy_(torch.rand(Cn) + 0.5)
        ref.bn.bias.copy_(torch.randn(Cn) * 0.2)
    mine = C.BatchNorm2d(Cn, activation=act, negative_slope=slope)
    name = f"bn N{N} C{Cn} {H}x{W} act={act} resid={resid}"
    try:
        ref = ref.to(dev)
        mine = mine.to(dev)
        mine.load_state_dict(ref.state_dict())
        xr = x.clone().requires_grad_(True)
        xm = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        rr = rm = None
        if resid:
            r0 = bf(torch.randn_like(x))
            rr = r0.clone().requires_grad_(True)
            rm = r0.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        yr = ref(xr, rr)
        ym = mine(xm, rm)
        report(name + " fwd", rel(ym, yr), 2e-2)
        report(name + " running_mean", rel(mine.running_mean, ref.bn.running_mean), 1e-2)
        report(name + " running_var", rel(mine.running_var, ref.bn.running_var), 1e-2)
        g = bf(torch.randn_like(yr))
        yr.backward(g)
        ym.backward(g.to(torch.bfloat16))
        report(name + " dx", rel(xm.grad, xr.grad), 3e-2)
        if resid:
            report(name + " dresid", rel(rm.grad, rr.grad), 2e-2)
        report(name + " dgamma", rel(mine.weight.grad, ref.bn.weight.grad), 2e-2)
        report(name + " dbeta", rel(mine.bias.grad, ref.bn.bias.grad), 2e-2)
        mine.eval(); ref.eval()
        ye = mine(xm.detach(), rm.detach() if resid else None)
        report(name + " eval fwd", rel(ye, ref(xr.detach(), rr.detach() if resid else None)), 2e-2)
    except Exception as e:  # noqa
        print("ERROR", name, repr(e)[:400], flush=True)
        results.append((name, False))


def misc_cases():
    torch.manual_seed(3)
    for (k, s, p, H) in [(3, 2, 1, 17), (2, 2, 0, 12), (3, 2, 1, 112)]:
        x = bf(torch.randn(3, 16, H, H, device=dev))
        check_module(f"maxpool k{k} s{s} p{p} H{H}", C.MaxPool2d(k, s, p), nn.MaxPool2d(k, s, p), x, tol=1e-6, gtol=1e-2)
    x = bf(torch.randn(5, 64, 7, 7, device=dev))

    class RefGAP(nn.Module):
        def forward(self, x):
            return x.mean(dim=(2, 3))
    check_module("global avgpool", C.GlobalAvgPool2d(), RefGAP(), x, tol=1e-2)
    for kind, refm in [("relu", nn.ReLU()), ("leaky_relu", nn.LeakyReLU(0.2)), ("tanh", nn.Tanh()), ("sigmoid", nn.Sigmoid())]:
        check_module(f"act {kind}", C.Activation(kind, 0.2), refm, bf(torch.randn(2, 8, 5, 5, device=dev)), tol=1e-2)
    d = C.Dropout(0.25).to(dev).train()
    xx = torch.ones(4, 64, 16, 16, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = d(xx)
    frac = (y.float() > 0).float().mean().item()
    print(f"dropout keep fraction {frac:.4f} (expect 0.75)")
    results.append(("dropout fraction", abs(frac - 0.75) < 0.01))
    y.backward(torch.ones_like(y))
    results.append(("dropout bwd mask matches fwd", bool(torch.equal(xx.grad.float() > 0, y.float() > 0))))
    xi = torch.randn(3, 3, 9, 11, device=dev)
    report("nchw->nhwc", rel(F.nchw_to_nhwc_bf16(xi).permute(0, 3, 1, 2), bf(xi)), 1e-6)


def timing():
    def bench(fn, iters=10):
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

    rows = []
    N = 128
    for (H, Cin, Cout, k, s, p) in [(56, 64, 64, 3, 1, 1), (28, 128, 128, 3, 1, 1), (14, 256, 256, 3, 1, 1), (7, 512, 512, 3, 1, 1),
                                    (56, 64, 256, 1, 1, 0), (56, 256, 64, 1, 1, 0), (28, 512, 128, 1, 1, 0), (14, 1024, 256, 1, 1, 0),
                                    (56, 128, 128, 3, 2, 1), (224, 3, 64, 7, 2, 3)]:
        try:
            mine = C.Conv2d(Cin, Cout, k, s, p, bias=False).to(dev)
            ref = nn.Conv2d(Cin, Cout, k, s, p, bias=False).to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
            x = torch.randn(N, Cin, H, H, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
            need_dx = Cin > 3
            x.requires_grad_(need_dx)
            flops = 2.0 * N * (C.F.conv_out_size(H, k, s, p) ** 2) * Cin * Cout * k * k
            row = {"N": N, "H": H, "Cin": Cin, "Cout": Cout, "k": k, "s": s}
            with torch.no_grad():
                t = bench(lambda: mine(x)); row["det_fwd_ms"] = round(t, 4); row["det_fwd_tflops"] = round(flops / t / 1e9, 1)
                t = bench(lambda: ref(x)); row["cudnn_fwd_ms"] = round(t, 4); row["cudnn_fwd_tflops"] = round(flops / t / 1e9, 1)
            ym = mine(x); g = torch.randn_like(ym)
            t = bench(lambda: torch.autograd.grad(ym, [x, mine.weight] if need_dx else [mine.weight], g, retain_graph=True, allow_unused=True))
            row["det_bwd_ms"] = round(t, 4)
            yr = ref(x)
            t = bench(lambda: torch.autograd.grad(yr, [x, ref.weight] if need_dx else [ref.weight], g, retain_graph=True))
            row["cudnn_bwd_ms"] = round(t, 4)
            rows.append(row)
            print(json.dumps(row), flush=True)
        except Exception as e:  # noqa
            print("TIMING ERROR", H, Cin, Cout, k, s, repr(e)[:300], flush=True)
    # BatchNorm + ReLU bandwidth
    for (H, Cn) in [(56, 256), (28, 512), (112, 64)]:
        x = torch.randn(N, Cn, H, H, device=dev, dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
        m = C.BatchNorm2d(Cn, activation="relu").to(dev)
        r = nn.BatchNorm2d(Cn).to(dev)
        with torch.no_grad():
            t = bench(lambda: m(x)); t2 = bench(lambda: torch.relu(r(x)))
        gb = x.numel() * 2 / 1e9
        row = {"bn_relu": True, "H": H, "C": Cn, "det_ms": round(t, 4), "det_GBps_3pass": round(3 * gb / t * 1e3, 1), "torch_ms": round(t2, 4)}
        rows.append(row)
        print(json.dumps(row), flush=True)
    # the four BatchNorm kernels one by one (effective GB/s = bytes each must move / time)
    for (H, Cn) in [(112, 64), (56, 64), (56, 256), (28, 512), (14, 1024), (7, 2048)]:
        M = N * H * H
        xx = torch.randn(M, Cn, device=dev, dtype=torch.bfloat16)
        dy = torch.randn_like(xx)
        rs_ = torch.randn_like(xx)
        gamma = torch.ones(Cn, device=dev); beta = torch.zeros(Cn, device=dev)
        mean, rstd = F.bn_stats(xx, Cn, None, None, 1e-5, 0.1)
        yy = F.bn_apply(xx, rs_, mean, rstd, gamma, beta, Cn, F.ACT_RELU)
        dg = torch.zeros(Cn, device=dev); db = torch.zeros(Cn, device=dev)
        gb = M * Cn * 2 / 1e9
        t_stats = bench(lambda: F.bn_stats(xx, Cn, None, None, 1e-5, 0.1))
        t_apply = bench(lambda: F.bn_apply(xx, None, mean, rstd, gamma, beta, Cn, F.ACT_RELU))
        t_apply_r = bench(lambda: F.bn_apply(xx, rs_, mean, rstd, gamma, beta, Cn, F.ACT_RELU))
        t_bwd = bench(lambda: F.bn_bwd(dy, xx, yy, mean, rstd, gamma, dg, db, Cn, F.ACT_RELU, 0.0, want_dresid=False))
        t_bwd_r = bench(lambda: F.bn_bwd(dy, xx, yy, mean, rstd, gamma, dg, db, Cn, F.ACT_RELU, 0.0, want_dresid=True))
        row = {"bn_kernels": True, "H": H, "C": Cn, "MB": round(gb * 1e3, 1),
               "stats_GBps": round(gb / t_stats * 1e3), "apply_GBps": round(2 * gb / t_apply * 1e3),
               "apply_resid_GBps": round(3 * gb / t_apply_r * 1e3), "bwd_GBps": round(7 * gb / t_bwd * 1e3),
               "bwd_resid_GBps": round(8 * gb / t_bwd_r * 1e3),
               "ms": [round(t, 4) for t in (t_stats, t_apply, t_apply_r, t_bwd, t_bwd_r)]}
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "conv_timing.json"), "w") as f:
        json.dump(rows, f, indent=1)


def summarize():
    bad = [n for n, ok in results if not ok]
    print(f"SUMMARY: {len(results) - len(bad)}/{len(results)} passed")
    for n in bad:
        print("  FAILED:", n)
    sys.exit(1 if bad else 0)


def main():
    print(torch.cuda.get_device_name(0), flush=True)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    # implicit GEMM (stride 1)
    conv_case(4, 56, 56, 64, 64, 3, 1, 1)
    conv_case(2, 28, 28, 128, 128, 3, 1, 1, bias=False)
    conv_case(3, 14, 14, 256, 256, 3, 1, 1, bias=False)
    conv_case(5, 7, 7, 512, 512, 3, 1, 1, bias=False)
    conv_case(2, 10, 13, 32, 48, 3, 1, 1, relu=True)
    conv_case(1, 20, 20, 16, 24, 5, 1, 2)
    conv_case(2, 12, 12, 32, 64, 3, 1, 0, relu=True)
    conv_case(1, 8, 200, 16, 16, 3, 1, 1)
    conv_case(3, 9, 9, 72, 136, 3, 1, 1)
    conv_case(2, 6, 6, 320, 40, 3, 1, 1)
    # plain GEMM (1x1)
    conv_case(4, 14, 14, 256, 1024, 1, 1, 0, bias=False)
    # gather path
    conv_case(2, 28, 28, 128, 128, 3, 2, 1, bias=False)
    conv_case(2, 64, 64, 3, 64, 7, 2, 3, bias=False)
    conv_case(2, 28, 28, 256, 512, 1, 2, 0, bias=False)
    conv_case(4, 28, 28, 1, 32, 3, 1, 0, relu=True)
    conv_case(4, 4, 4, 512, 1, 4, 1, 0, bias=False)
    conv_case(2, 32, 32, 3, 32, 3, 1, 1, relu=True)
    convt_case(4, 1, 1, 100, 512, 4, 1, 0)
    convt_case(2, 8, 8, 256, 128, 4, 2, 1)
    convt_case(2, 16, 16, 64, 3, 4, 2, 1)
    bn_case(4, 64, 14, 14, "relu", True)
    bn_case(3, 256, 7, 9, None, False)
    bn_case(2, 128, 8, 8, "leaky_relu", False)
    bn_case(2, 2048, 7, 7, "relu", True)
    bn_case(8, 24, 5, 5, "relu", False)
    misc_cases()
    if "--no-timing" not in sys.argv:
        timing()
    summarize()


if __name__ == "__main__":
    main()
