"""ResNet-50 training step (fwd + bwd + fused SGD) on one GPU: eager timing of ours vs torchvision (bf16 autocast,
channels_last, cuDNN) and a per-kernel breakdown of ours from the torch profiler. Diagnostic only."""
import json
import os
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.models.resnet import resnet50  # noqa: E402
from determined_b200.ops.modules import cross_entropy  # noqa: E402
from determined_b200.parallel.fused_optimizer import FusedOptimizer  # noqa: E402

B = int(os.environ.get("B", 128))
x = torch.randn(B, 3, 224, 224, device="cuda")
y = torch.randint(0, 1000, (B,), device="cuda")
out = {"batch": B}


def bench(fn, iters=8):
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


model = resnet50().cuda().train()
opt = FusedOptimizer(torch.optim.SGD(model.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4))
losses = []


def step():
    loss = cross_entropy(model(x), y)
    loss.backward()
    opt.step(zero_grad=True)
    losses.append(loss.detach())


ms = bench(step)
out["det_eager_ms"] = round(ms, 3)
out["det_eager_img_s"] = round(B / ms * 1e3, 1)
print("loss trajectory (same batch):", [round(float(l), 4) for l in losses[:3]], "...", round(float(losses[-1]), 4), flush=True)
print(json.dumps(out), flush=True)

# CUDA-graph replay of the same step
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
import gc
gc.collect()
try:
    with torch.cuda.stream(s):
        step()
        torch.cuda.synchronize()
        opt.host_prepare()
        with torch.cuda.graph(g, stream=s):
            loss = cross_entropy(model(x), y)
            loss.backward()
            opt.device_step(zero_grad=True)
    torch.cuda.current_stream().wait_stream(s)

    def replay():
        opt.host_prepare()
        g.replay()
    ms = bench(replay)
    out["det_graph_ms"] = round(ms, 3)
    out["det_graph_img_s"] = round(B / ms * 1e3, 1)
    print(json.dumps(out), flush=True)
except Exception as e:  # noqa
    print("graph capture failed:", repr(e)[:500], flush=True)

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step()
    torch.cuda.synchronize()
tab = prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=100)
print(tab)
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/resnet_kernel_breakdown.txt", "w").write(tab)

del model, opt, g
torch.cuda.empty_cache()
try:
    import torchvision

    ref = torchvision.models.resnet50().cuda().train().to(memory_format=torch.channels_last)
    ropt = torch.optim.SGD(ref.parameters(), lr=0.01, momentum=0.9, weight_decay=1e-4)
    xc = x.contiguous(memory_format=torch.channels_last)

    def ref_step_amp():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = torch.nn.functional.cross_entropy(ref(xc), y)
        loss.backward()
        ropt.step()
        ropt.zero_grad(set_to_none=True)

    def ref_step_fp32():
        loss = torch.nn.functional.cross_entropy(ref(x), y)
        loss.backward()
        ropt.step()
        ropt.zero_grad(set_to_none=True)

    ms = bench(ref_step_amp)
    out["torchvision_bf16_cl_ms"] = round(ms, 3)
    out["torchvision_bf16_cl_img_s"] = round(B / ms * 1e3, 1)
    ref = ref.to(memory_format=torch.contiguous_format)
    ms = bench(ref_step_fp32)
    out["torchvision_fp32_ms"] = round(ms, 3)
    out["torchvision_fp32_img_s"] = round(B / ms * 1e3, 1)
except Exception as e:  # noqa
    print("torchvision comparison failed:", repr(e)[:300])
print(json.dumps(out), flush=True)
json.dump(out, open("gpurun_out/resnet_timing.json", "w"), indent=1)
