"""GPU check of the fused BERT path: (1) loss and gradients of a 2-layer model against the fp32 CPU
statement of the same math, (2) BERT-base step timing, eager and CUDA-graph, with FusedOptimizer."""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.models.bert import BertConfig, BertForSequenceClassification  # noqa: E402
from determined_b200.ops import _native  # noqa: E402
from determined_b200.ops import functional as F  # noqa: E402
from determined_b200.parallel.fused_optimizer import FusedOptimizer  # noqa: E402


def numerics():
    torch.manual_seed(0)
    cfg = BertConfig(num_hidden_layers=2, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, vocab_size=1000)
    ref = BertForSequenceClassification(cfg)
    gpu = BertForSequenceClassification(cfg)
    gpu.load_state_dict(ref.state_dict())
    gpu.cuda()
    B, S = 4, 128
    ids = torch.randint(0, 1000, (B, S))
    am = torch.ones(B, S, dtype=torch.long)
    am[:, 100:] = 0
    tt = torch.randint(0, 2, (B, S))
    y = torch.randint(0, 2, (B,))
    ref.train()
    gpu.train()
    loss_r, logit_r = ref(ids, am, tt, y)
    loss_r.backward()
    loss_g, logit_g = gpu(ids.cuda(), am.cuda(), tt.cuda(), y.cuda())
    loss_g.backward()
    torch.cuda.synchronize()
    print(f"loss ref {loss_r.item():.5f} gpu {loss_g.item():.5f}")
    ok = abs(loss_r.item() - loss_g.item()) < 2e-2
    worst = 0.0
    for (n, pr), (_, pg) in zip(ref.named_parameters(), gpu.named_parameters()):
        if n.endswith("_anchor"):
            continue
        if pg.grad is None:
            print("NO GRAD", n)
            ok = False
            continue
        gr, gg = pr.grad, pg.grad.cpu()
        denom = gr.abs().max().clamp_min(1e-6)
        err = ((gr - gg).abs().max() / denom).item()
        cos = torch.nn.functional.cosine_similarity(gr.flatten(), gg.flatten(), dim=0).item()
        worst = max(worst, err)
        flag = "ok" if (err < 6e-2 and cos > 0.995) else "BAD"
        if flag == "BAD":
            ok = False
        print(f"  {flag} {n:40s} max_rel_err={err:.3e} cos={cos:.5f}")
    print("NUMERICS", "PASS" if ok else "FAIL", f"worst={worst:.3e}")
    return ok


def timing():
    torch.manual_seed(0)
    cfg = BertConfig()
    model = BertForSequenceClassification(cfg).cuda().train()
    opt = FusedOptimizer(torch.optim.AdamW(model.parameters(), lr=2e-5, eps=1e-8, weight_decay=0.0))
    print("fused optimizer:", opt.fused, "params", sum(p.numel() for p in model.parameters()) / 1e6, "M")
    B, S = 48, 128
    ids = torch.randint(0, cfg.vocab_size, (B, S), device="cuda")
    am = torch.ones(B, S, dtype=torch.long, device="cuda")
    tt = torch.zeros(B, S, dtype=torch.long, device="cuda")
    y = torch.randint(0, 2, (B,), device="cuda")

    def micro():
        loss, _ = model(ids, am, tt, y)
        loss.backward()
        return loss.detach()

    def full_step(agg=4):
        for _ in range(agg):
            loss = micro()
        F.advance_step_seed()
        opt.step(max_grad_norm=1.0, zero_grad=True)
        return loss

    for _ in range(3):
        full_step()
    torch.cuda.synchronize()
    n0 = _native.launch_count
    full_step()
    launches = _native.launch_count - n0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 5
    e0.record()
    for _ in range(iters):
        loss = full_step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    out = {"mode": "eager", "ms_per_opt_step": round(ms, 3), "samples_per_s": round(4 * B / ms * 1e3, 1),
           "loss": float(loss), "det_launches_per_opt_step": launches}
    print(json.dumps(out), flush=True)

    # CUDA graph of the whole optimizer step (4 micro-batches + clip + update)
    del loss
    import gc
    gc.collect()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        full_step()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    opt.host_prepare()
    with torch.cuda.graph(g):
        for _ in range(4):
            static_loss = micro()
        F.advance_step_seed()
        opt.device_step(max_grad_norm=1.0, zero_grad=True)
    for _ in range(3):
        opt.host_prepare()
        g.replay()
    torch.cuda.synchronize()
    iters = 10
    e0.record()
    for _ in range(iters):
        opt.host_prepare()
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    flops = model.flops_per_sample(S) * 4 * B
    out2 = {"mode": "cuda_graph", "ms_per_opt_step": round(ms, 3), "samples_per_s": round(4 * B / ms * 1e3, 1),
            "loss": float(static_loss), "model_tflops": round(flops / ms / 1e9, 1)}
    print(json.dumps(out2), flush=True)
    # stock torch reference point: HF-style BERT via torch SDPA is not available offline; skip here.
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/bert_timing.json", "w") as f:
        json.dump([out, out2], f, indent=1)


if __name__ == "__main__":
    ok = numerics()
    try:
        timing()
    except Exception:
        import traceback
        traceback.print_exc()
        ok = False
    sys.exit(0 if ok else 1)
