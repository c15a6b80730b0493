"""DARTS recurrent cell at the published sizes (T=35, B=64, width 850, bf16): the cell as published (one GEMM for the
initial state and one per node, every time step) against the regrouped schedule of models/darts_rnn.py (input GEMM
hoisted out of the recurrence, sibling nodes batched) with torch's elementwise gates and with the fused gate kernels of
csrc/kernels/rnn_ops.cu, forward + backward, eager and as a replayed CUDA graph.
Writes gpurun_out/darts_rnn_cell.json."""
import copy
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.models import darts_rnn as R  # noqa: E402


def naive_cell(cell, inputs, hidden, x_mask, h_mask):
    nh = cell.nhid
    w0 = cell._W0.to(inputs.dtype)
    ws = [w.to(inputs.dtype) for w in cell._Ws]
    h = hidden[0]
    outs = []
    for t in range(inputs.shape[0]):
        c0, h0 = (torch.cat([inputs[t] * x_mask, h * h_mask], -1) @ w0).split(nh, -1)
        states = [h + c0.sigmoid() * (torch.tanh(h0) - h)]
        for i, (name, pred) in enumerate(cell.genotype.recurrent):
            sp = states[pred]
            c, hh = ((sp * h_mask) @ ws[i]).split(nh, -1)
            states.append(sp + c.sigmoid() * (R.ACTIVATIONS[name](hh) - sp))
        h = torch.stack([states[i] for i in cell.genotype.concat], -1).mean(-1)
        outs.append(h)
    return torch.stack(outs)


def timed(fn, iters=20, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    torch.manual_seed(0)
    T, B, H = 35, 64, 850
    dev = torch.device("cuda")
    cell = R.DARTSCell(H, H, 0.25, 0.75, R.DARTS_V2).to(dev)
    x = torch.randn(T, B, H, device=dev, dtype=torch.bfloat16)
    h0 = torch.zeros(1, B, H, device=dev, dtype=torch.bfloat16)
    xm, hm = R.locked_mask(x, B, H, 0.75), R.locked_mask(x, B, H, 0.25)

    def step(kind, cell=cell):
        for p in cell.parameters():
            p.grad = None
        if kind == "naive":
            out = naive_cell(cell, x, h0, xm, hm)
        else:
            out = cell(x, h0, masks=(xm, hm), fused=(kind == "fused"))[0]
        out.float().pow(2).mean().backward()
        return out

    a, b, c = step("regrouped").float(), step("naive").float(), step("fused").float()
    res = {"T": T, "B": B, "width": H, "dtype": "bf16", "max_abs_diff_outputs": (a - b).abs().max().item(),
           "max_abs_diff_fused_vs_naive": (c - b).abs().max().item(), "output_scale": b.abs().max().item()}
    for kind in ("naive", "regrouped", "fused"):
        res[f"{kind}_eager_ms"] = round(timed(lambda: step(kind)), 3)
        # a fresh copy of the cell: its gradient accumulators are created on the capture stream (the ones of `cell` are
        # bound to the default stream by the eager runs above, which capture refuses to mix in)
        cell_g = copy.deepcopy(cell)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                step(kind, cell_g)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        for p in cell_g.parameters():
            p.grad = None
        try:
            with torch.cuda.graph(g, stream=s):
                step(kind, cell_g)
            res[f"{kind}_graph_ms"] = round(timed(g.replay), 3)
        except Exception as e:      # keep the eager numbers if capture is refused
            res[f"{kind}_graph_error"] = str(e).splitlines()[0]
    res["speedup_eager"] = round(res["naive_eager_ms"] / res["regrouped_eager_ms"], 2)
    if "naive_graph_ms" in res and "regrouped_graph_ms" in res:
        res["speedup_graph"] = round(res["naive_graph_ms"] / res["regrouped_graph_ms"], 2)
    res["speedup_fused_eager"] = round(res["naive_eager_ms"] / res["fused_eager_ms"], 2)
    if "naive_graph_ms" in res and "fused_graph_ms" in res:
        res["speedup_fused_graph"] = round(res["naive_graph_ms"] / res["fused_graph_ms"], 2)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/darts_rnn_cell.json", "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
