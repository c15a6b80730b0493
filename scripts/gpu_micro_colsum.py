"""Column-sum (bias gradient) kernel on the four BERT-base shapes of one layer's backward; prints per-call microseconds and
the effective bandwidth. The mixture averaged 5.44 us per call before the kernel kept four loads in flight per thread."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import functional as F  # noqa: E402

T = 6144
rows = []
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for N in (768, 3072, 768, 2304):
    x = torch.randn(T, N, device="cuda").bfloat16()
    out = torch.zeros(N, device="cuda")
    F.colsum_accum(x, out)
    torch.cuda.synchronize()
    ref = x.float().sum(0)
    err = ((out - ref).abs().max() / ref.abs().max()).item()
    ts = []
    for _ in range(20):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        F.colsum_accum(x, out)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    us = ts[len(ts) // 2]
    rows.append({"T": T, "N": N, "us": round(us, 2), "GBps": round(T * N * 2 / us / 1e3, 1), "rel_err": err})
    print(json.dumps(rows[-1]), flush=True)
print(json.dumps({"mean_us": round(sum(r["us"] for r in rows) / len(rows), 2)}))
