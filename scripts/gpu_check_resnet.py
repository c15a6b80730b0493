"""Whole-network check: our ResNet-50 against torchvision's (fp32, same weights through load_state_dict): loss, logits,
per-parameter gradient cosine similarity, then loss trajectories of 10 SGD steps on one repeated batch."""
import os
import sys

import torch
import torchvision

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.models.resnet import resnet50  # noqa: E402
from determined_b200.ops.modules import cross_entropy  # noqa: E402
from determined_b200.parallel.fused_optimizer import FusedOptimizer  # noqa: E402

torch.manual_seed(0)
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
NC = 16
ref = torchvision.models.resnet50(num_classes=NC).cuda().train()
mine = resnet50(NC).cuda().train()
mine.load_state_dict(ref.state_dict())
res = int(os.environ.get("RES", 64))
x = torch.randn(16, 3, res, res, device="cuda")
y = torch.randint(0, NC, (16,), device="cuda")

def grad_report(tag):
    for p in list(ref.parameters()) + list(mine.parameters()):
        p.grad = None
    lr_ = torch.nn.functional.cross_entropy(ref(x), y)
    lr_.backward()
    lm = cross_entropy(mine(x), y)
    lm.backward()
    print(f"[{tag}] loss ref {lr_.item():.5f} mine {lm.item():.5f}")
    rows = []
    mp = dict(mine.named_parameters())
    for k, p in ref.named_parameters():
        g = mp[k].grad
        if g is None:
            print("NO GRAD", k)
            continue
        if g.dim() == 4:
            g = g.permute(0, 3, 1, 2)
        cos = torch.nn.functional.cosine_similarity(g.flatten().float(), p.grad.flatten(), dim=0).item()
        ratio = (g.float().norm() / p.grad.norm().clamp_min(1e-12)).item()
        rows.append((cos, ratio, k))
    rows.sort()
    print(f"[{tag}] lowest cosine similarities (cos, |mine|/|ref|, name):")
    for w in rows[:6]:
        print(f"  {w[0]:.4f} {w[1]:.3f} {w[2]}")
    convs = [r for r in rows if r[2].endswith("conv1.weight") or r[2].endswith("conv2.weight") or r[2].endswith("conv3.weight") or r[2] == "fc.weight"]
    med_all, med_conv = rows[len(rows) // 2][0], convs[len(convs) // 2][0]
    print(f"[{tag}] median cos all {med_all:.4f}, conv/fc weights {med_conv:.4f}, min conv {convs[0][0]:.4f} {convs[0][2]}")
    return med_all, med_conv


# frozen statistics first: without the batch-statistics feedback the comparison is well conditioned
ref.eval(); mine.eval()
med_eval, med_eval_conv = grad_report("eval-mode BN")
ref.train(); mine.train()
grad_report("train-mode BN")

ropt = torch.optim.SGD(ref.parameters(), lr=0.02, momentum=0.9)
mopt = FusedOptimizer(torch.optim.SGD(mine.parameters(), lr=0.02, momentum=0.9))
ropt.zero_grad()
mopt.zero_grad()
tr, tm = [], []
for _ in range(10):
    l1 = torch.nn.functional.cross_entropy(ref(x), y)
    l1.backward()
    ropt.step()
    ropt.zero_grad()
    l2 = cross_entropy(mine(x), y)
    l2.backward()
    mopt.step(zero_grad=True)
    tr.append(round(l1.item(), 4))
    tm.append(round(l2.item(), 4))
print("ref :", tr)
print("mine:", tm)
close = max(abs(a - b) / max(abs(a), 1e-6) for a, b in zip(tr, tm))
print(f"max relative gap between the trajectories: {close:.3f}")
ok = med_eval_conv > 0.98 and close < 0.35
print("RESNET CHECK", "PASS" if ok else "FAIL")
sys.exit(0 if ok else 1)
