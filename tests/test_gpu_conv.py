"""sm_100a convolution stack against fp32 torch references (forward, input / weight / bias gradients)."""
import pytest
import torch
import torch.nn as nn

pytestmark = pytest.mark.gpu


def _bf(t):
    return t.to(torch.bfloat16).float()


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


def _compare(mine, ref, x, tol=2e-2):
    from determined_b200.ops import _native

    before = _native.launch_count
    ref = ref.cuda().float()
    mine = mine.cuda()
    mine.load_state_dict(ref.state_dict())
    xr = x.clone().requires_grad_(True)
    xm = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    yr, ym = ref(xr), mine(xm)
    assert _rel(ym, yr) < tol
    g = _bf(torch.randn_like(yr))
    yr.backward(g)
    ym.backward(g.to(torch.bfloat16))
    assert _rel(xm.grad, xr.grad) < tol
    for (k, pm), (_, pr) in zip(mine.named_parameters(), ref.named_parameters()):
        gm = pm.grad.permute(0, 3, 1, 2) if pm.grad.dim() == 4 else pm.grad
        assert _rel(gm, pr.grad) < tol, k
    assert _native.launch_count > before, "the native kernels did not run"


@pytest.mark.parametrize("N,H,W,Cin,Cout,k,s,p", [
    (4, 56, 56, 64, 64, 3, 1, 1),      # implicit GEMM, box 56x2
    (3, 14, 14, 256, 256, 3, 1, 1),    # implicit GEMM, box 14x7
    (5, 7, 7, 512, 512, 3, 1, 1),      # implicit GEMM, boxes span images
    (2, 10, 13, 32, 48, 3, 1, 1),      # channel tails
    (2, 12, 12, 32, 64, 3, 1, 0),      # no padding
    (4, 14, 14, 256, 1024, 1, 1, 0),   # 1x1 → plain GEMM
    (2, 28, 28, 128, 128, 3, 2, 1),    # strided → gather path
    (2, 64, 64, 3, 64, 7, 2, 3),       # RGB stem
    (4, 4, 4, 512, 1, 4, 1, 0),        # single output channel (DCGAN discriminator head)
])
def test_conv2d(N, H, W, Cin, Cout, k, s, p):
    from determined_b200.ops import conv as C

    torch.manual_seed(N + H + Cin + Cout)
    x = _bf(torch.randn(N, Cin, H, W, device="cuda"))
    ref = nn.Conv2d(Cin, Cout, k, s, p)
    with torch.no_grad():
        ref.weight.copy_(_bf(ref.weight))
    _compare(C.Conv2d(Cin, Cout, k, s, p), ref, x)


@pytest.mark.parametrize("N,H,Cin,Cout,k,s,p", [(4, 1, 100, 512, 4, 1, 0), (2, 8, 256, 128, 4, 2, 1), (2, 16, 64, 3, 4, 2, 1)])
def test_conv_transpose2d(N, H, Cin, Cout, k, s, p):
    from determined_b200.ops import conv as C

    torch.manual_seed(H + Cin)
    x = _bf(torch.randn(N, Cin, H, H, device="cuda"))
    ref = nn.ConvTranspose2d(Cin, Cout, k, s, p)
    with torch.no_grad():
        ref.weight.copy_(_bf(ref.weight))
    _compare(C.ConvTranspose2d(Cin, Cout, k, s, p), ref, x)


@pytest.mark.parametrize("C_,act,resid", [(64, "relu", True), (256, None, False), (128, "leaky_relu", False), (2048, "relu", True)])
def test_batchnorm_fused(C_, act, resid):
    from determined_b200.ops import conv as C

    torch.manual_seed(C_)
    x = _bf(torch.randn(4, C_, 7, 9, device="cuda") * 2 + 0.5)
    r = _bf(torch.randn_like(x)) if resid else None
    ref = nn.BatchNorm2d(C_).cuda()
    mine = C.BatchNorm2d(C_, activation=act, negative_slope=0.2).cuda()
    with torch.no_grad():
        ref.weight.copy_(torch.rand(C_) + 0.5)
        ref.bias.copy_(torch.randn(C_) * 0.2)
    mine.load_state_dict(ref.state_dict())
    xr = x.clone().requires_grad_(True)
    xm = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    rr = r.clone().requires_grad_(True) if resid else None
    rm = r.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True) if resid else None
    yr = ref(xr)
    if resid:
        yr = yr + rr
    if act == "relu":
        yr = torch.relu(yr)
    elif act == "leaky_relu":
        yr = nn.functional.leaky_relu(yr, 0.2)
    ym = mine(xm, rm)
    assert _rel(ym, yr) < 2e-2
    assert _rel(mine.running_var, ref.running_var) < 1e-2 and _rel(mine.running_mean, ref.running_mean) < 1e-2
    g = _bf(torch.randn_like(yr))
    yr.backward(g)
    ym.backward(g.to(torch.bfloat16))
    assert _rel(xm.grad, xr.grad) < 3e-2
    assert _rel(mine.weight.grad, ref.weight.grad) < 2e-2 and _rel(mine.bias.grad, ref.bias.grad) < 2e-2
    if resid:
        assert _rel(rm.grad, rr.grad) < 2e-2


def test_pools_activations_dropout():
    from determined_b200.ops import conv as C

    torch.manual_seed(5)
    x = _bf(torch.randn(3, 16, 17, 17, device="cuda"))
    _compare(C.MaxPool2d(3, 2, 1), nn.MaxPool2d(3, 2, 1), x, tol=1e-2)
    _compare(C.MaxPool2d(2), nn.MaxPool2d(2), x, tol=1e-2)

    class GAP(nn.Module):
        def forward(self, t):
            return t.mean(dim=(2, 3))
    _compare(C.GlobalAvgPool2d(), GAP(), x, tol=1e-2)
    for kind, refm in [("relu", nn.ReLU()), ("leaky_relu", nn.LeakyReLU(0.2)), ("tanh", nn.Tanh()), ("sigmoid", nn.Sigmoid())]:
        _compare(C.Activation(kind, 0.2), refm, x, tol=1e-2)
    d = C.Dropout(0.25).cuda().train()
    ones = torch.ones(4, 64, 16, 16, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    y = d(ones)
    assert abs((y.float() > 0).float().mean().item() - 0.75) < 0.01
    y.backward(torch.ones_like(y))
    assert torch.equal(ones.grad.float() > 0, y.float() > 0)


def test_resnet50_tracks_torchvision():
    """Same weights (state_dict interchange), same batch, same SGD: the loss trajectory of our ResNet-50 must follow
    torchvision's fp32 one, and with frozen BatchNorm statistics the weight gradients must point the same way."""
    import torchvision

    from determined_b200.models.resnet import resnet50
    from determined_b200.ops.modules import cross_entropy
    from determined_b200.parallel.fused_optimizer import FusedOptimizer

    torch.manual_seed(0)
    ref = torchvision.models.resnet50(num_classes=16).cuda()
    mine = resnet50(num_classes=16).cuda()
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(16, 3, 64, 64, device="cuda")
    y = torch.randint(0, 16, (16,), device="cuda")
    ref.eval(); mine.eval()
    torch.nn.functional.cross_entropy(ref(x), y).backward()
    cross_entropy(mine(x), y).backward()
    cos = []
    mp = dict(mine.named_parameters())
    for k, p in ref.named_parameters():
        if p.dim() == 4 and mp[k].grad is not None:
            g = mp[k].grad.permute(0, 3, 1, 2).flatten().float()
            cos.append(torch.nn.functional.cosine_similarity(g, p.grad.flatten(), dim=0).item())
    cos.sort()
    assert cos[len(cos) // 2] > 0.98 and cos[0] > 0.9
    ref.train(); mine.train()
    ropt = torch.optim.SGD(ref.parameters(), lr=0.002, momentum=0.9)
    mopt = FusedOptimizer(torch.optim.SGD(mine.parameters(), lr=0.002, momentum=0.9))
    ropt.zero_grad()
    mopt.zero_grad()
    tr, tm = [], []
    for _ in range(6):
        l1 = torch.nn.functional.cross_entropy(ref(x), y)
        l1.backward()
        ropt.step()
        ropt.zero_grad()
        l2 = cross_entropy(mine(x), y)
        l2.backward()
        mopt.step(zero_grad=True)
        tr.append(l1.item())
        tm.append(l2.item())
    assert all(abs(a - b) / abs(a) < 0.3 for a, b in zip(tr, tm)), (tr, tm)


@pytest.mark.parametrize("N,H,Cin,Cout,k,s,p", [(4, 14, 64, 256, 1, 1, 0), (3, 14, 64, 64, 3, 1, 1), (2, 28, 128, 128, 3, 2, 1), (2, 9, 72, 136, 3, 1, 1)])
def test_conv_bn_fused_statistics(N, H, Cin, Cout, k, s, p):
    """conv_bn takes the BatchNorm batch statistics from the convolution's epilogue (GEMM, implicit-GEMM and gather
    paths). Against torch (fp32): output and running statistics. Against the SAME modules run unfused (identical bf16
    numerics, statistics from the separate reduction kernel): output and every gradient, tightly — comparing gradients
    with fp32 torch here would mostly measure ReLU-mask flips of near-zero activations, not the fusion."""
    from determined_b200.ops import conv as C

    torch.manual_seed(Cin + Cout + k)
    x = _bf(torch.randn(N, Cin, H, H, device="cuda"))
    rconv = nn.Conv2d(Cin, Cout, k, s, p, bias=False).cuda()
    rbn = nn.BatchNorm2d(Cout).cuda()
    with torch.no_grad():
        rconv.weight.copy_(_bf(rconv.weight))
        rbn.weight.copy_(torch.rand(Cout, device="cuda") + 0.5)
        rbn.bias.copy_(torch.randn(Cout, device="cuda") * 0.2)

    def mine():
        conv = C.Conv2d(Cin, Cout, k, s, p, bias=False).cuda()
        bn = C.BatchNorm2d(Cout, activation="relu").cuda()
        conv.load_state_dict(rconv.state_dict())
        bn.load_state_dict(rbn.state_dict())
        return conv, bn

    fconv, fbn = mine()
    uconv, ubn = mine()
    xf = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    xu = xf.detach().clone().requires_grad_(True)
    yr = torch.relu(rbn(rconv(x)))
    yf = C.conv_bn(fconv, fbn, xf)
    yu = ubn(uconv(xu))
    assert _rel(yf, yr) < 3e-2
    assert _rel(fbn.running_mean, rbn.running_mean) < 2e-2 and _rel(fbn.running_var, rbn.running_var) < 2e-2
    assert _rel(yf, yu) < 1e-2 and _rel(fbn.running_var, ubn.running_var) < 1e-3
    g = _bf(torch.randn_like(yr)).to(torch.bfloat16)
    yf.backward(g)
    yu.backward(g)
    assert _rel(xf.grad, xu.grad) < 2e-2
    assert _rel(fconv.weight.grad, uconv.weight.grad) < 2e-2
    assert _rel(fbn.weight.grad, ubn.weight.grad) < 2e-2 and _rel(fbn.bias.grad, ubn.bias.grad) < 2e-2


def test_protonet_meta_learns_on_gpu():
    """Few-shot episodes through the conv stack (1-channel stem → gather path, fused BN+ReLU, pooling) and the controller."""
    from determined_b200.models import protonet_trial as P
    from determined_b200.ops import _native
    from tests.harness_utils import TrainAndValidate, make_controller

    before = _native.launch_count
    hp = {"global_batch_size": 2, "val_batch_size": 2, "learning_rate": 2e-3, "num_classes_train": 20, "num_support_train": 1,
          "num_query_train": 5, "num_classes_val": 5, "num_support_val": 1, "num_query_val": 5, "hidden_dim": 64, "embedding_dim": 64}
    tv = TrainAndValidate()
    make_controller(P.OmniglotProtoNetTrial, hp, tv.send(steps=3, validation_freq=3, scheduling_unit=10),
                    exp_config={"data": {"tasks_per_epoch_train": 60, "tasks_per_epoch_val": 20, "num_characters": 200},
                                "searcher": {"name": "single", "metric": "loss", "max_length": {"batches": 100}}}).run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert sum(losses[-5:]) < sum(losses[:5]) and tv.validation_metrics[-1]["acc"] > 0.5
    assert _native.launch_count > before


@pytest.mark.parametrize("C_,k,s,p,d,H", [(64, 3, 1, 1, 1, 16), (32, 5, 1, 2, 1, 9), (128, 3, 2, 1, 1, 17), (16, 3, 1, 2, 2, 12), (72, 5, 2, 4, 2, 15),
                                          (512, 3, 1, 1, 1, 8)])
def test_depthwise_conv2d(C_, k, s, p, d, H):
    from determined_b200.ops import conv as C

    torch.manual_seed(C_ + k + d)
    x = _bf(torch.randn(3, C_, H, H + 2, device="cuda"))
    ref = nn.Conv2d(C_, C_, k, s, p, d, groups=C_, bias=True).cuda()
    mine = C.DepthwiseConv2d(C_, k, stride=s, padding=p, dilation=d, bias=True).cuda()
    mine.load_state_dict(ref.state_dict())
    xr = x.clone().requires_grad_(True)
    xm = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    yr, ym = ref(xr), mine(xm)
    assert ym.shape == yr.shape and _rel(ym, yr) < 1e-2
    g = _bf(torch.randn_like(yr))
    yr.backward(g)
    ym.backward(g.to(torch.bfloat16))
    assert _rel(xm.grad, xr.grad) < 2e-2
    assert _rel(mine.weight.grad.permute(2, 0, 1).unsqueeze(1), ref.weight.grad) < 2e-2
    assert _rel(mine.bias.grad, ref.bias.grad) < 2e-2


@pytest.mark.parametrize("k,s,p,inc", [(3, 1, 1, False), (3, 2, 1, False), (5, 3, 0, False), (2, 2, 0, True), (3, 2, 1, True)])
def test_avgpool2d(k, s, p, inc):
    from determined_b200.ops import conv as C

    torch.manual_seed(k + s)
    x = _bf(torch.randn(3, 24, 17, 14, device="cuda"))
    _compare(C.AvgPool2d(k, s, p, count_include_pad=inc), nn.AvgPool2d(k, s, p, count_include_pad=inc), x, tol=1e-2)


def test_darts_network_matches_cpu_and_trains():
    """Same weights, same batch: the sm_100a network follows the fp32 CPU statement of the same graph (every primitive in
    use), and the trial learns through the controller."""
    from determined_b200.models import darts as D
    from determined_b200.models.darts_trial import DARTSCNNTrial
    from determined_b200.ops import _native
    from tests.harness_utils import TrainAndValidate, make_controller

    torch.manual_seed(0)
    hp, ops, j = {}, D.PRIMITIVES[1:], 0
    for cell in ("normal", "reduce"):
        for i in range(1, 5):
            for e in range(1, 3):
                hp[f"{cell}_node{i}_edge{e}_op"] = ops[j % len(ops)]
                hp[f"{cell}_node{i}_edge{e}"] = j % (i + 1)
                j += 1
    g = D.genotype_from_hparams(hp)
    cpu = D.DartsNetworkCIFAR(16, 10, layers=3, auxiliary=True, genotype=g)
    gpu = D.DartsNetworkCIFAR(16, 10, layers=3, auxiliary=True, genotype=g)
    gpu.load_state_dict(cpu.state_dict())
    gpu.cuda()
    x = torch.randn(16, 3, 32, 32)
    before = _native.launch_count
    lc, ac = cpu(x)
    lg, ag = gpu(x.cuda())
    assert _native.launch_count - before > 50
    assert _rel(lg.cpu(), lc) < 0.1 and _rel(ag.cpu(), ac) < 0.1
    (lc.square().mean() + ac.square().mean()).backward()
    (lg.float().square().mean() + ag.float().square().mean()).backward()
    cos = []
    for (n, pc), (_, pg) in zip(cpu.named_parameters(), gpu.named_parameters()):
        if pc.grad is None or pc.grad.abs().max() < 1e-7:
            continue
        a, b = pc.grad.flatten(), pg.grad.cpu().flatten().float()
        cos.append(torch.nn.functional.cosine_similarity(a, b, dim=0).item())
    cos.sort()
    assert cos[len(cos) // 2] > 0.93 and cos[len(cos) // 10] > 0.8, cos[:10]      # bf16 through ~20 conv / BN layers, batch of 16
    hp2 = dict(hp, global_batch_size=32, learning_rate=0.05, init_channels=16, layers=3, train_epochs=5, drop_path_prob=0.1, cutout_length=8,
               use_cuda_graphs=True)
    tv = TrainAndValidate()
    make_controller(DARTSCNNTrial, hp2, tv.send(steps=2, validation_freq=2, scheduling_unit=10),
                    exp_config={"data": {"train_size": 128, "validation_size": 64},
                                "searcher": {"name": "single", "metric": "top1_accuracy", "smaller_is_better": False, "max_length": {"batches": 100}}}).run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert sum(losses[-5:]) < sum(losses[:5]), losses


def test_gaea_supernet_on_gpu():
    """The weight-sharing supernet (every candidate operation on every edge) trains on the sm_100a modules: both the
    shared weights and the simplex-constrained architecture parameters move."""
    from determined_b200.models import gaea_search as G
    from determined_b200.ops import _native
    from tests.harness_utils import TrainAndValidate, make_controller

    before = _native.launch_count
    hp = {"global_batch_size": 32, "learning_rate": 0.05, "arch_learning_rate": 0.5, "init_channels": 16, "layers": 3, "train_epochs": 5}
    tv = TrainAndValidate()
    c = make_controller(G.GAEASearchTrial, hp, tv.send(steps=2, validation_freq=2, scheduling_unit=5),
                        exp_config={"data": {"train_size": 128, "validation_size": 64},
                                    "searcher": {"name": "single", "metric": "top1_accuracy", "smaller_is_better": False, "max_length": {"batches": 100}}})
    c.run()
    a = c.trial.model.alphas_normal.detach().float().cpu()
    assert torch.allclose(a.sum(-1), torch.ones(14), atol=1e-4) and (a - 1 / 8).abs().max() > 1e-4
    assert all(torch.isfinite(torch.tensor(m["loss"])) for m in tv.training_metrics)
    assert _native.launch_count - before > 500
    assert len(c.trial.model.genotype().normal) == 8
