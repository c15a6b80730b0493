"""Tiled (online-softmax) attention for sequences longer than 128 tokens, the any-width LayerNorm kernels and the
SQuAD-shape models, each against a plain fp32 PyTorch statement of the same math."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _F():
    from determined_b200.ops import _native
    from determined_b200.ops import functional as F

    _native.load(build_if_missing=False)
    return F


def rel(out, ref):
    return ((out.float() - ref.float()).abs().max() / ref.float().abs().max().clamp_min(1e-6)).item()


def _ref_attention(qkv, mask, B, S, H):
    hid = H * 64
    q, k, v = [t.float().view(B, S, H, 64).transpose(1, 2).requires_grad_(True) for t in qkv.split(hid, dim=1)]
    sc = q @ k.transpose(-1, -2) * 0.125
    if mask is not None:
        sc = sc + mask[:, None, None, :]
    o = torch.softmax(sc, dim=-1) @ v
    return (q, k, v), sc, o.transpose(1, 2).reshape(B * S, hid)


@pytest.mark.parametrize("B,S,H,use_mask", [(2, 256, 2, False), (2, 384, 12, True), (3, 200, 3, True), (1, 129, 1, False),
                                            (1, 1024, 2, True)])
def test_flash_attention_fwd_bwd(B, S, H, use_mask):
    F = _F()
    torch.manual_seed(S + H)
    hid = H * 64
    qkv = (torch.randn(B * S, 3 * hid, device="cuda") * 0.5).bfloat16()
    mask = None
    if use_mask:
        mask = torch.zeros(B, S, device="cuda")
        mask[:, S - S // 4:] = -10000.0
        mask[0, 5:9] = -10000.0
    ctx, lse = F.attention_fwd(qkv, mask, B, S, H, 0.0, 0)
    (q, k, v), sc, ref = _ref_attention(qkv, mask, B, S, H)
    assert rel(ctx, ref) < 2e-2
    assert rel(lse, torch.logsumexp(sc, dim=-1)) < 1e-2
    dctx = (torch.randn(B * S, hid, device="cuda") * 0.5).bfloat16()
    dqkv = F.attention_bwd(qkv, dctx, mask, lse, B, S, H, 0.0, 0, ctx)
    ref.backward(dctx.float())
    ref_d = torch.cat([t.grad.transpose(1, 2).reshape(B * S, hid) for t in (q, k, v)], dim=1)
    for name, a, b in zip("qkv", dqkv.split(hid, dim=1), ref_d.split(hid, dim=1)):
        assert rel(a, b) < 3e-2, name


def test_flash_attention_matches_single_tile_kernel_at_128():
    """At S = 128 both kernels apply; they must agree (including the dropout mask, which hashes the same indices)."""
    from determined_b200.ops import _native
    from determined_b200.ops.functional import _p, _stream, step_seed_tensor

    F = _F()
    torch.manual_seed(1)
    B, S, H = 2, 128, 3
    hid = H * 64
    qkv = (torch.randn(B * S, 3 * hid, device="cuda") * 0.5).bfloat16()
    F.set_step_seed(5)
    for p in (0.0, 0.1):
        a, lse_a = F.attention_fwd(qkv, None, B, S, H, p, 77)
        b = torch.empty_like(a)
        lse_b = torch.empty_like(lse_a)
        _native.call("det_attention_flash_fwd", _p(qkv), None, _p(b), _p(lse_b), B, S, H, float(p), 77,
                     _p(step_seed_tensor(qkv.device)) if p > 0 else None, _stream())
        assert rel(b, a) < 1e-2 and rel(lse_b, lse_a) < 1e-3
        dctx = torch.randn_like(a)
        da = F.attention_bwd(qkv, dctx, None, lse_a, B, S, H, p, 77)
        db = torch.empty_like(qkv)
        acc = torch.empty(B * S, hid, device="cuda")
        delta = torch.empty(B, H, S, device="cuda")
        _native.call("det_attention_flash_bwd", _p(qkv), _p(dctx), _p(a), None, _p(lse_a), _p(db), _p(acc), _p(delta), B, S, H,
                     float(p), 77, _p(step_seed_tensor(qkv.device)) if p > 0 else None, _stream())
        assert rel(db, da) < 2e-2


def test_flash_attention_dropout_adjoint():
    F = _F()
    torch.manual_seed(4)
    B, S, H = 2, 320, 2
    hid = H * 64
    qkv = (torch.randn(B * S, 3 * hid, device="cuda") * 0.5).bfloat16()
    F.set_step_seed(11)
    c1, lse = F.attention_fwd(qkv, None, B, S, H, 0.1, 55)
    dctx = torch.randn_like(c1)
    dV = F.attention_bwd(qkv, dctx, None, lse, B, S, H, 0.1, 55, c1)[:, 2 * hid:].float()
    qkv2 = qkv.clone()
    pert = (torch.randn(B * S, hid, device="cuda") * 0.25).bfloat16()
    qkv2[:, 2 * hid:] = pert
    c2, _ = F.attention_fwd(qkv2, None, B, S, H, 0.1, 55)
    lhs = (c2.float() * dctx.float()).sum().item()      # O is linear in V for a fixed mask
    rhs = (dV * pert.float()).sum().item()
    assert abs(lhs - rhs) / max(abs(lhs), 1e-3) < 3e-2


@pytest.mark.parametrize("H", [128, 1536, 4096])
def test_layernorm_any_width(H):
    F = _F()
    torch.manual_seed(5)
    T = 777
    x = torch.randn(T, H, device="cuda").bfloat16()
    g = torch.randn(H, device="cuda") * 0.1 + 1.0
    b = torch.randn(H, device="cuda") * 0.1
    y, mean, rstd = F.layernorm_fwd(x, g, b, 1e-12)
    xf, gf, bf = x.float().requires_grad_(True), g.clone().requires_grad_(True), b.clone().requires_grad_(True)
    ref = torch.nn.functional.layer_norm(xf, (H,), gf, bf, 1e-12)
    assert rel(y, ref) < 1e-2
    dy = torch.randn(T, H, device="cuda").bfloat16()
    dg, db = torch.zeros(H, device="cuda"), torch.zeros(H, device="cuda")
    dxs = torch.zeros(H, device="cuda")
    dx, _ = F.layernorm_bwd(dy, x, g, mean, rstd, dg, db, dxsum=dxs)
    ref.backward(dy.float())
    assert rel(dx, xf.grad) < 1.5e-2 and rel(dg, gf.grad) < 1e-2 and rel(db, bf.grad) < 1e-2
    want = xf.grad.sum(0)
    assert (dxs - want).abs().max().item() < 2e-2 * max(want.abs().max().item(), 1.0) + 0.05


@pytest.mark.parametrize("kind", ["bert", "albert"])
def test_squad_models_match_cpu_reference(kind):
    from determined_b200.models.bert import AlbertConfig, AlbertForQuestionAnswering, BertConfig, BertForQuestionAnswering

    _F()
    torch.manual_seed(0)
    if kind == "bert":
        cfg = BertConfig(num_hidden_layers=2, hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0, vocab_size=1000)
        make = lambda: BertForQuestionAnswering(cfg)
    else:
        cfg = AlbertConfig(num_hidden_layers=3, vocab_size=1000, hidden_size=256, num_attention_heads=4, intermediate_size=512)
        make = lambda: AlbertForQuestionAnswering(cfg)
    ref, gpu = make(), make()
    gpu.load_state_dict(ref.state_dict())
    gpu.cuda()
    B, S = 3, 384
    ids = torch.randint(0, 1000, (B, S))
    am = torch.ones(B, S, dtype=torch.long)
    am[:, 300:] = 0
    tt = torch.randint(0, 2, (B, S))
    st, en = torch.tensor([10, 120, 250]), torch.tensor([14, 130, 251])
    out_r = ref(ids, am, tt, st, en)
    out_r[0].backward()
    out_g = gpu(ids.cuda(), am.cuda(), tt.cuda(), st.cuda(), en.cuda())
    out_g[0].backward()
    assert abs(out_r[0].item() - out_g[0].item()) < 3e-2
    assert rel(out_g[1][:, :300].cpu(), out_r[1][:, :300]) < 5e-2
    for (n, pr), (_, pg) in zip(ref.named_parameters(), gpu.named_parameters()):
        # the span softmax is shift invariant, so the gradients of qa_outputs.bias and of the last LayerNorm bias are
        # identically zero (rounding noise in both implementations): nothing to compare
        if n.endswith("_anchor") or pr.grad.abs().max() < 1e-6:
            continue
        cos = torch.nn.functional.cosine_similarity(pr.grad.flatten(), pg.grad.cpu().flatten(), dim=0).item()
        assert cos > 0.99, (n, cos)


def test_squad_trial_trains_on_gpu():
    """A small BERT span model through the PyTorchTrial controller at seq 384: the loss must fall."""
    from determined_b200.models.squad_trials import BertSquadTrial
    from tests.harness_utils import TrainAndValidate, make_controller

    hp = {"global_batch_size": 8, "learning_rate": 1e-3, "num_hidden_layers": 2, "hidden_size": 256, "num_attention_heads": 4,
          "intermediate_size": 512, "vocab_size": 2000, "max_seq_length": 384, "dataset_size": 64, "validation_size": 16,
          "dropout": 0.0, "num_training_steps": 100}
    tv = TrainAndValidate()
    make_controller(BertSquadTrial, hp, tv.send(steps=3, validation_freq=3, scheduling_unit=10)).run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert len(losses) == 30 and sum(losses[-5:]) < sum(losses[:5]) * 0.9, losses
    assert {"validation_loss", "exact", "f1"} <= set(tv.validation_metrics[-1])
