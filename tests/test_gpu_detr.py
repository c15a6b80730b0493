"""DETR on the GPU: the network built from the sm_100a modules against the same weights evaluated by plain PyTorch on the
CPU, a short training run through the controller, and the per-batch advance of the dropout step counter."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _load():
    from determined_b200.ops import _native

    _native.load(build_if_missing=False)


def test_detr_forward_backward_tracks_cpu_reference():
    from determined_b200.models import detr

    _load()
    torch.manual_seed(0)
    make = lambda: detr.DETR(5, 12, 128, 2, 2, 2, 256, 0.0, False, "resnet18", True)      # noqa: E731
    ref, gpu = make(), make()
    gpu.load_state_dict(ref.state_dict())
    gpu.cuda()
    ref.train()
    gpu.train()
    images = torch.rand(2, 3, 160, 192)
    mask = torch.zeros(2, 160, 192, dtype=torch.bool)
    mask[1, :, 128:] = True
    images[1, :, :, 128:] = 0
    targets = [{"labels": torch.tensor([1, 3]), "boxes": torch.tensor([[0.3, 0.3, 0.2, 0.2], [0.6, 0.5, 0.3, 0.4]])},
               {"labels": torch.tensor([0]), "boxes": torch.tensor([[0.25, 0.5, 0.2, 0.5]])}]
    out_r = ref(images, mask)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out_g = gpu(images.cuda(), mask.cuda())
    assert (out_g["pred_boxes"].cpu() - out_r["pred_boxes"]).abs().max() < 4e-2
    assert (out_g["pred_logits"].cpu() - out_r["pred_logits"]).abs().max() < 0.15
    assert len(out_g["aux_outputs"]) == 1
    wd = detr.build_weight_dict(5, 2, 2, True)
    # a fixed assignment: at random initialisation the matching costs are near ties, and bf16 noise could break them
    # differently on the two sides
    fixed = lambda outputs, tgts: [(torch.tensor([0, 1]), torch.tensor([0, 1])), (torch.tensor([2]), torch.tensor([0]))]  # noqa: E731
    crit = detr.SetCriterion(5, fixed, wd)
    loss_r = sum(v * wd[k] for k, v in crit(out_r, targets).items() if k in wd)
    crit_g = crit.cuda()
    tg = [{k: v.cuda() for k, v in t.items()} for t in targets]
    loss_g = sum(v * wd[k] for k, v in crit_g(out_g, tg).items() if k in wd)
    assert abs(loss_r.item() - loss_g.item()) < 0.05 * abs(loss_r.item()) + 0.05
    loss_r.backward()
    loss_g.backward()
    from determined_b200.ops.modules import grad_buffer

    checked = 0
    for (n, pr), (_, pg) in zip(ref.named_parameters(), gpu.named_parameters()):
        if not pr.requires_grad or pr.grad is None or pr.grad.abs().max() < 1e-5:
            continue
        g = pg.grad if pg.grad is not None else grad_buffer(pg)
        a, b = pr.grad.flatten(), g.detach().float().cpu().flatten()
        cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
        if any(k in n for k in ("query_embed", "class_embed", "bbox_embed", "decoder.1", "decoder_norm")):
            assert cos > 0.97, (n, cos)
            checked += 1
    assert checked >= 8


def test_detr_trial_trains_on_gpu_and_step_counter_advances():
    from determined_b200.models.detr_trial import DETRTrial
    from determined_b200.ops import functional as F
    from tests.harness_utils import TrainAndValidate, make_controller

    _load()
    hp = {"global_batch_size": 8, "lr": 1e-3, "lr_backbone": 1e-4, "backbone": "resnet18", "enc_layers": 2, "dec_layers": 2,
          "dim_feedforward": 256, "hidden_dim": 128, "nheads": 2, "dropout": 0.1, "num_queries": 16, "num_classes": 4,
          "image_size": 160, "dataset_size": 64, "validation_size": 16, "max_objects": 2, "clip_max_norm": 0.1}
    tv = TrainAndValidate()
    controller = make_controller(DETRTrial, hp, tv.send(steps=4, validation_freq=4, scheduling_unit=8))
    seed0 = int(F.step_seed_tensor(controller.context.device).item())
    controller.run()
    losses = [m["loss"] for m in tv.training_metrics]
    assert len(losses) == 32 and sum(losses[-8:]) < sum(losses[:8]), losses
    assert {"mAP", "mAP_50", "loss_giou"} <= set(tv.validation_metrics[-1])
    # one advance per trained batch: every batch drew its own dropout masks
    assert int(F.step_seed_tensor(controller.context.device).item()) == seed0 + 32


@pytest.mark.parametrize("graphs", [False, True])
def test_darts_rnn_trial_trains_on_gpu(graphs):
    """The DARTS recurrent-cell language model (bf16 autocast), eagerly with variable window lengths and as a replayed
    CUDA graph with fixed ones: perplexity on the synthetic corpus must fall."""
    from determined_b200.models.darts_rnn_trial import DARTSRNNTrial
    from tests.harness_utils import TrainAndValidate, make_controller
    from tests.test_darts_rnn import GENO_HP

    _load()
    hp = dict(GENO_HP, global_batch_size=16, learning_rate=20.0, emsize=64, nhid=64, nhidlast=64, bptt=16, max_seq_length_delta=8,
              dropout=0.1, dropouth=0.05, dropoutx=0.1, dropouti=0.05, dropoute=0.05, vocab_size=64, train_tokens=40000,
              valid_tokens=2000, eval_batch_size=4, fixed_seq_len=graphs, use_cuda_graphs=graphs)
    tv = TrainAndValidate()
    controller = make_controller(DARTSRNNTrial, hp, tv.send(steps=5, validation_freq=5, scheduling_unit=30))
    controller.run()
    ppl = [m["perplexity"] for m in tv.training_metrics]
    assert len(ppl) == 150 and sum(ppl[-10:]) < 0.6 * sum(ppl[:10]), (ppl[:3], ppl[-3:])
    assert tv.validation_metrics[-1]["perplexity"] < 64 * 0.6
    if graphs:
        assert controller._graphed is not None and controller._graphed.replays > 100


@pytest.mark.parametrize("k,use_mask", [(1, False), (3, True), (2, True)])
def test_highway_gate_kernels_match_fp32_reference(k, use_mask):
    from determined_b200.ops import functional as F

    _load()
    torch.manual_seed(k)
    B, nh = 5, 850
    acts = tuple((i * 3 + 1) % 4 for i in range(k))
    ch = torch.randn(B, 2 * nh * k, device="cuda").bfloat16()
    sp = torch.randn(B, nh, device="cuda").bfloat16()
    mask = (torch.bernoulli(torch.full((B, nh), 0.75, device="cuda")) / 0.75).bfloat16() if use_mask else None
    fns = [torch.tanh, torch.relu, torch.sigmoid, lambda t: t]
    chf, spf = ch.float().requires_grad_(True), sp.float().requires_grad_(True)
    ref_s = []
    for j, a in enumerate(acts):
        c, h = chf[:, 2 * nh * j:2 * nh * (j + 1)].split(nh, -1)
        ref_s.append(spf + torch.sigmoid(c) * (fns[a](h) - spf))
    s, sm = F.highway_gate_fwd(ch, sp, mask, acts)
    for j in range(k):
        assert (s[j].float() - ref_s[j]).abs().max() < 2e-2
        if use_mask:
            assert (sm[j].float() - ref_s[j] * mask.float()).abs().max() < 4e-2
    ds = [torch.randn(B, nh, device="cuda").bfloat16() if j != 1 else None for j in range(k)]
    dsm = [torch.randn(B, nh, device="cuda").bfloat16() for _ in range(k)] if use_mask else [None] * k
    loss = sum((ref_s[j] * ds[j].float()).sum() for j in range(k) if ds[j] is not None)
    if use_mask:
        loss = loss + sum((ref_s[j] * mask.float() * dsm[j].float()).sum() for j in range(k))
    loss.backward()
    dch, dsp = F.highway_gate_bwd(ch, sp, mask, ds, dsm, acts)
    assert (dch.float() - chf.grad).abs().max() < 3e-2 * max(1.0, chf.grad.abs().max().item())
    assert (dsp.float() - spf.grad).abs().max() < 3e-2 * max(1.0, spf.grad.abs().max().item())


def test_darts_cell_fused_gates_match_torch_path():
    from determined_b200.models import darts_rnn as R

    _load()
    torch.manual_seed(0)
    T, B, H = 9, 6, 850
    cell = R.DARTSCell(H, H, 0.25, 0.75, R.DARTS_V2).cuda()
    x = torch.randn(T, B, H, device="cuda").bfloat16()
    h0 = (torch.randn(1, B, H, device="cuda") * 0.1).bfloat16()
    xm, hm = R.locked_mask(x, B, H, 0.75), R.locked_mask(x, B, H, 0.25)
    outs, grads = {}, {}
    for fused in (False, True):
        for p in cell.parameters():
            p.grad = None
        out = cell(x, h0, masks=(xm, hm), fused=fused)[0]
        out.float().pow(2).mean().backward()
        outs[fused], grads[fused] = out.float(), [p.grad.clone() for p in cell.parameters()]
    assert (outs[True] - outs[False]).abs().max() < 3e-2 * max(outs[False].abs().max().item(), 0.1)
    for a, b in zip(grads[True], grads[False]):
        cos = torch.nn.functional.cosine_similarity(a.flatten(), b.flatten(), dim=0).item()
        assert cos > 0.995, cos
    cell.eval()                                             # no masks: the gate feeds the next GEMM directly
    a, b = cell(x, h0, fused=True)[0].float(), cell(x, h0, fused=False)[0].float()
    assert (a - b).abs().max() < 3e-2 * max(b.abs().max().item(), 0.1)
