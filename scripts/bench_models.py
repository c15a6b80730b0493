#!/usr/bin/env python
"""Secondary benchmarks (BASELINE configs #2 and #4): ResNet-50 / DCGAN / CIFAR / MNIST trials driven end to end through
the PyTorchTrialController (pinned host batch -> device copy every batch, metrics read back per workload).

    python scripts/bench_models.py --model resnet50 --batch 128 --steps 20 --warmup 5
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/bench_models.py --model resnet50 --gpus 8

`--impl torch` runs the same trial shape on stock torch / torchvision modules (bf16 autocast + channels_last when
`--amp`, else fp32) inside the same controller, as the "what the reference's user code would get from cuDNN" arm.
One step = one batch = one optimizer update. Prints one JSON line on rank 0 (max over ranks, device-timed)."""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from bench import ClockSampler, _barrier_sync, _dist_env, _max_over_ranks  # noqa: E402
from determined_b200 import _execution, load  # noqa: E402
from determined_b200 import pytorch as dp  # noqa: E402
from determined_b200 import workload as wl  # noqa: E402
from determined_b200.models import darts_trial as DT
from determined_b200.models import darts_rnn_trial as DR
from determined_b200.models import detr_trial as DETR
from determined_b200.models import squad_trials as ST
from determined_b200.models import vision_trials as VT  # noqa: E402
from determined_b200.ops import _native  # noqa: E402


class TorchResNet50Trial(VT.ResNet50Trial):
    """Same recipe on torchvision's ResNet-50 + torch.optim.SGD (cuDNN / cuBLAS kernels)."""

    def __init__(self, context: dp.PyTorchTrialContext) -> None:
        import torchvision

        self.context = context
        hp = context.get_hparams()
        self.amp = bool(hp.get("amp", True))
        self.shape, self.num_classes = (3, 224, 224), 1000
        net = torchvision.models.resnet50()
        if self.amp:
            net = net.to(memory_format=torch.channels_last)
        self.model = context.wrap_model(net)
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4),
                                                fused=False)

    def train_batch(self, batch, epoch_idx, batch_idx):
        data, labels = batch
        if data.dtype == torch.uint8:   # same uint8 NHWC bytes as our arm, normalised with stock torch ops
            mean = torch.tensor(VT.IMAGENET_MEAN, device=data.device).view(1, 3, 1, 1)
            std = torch.tensor(VT.IMAGENET_STD, device=data.device).view(1, 3, 1, 1)
            data = (data.permute(0, 3, 1, 2).float() / 255.0 - mean) / std
        if self.amp:
            data = data.contiguous(memory_format=torch.channels_last)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.amp):
            loss = torch.nn.functional.cross_entropy(self.model(data), labels)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss.detach()}


MODELS = {
    "resnet50": (VT.ResNet50Trial, {"learning_rate": 0.1, "use_cuda_graphs": True}, (3, 224, 224), "images/s"),
    "dcgan": (VT.DCGANTrial, {"latent_dim": 128, "use_cuda_graphs": True}, (3, 64, 64), "images/s"),
    "cifar": (VT.CIFARTrial, {"learning_rate": 1e-4}, (3, 32, 32), "images/s"),
    "mnist": (VT.MNistTrial, {"learning_rate": 1.0}, (1, 28, 28), "images/s"),
    # DARTS CIFAR-10 evaluation network (20 cells, separable / dilated convolutions), reference batch 96
    "darts": (DT.DARTSCNNTrial, {"learning_rate": 0.025, "init_channels": 32, "layers": 20, "use_cuda_graphs": True}, (3, 32, 32), "images/s"),
    # DARTS recurrent cell language model at the published sizes (850 wide, vocabulary 10,000, 35-token windows, batch 64)
    "darts_rnn": (DR.DARTSRNNTrial, dict({f"node{i + 1}_edge": e for i, e in enumerate([0, 1, 1, 1, 2, 5, 3, 5])},
                                         **{f"node{i + 1}_op": o for i, o in enumerate(["sigmoid", "relu", "relu", "identity", "tanh", "sigmoid", "tanh", "relu"])},
                                         learning_rate=20.0, fixed_seq_len=True, use_cuda_graphs=True, train_tokens=400_000),
                  (1, 1, 35), "windows/s"),
    # DETR (ResNet-50 trunk, 6 + 6 transformer layers, 100 queries, Hungarian matching on the host) on padded ~512 px scenes
    "detr": (DETR.DETRTrial, {"lr": 1e-4, "lr_backbone": 1e-5, "image_size": 512, "num_classes": 91, "dataset_size": 512}, (3, 512, 512), "images/s"),
    # BERT-base span model at the SQuAD shape (seq 384): the tiled attention kernels + span head
    "bert_squad": (ST.BertSquadTrial, {"learning_rate": 3e-5, "max_seq_length": 384, "use_cuda_graphs": True}, (1, 1, 384), "samples/s"),
    # ALBERT-xxlarge sizes (hidden 4096, 64 heads, one shared layer applied 12 times)
    "albert_xxlarge_squad": (ST.AlbertSquadTrial, {"learning_rate": 5e-5, "max_seq_length": 384, "hidden_size": 4096,
                                                   "num_attention_heads": 64, "intermediate_size": 16384}, (1, 1, 384), "samples/s"),
}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet50", choices=sorted(MODELS))
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--batch", type=int, default=128, help="per-slot batch size")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "torch"])
    ap.add_argument("--amp", type=int, default=1)
    ap.add_argument("--graphs", type=int, default=1)
    args = ap.parse_args()
    rank, local_rank, world = _dist_env()
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    trial_cls, hp, shape, unit = MODELS[args.model]
    hp = dict(hp, global_batch_size=args.batch * world)
    if not args.graphs:
        hp.pop("use_cuda_graphs", None)
    if args.impl == "torch":
        if args.model != "resnet50":
            raise SystemExit("--impl torch is implemented for resnet50")
        trial_cls, hp = TorchResNet50Trial, dict(hp, amp=bool(args.amp))
        hp.pop("use_cuda_graphs", None)
    cfg = {"description": f"{args.model}_bench", "hyperparameters": hp,
           "data": {"train_size": max(4 * args.batch * world, 512), "validation_size": args.batch * world},
           "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 1000}},
           "resources": {"slots_per_trial": world}}
    env, dist_ctx = _execution.make_local_execution_env(True, cfg, hparams=hp, seed=7)
    env.experiment_config["resources"]["slots_per_trial"] = world
    env._per_slot_batch_size, env._global_batch_size = env._calculate_batch_sizes()
    _native.load()
    timing = {}
    sampler = ClockSampler(local_rank) if rank == 0 else None
    box = {}

    def launches() -> int:
        g = getattr(box.get("c"), "_graphed", None)
        return _native.launch_count + (g.replayed_launches if g is not None else 0)

    def workloads():
        icpt = wl.WorkloadResponseInterceptor()
        total = 0
        for i in range(args.warmup + 6):   # graph capture needs a few eager batches first
            yield from icpt.send(wl.train_workload(i + 1, num_batches=1, total_batches_processed=total), [])
            total += 1
        _barrier_sync(device)
        if sampler:
            sampler.start()
        n0 = launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        losses = []
        for i in range(args.steps):
            yield from icpt.send(wl.train_workload(args.warmup + 7 + i, num_batches=1, total_batches_processed=total), [])
            total += 1
            r = icpt.result()
            if not isinstance(r, wl.Skipped):
                losses.append(r["metrics"]["avg_metrics"]["loss"])
        e1.record()
        _barrier_sync(device)
        timing["ms"] = e0.elapsed_time(e1)
        timing["launches"] = launches() - n0
        timing["losses"] = losses
        yield from icpt.send(wl.terminate_workload(), [])

    controller = load.load_controller_from_trial(trial_cls, env, workloads(), None, None, dist_ctx)
    box["c"] = controller
    controller.run()
    clocks = sampler.stop() if sampler else None
    ms = _max_over_ranks(timing["ms"], device)
    per_step = args.batch * world
    n_in = shape[0] * shape[1] * shape[2]
    out = {"metric": f"{unit} per trial ({args.model}, synthetic {shape[1]}x{shape[2]} batches, end to end through PyTorchTrialController)",
           "value": round(per_step * args.steps / (ms / 1e3), 1), "unit": unit, "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True, "scaling": "weak",
           "dtype": "bf16" if (args.impl == "ours" or args.amp) else "fp32", "data": "synthetic", "impl": args.impl,
           "config": {"model": args.model, "global_batch": per_step, "per_slot_batch": args.batch, "parallelism": f"dp{world}",
                      "cuda_graphs": bool(getattr(controller, "_graphed", None))},
           "e2e": {"h2d_bytes_per_step": args.batch * (3 * 384 * 8 + 16 if "squad" in args.model else n_in * (1 if args.model == "resnet50" else 4) + 8),
                   "d2h_bytes_per_step": 4},
           "gpu_launches": int(timing["launches"]), "clocks": clocks,
           "first_loss": timing["losses"][0] if timing["losses"] else None,
           "final_loss": timing["losses"][-1] if timing["losses"] else None}
    if rank == 0:
        print(json.dumps(out), flush=True)
    try:
        import torch.distributed as dist

        if dist.is_initialized():
            dist.destroy_process_group()
    except Exception:
        pass


if __name__ == "__main__":
    main()
