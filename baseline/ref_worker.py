"""Reference arm worker: runs the UNMODIFIED reference harness (``baseline/_ref``: determined 0.13.10.dev0) through
its own public API — a ``determined.pytorch.PyTorchTrial`` driven by the reference's ``PyTorchTrialController`` over
RUN_STEP workloads — on the same model/config as ``bench.py``'s own arm.

Executed as a subprocess by ``baseline/reference_arm.py`` with ``PYTHONPATH=baseline/shims:baseline/_ref`` and a
working directory outside the repository, so nothing of determined_b200 is importable here. Models are stock
``torch.nn`` / ``transformers`` / ``torchvision`` modules (cuBLAS GEMMs, cuDNN convolutions, torch SDPA), stock
``torch.optim`` optimizers, ``torch.nn.utils.clip_grad_norm_``. apex is not installable offline, so mixed precision is
``torch.autocast(bfloat16)`` inside ``train_batch`` (the reference's ``configure_apex_amp`` path is unavailable).

The ``EnvContext`` / ``HorovodContext`` are built the way the reference's own entry point builds them from the master's
task spec (``harness/determined/exec/harness.py``), with the FULL experiment config: the local-test helper
(``_execution._make_local_execution_env``) silently drops the ``optimizations`` and ``resources`` sections and would run
``aggregation_frequency: 1`` on one slot whatever the config says.

* 1 rank: ``hvd_config.use == False`` — the controller's single-process path.
* N ranks (one process per GPU under ``torch.distributed.run``): ``resources.slots_per_trial = N`` makes
  ``HorovodContext.from_configs`` choose ``use == True``; the reference then calls ``hvd.init``,
  ``hvd.DistributedOptimizer``, ``hvd.broadcast_parameters`` / ``broadcast_optimizer_state``, ``optimizer.synchronize`` and
  its own chief <-> worker ZMQ channel. ``horovod.torch`` resolves to ``baseline/shims/horovod`` — a stand-in over NCCL
  for the third-party package that cannot be built offline.
"""
import det_ref_compat  # noqa: F401  (must precede the reference imports)

import argparse
import json
import os
import time

import torch

import determined as det
from determined import _execution, constants, horovod, load, workload
from determined.pytorch import DataLoader, LRScheduler, PyTorchTrial, PyTorchTrialContext

SEQ_LEN = 128
TINY = os.environ.get("REF_TINY") == "1"      # CPU smoke test of the wiring (gloo, toy sizes)


class SyntheticGlue(torch.utils.data.Dataset):
    def __init__(self, size, seq_len, vocab, seed):
        g = torch.Generator().manual_seed(seed)
        self.ids = torch.randint(1000 if vocab > 2000 else 1, vocab, (size, seq_len), generator=g, dtype=torch.int64)
        lengths = torch.randint(seq_len // 2, seq_len + 1, (size,), generator=g)
        pos = torch.arange(seq_len)[None, :]
        self.mask = (pos < lengths[:, None]).to(torch.int64)
        self.ids = self.ids * self.mask
        first = torch.randint(seq_len // 4, seq_len // 2, (size,), generator=g)
        self.tt = ((pos >= first[:, None]) & (pos < lengths[:, None])).to(torch.int64)
        self.labels = torch.randint(0, 2, (size,), generator=g, dtype=torch.int64)

    def __len__(self):
        return self.ids.shape[0]

    def __getitem__(self, i):
        return self.ids[i], self.mask[i], self.tt[i], self.labels[i]


def _autocast():
    return torch.autocast("cuda", dtype=torch.bfloat16, enabled=torch.cuda.is_available())


class RefBertTrial(PyTorchTrial):
    def __init__(self, context: PyTorchTrialContext) -> None:
        from transformers import BertConfig, BertForSequenceClassification

        self.context = context
        cfg = BertConfig(num_labels=2)  # bert-base-uncased architecture, random init
        if TINY:
            cfg = BertConfig(num_labels=2, hidden_size=64, num_hidden_layers=2, num_attention_heads=2, intermediate_size=128,
                             vocab_size=512)
        self.vocab = cfg.vocab_size
        self.model = context.wrap_model(BertForSequenceClassification(cfg))
        no_decay = ["bias", "LayerNorm.weight"]
        groups = [
            {"params": [p for n, p in self.model.named_parameters() if not any(nd in n for nd in no_decay)],
             "weight_decay": 0.0},
            {"params": [p for n, p in self.model.named_parameters() if any(nd in n for nd in no_decay)],
             "weight_decay": 0.0},
        ]
        self.optimizer = context.wrap_optimizer(torch.optim.AdamW(groups, lr=2e-5, eps=1e-8))
        sched = torch.optim.lr_scheduler.LambdaLR(self.optimizer, lambda s: max(0.0, (100000 - s) / 100000.0))
        self.lr_scheduler = context.wrap_lr_scheduler(sched, LRScheduler.StepMode.STEP_EVERY_BATCH)

    def build_training_data_loader(self):
        return DataLoader(SyntheticGlue(512 if TINY else 4096, SEQ_LEN, self.vocab, 1),
                          batch_size=self.context.get_per_slot_batch_size(), shuffle=True, drop_last=True,
                          pin_memory=torch.cuda.is_available())

    def build_validation_data_loader(self):
        return DataLoader(SyntheticGlue(96, SEQ_LEN, self.vocab, 2), batch_size=self.context.get_per_slot_batch_size(),
                          drop_last=True, pin_memory=torch.cuda.is_available())

    def train_batch(self, batch, epoch_idx, batch_idx):
        ids, mask, tt, labels = batch
        with _autocast():
            out = self.model(input_ids=ids, attention_mask=mask, token_type_ids=tt, labels=labels)
        loss = out[0]
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer,
                                    clip_grads=lambda params: torch.nn.utils.clip_grad_norm_(params, 1.0))
        return {"loss": loss}

    def evaluate_batch(self, batch):
        ids, mask, tt, labels = batch
        with _autocast():
            out = self.model(input_ids=ids, attention_mask=mask, token_type_ids=tt, labels=labels)
        return {"validation_loss": out[0], "acc": (out[1].argmax(-1) == labels).float().mean()}


class SyntheticImages(torch.utils.data.Dataset):
    """uint8 NHWC bytes (what a JPEG decoder hands over) or float NCHW images, plus labels."""

    def __init__(self, size, shape, num_classes, seed, uint8_nhwc):
        g = torch.Generator().manual_seed(seed)
        C, H, W = shape
        self.labels = torch.randint(0, num_classes, (size,), generator=g, dtype=torch.int64)
        if uint8_nhwc:
            self.images = torch.randint(0, 256, (size, H, W, C), generator=g, dtype=torch.uint8)
        else:
            self.images = torch.randn(size, C, H, W, generator=g) * 0.5

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        return self.images[i], self.labels[i]


class RefResNet50Trial(PyTorchTrial):
    """torchvision ResNet-50 (cuDNN), bf16 autocast + channels_last, SGD momentum 0.9 — the recipe of bench.py's resnet50 block."""

    def __init__(self, context: PyTorchTrialContext) -> None:
        import torchvision

        self.context = context
        self.res = 32 if TINY else 224
        net = torchvision.models.resnet18(num_classes=10) if TINY else torchvision.models.resnet50()
        if torch.cuda.is_available():
            net = net.to(memory_format=torch.channels_last)
        self.model = context.wrap_model(net)
        self.optimizer = context.wrap_optimizer(torch.optim.SGD(self.model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4))
        self.lr_scheduler = context.wrap_lr_scheduler(torch.optim.lr_scheduler.StepLR(self.optimizer, step_size=30, gamma=0.1),
                                                      LRScheduler.StepMode.STEP_EVERY_EPOCH)
        self.mean = torch.tensor((0.485, 0.456, 0.406)).view(1, 3, 1, 1)
        self.std = torch.tensor((0.229, 0.224, 0.225)).view(1, 3, 1, 1)

    def build_training_data_loader(self):
        b = self.context.get_per_slot_batch_size()
        return DataLoader(SyntheticImages(max(4 * b, 64), (3, self.res, self.res), 10 if TINY else 1000, 1, True), batch_size=b,
                          shuffle=True, drop_last=True, pin_memory=torch.cuda.is_available())

    def build_validation_data_loader(self):
        b = self.context.get_per_slot_batch_size()
        return DataLoader(SyntheticImages(b, (3, self.res, self.res), 10 if TINY else 1000, 2, True), batch_size=b, drop_last=True)

    def _inputs(self, data):
        x = (data.permute(0, 3, 1, 2).float() / 255.0 - self.mean.to(data.device)) / self.std.to(data.device)
        return x.contiguous(memory_format=torch.channels_last) if x.is_cuda else x

    def train_batch(self, batch, epoch_idx, batch_idx):
        data, labels = batch
        with _autocast():
            loss = torch.nn.functional.cross_entropy(self.model(self._inputs(data)), labels)
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer)
        return {"loss": loss}

    def evaluate_batch(self, batch):
        data, labels = batch
        with _autocast():
            loss = torch.nn.functional.cross_entropy(self.model(self._inputs(data)), labels)
        return {"validation_loss": loss}


def _dcgan_nets(nz, ngf, ndf, nc):
    nn = torch.nn
    g = nn.Sequential(
        nn.ConvTranspose2d(nz, ngf * 8, 4, 1, 0, bias=False), nn.BatchNorm2d(ngf * 8), nn.ReLU(True),
        nn.ConvTranspose2d(ngf * 8, ngf * 4, 4, 2, 1, bias=False), nn.BatchNorm2d(ngf * 4), nn.ReLU(True),
        nn.ConvTranspose2d(ngf * 4, ngf * 2, 4, 2, 1, bias=False), nn.BatchNorm2d(ngf * 2), nn.ReLU(True),
        nn.ConvTranspose2d(ngf * 2, ngf, 4, 2, 1, bias=False), nn.BatchNorm2d(ngf), nn.ReLU(True),
        nn.ConvTranspose2d(ngf, nc, 4, 2, 1, bias=False), nn.Tanh())
    d = nn.Sequential(
        nn.Conv2d(nc, ndf, 4, 2, 1, bias=False), nn.LeakyReLU(0.2, True),
        nn.Conv2d(ndf, ndf * 2, 4, 2, 1, bias=False), nn.BatchNorm2d(ndf * 2), nn.LeakyReLU(0.2, True),
        nn.Conv2d(ndf * 2, ndf * 4, 4, 2, 1, bias=False), nn.BatchNorm2d(ndf * 4), nn.LeakyReLU(0.2, True),
        nn.Conv2d(ndf * 4, ndf * 8, 4, 2, 1, bias=False), nn.BatchNorm2d(ndf * 8), nn.LeakyReLU(0.2, True),
        nn.Conv2d(ndf * 8, 1, 4, 1, 0, bias=False))
    for m in list(g.modules()) + list(d.modules()):
        if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)):
            nn.init.normal_(m.weight, 0.0, 0.02)
        elif isinstance(m, nn.BatchNorm2d):
            nn.init.normal_(m.weight, 1.0, 0.02)
            nn.init.zeros_(m.bias)
    return g, d


def _bce_logits(logits, target):
    return torch.nn.functional.binary_cross_entropy_with_logits(logits.float(), torch.full_like(logits, target, dtype=torch.float32))


class RefDCGANTrial(PyTorchTrial):
    """The multi-model / multi-optimizer pattern of the reference's gan_mnist_pytorch example
    (examples/gan/gan_mnist_pytorch/model_def.py:74-184) on the convolutional DCGAN pair, stock torch.nn."""

    def __init__(self, context: PyTorchTrialContext) -> None:
        self.context = context
        self.nz = 128
        ngf = ndf = 16 if TINY else 64
        g, d = _dcgan_nets(self.nz, ngf, ndf, 3)
        if torch.cuda.is_available():
            g, d = g.to(memory_format=torch.channels_last), d.to(memory_format=torch.channels_last)
        self.generator = context.wrap_model(g)
        self.discriminator = context.wrap_model(d)
        self.opt_g = context.wrap_optimizer(torch.optim.Adam(self.generator.parameters(), lr=2e-4, betas=(0.5, 0.999)))
        self.opt_d = context.wrap_optimizer(torch.optim.Adam(self.discriminator.parameters(), lr=2e-4, betas=(0.5, 0.999)))
        self.lr_g = context.wrap_lr_scheduler(torch.optim.lr_scheduler.LambdaLR(self.opt_g, lr_lambda=lambda e: 0.95 ** e),
                                              LRScheduler.StepMode.STEP_EVERY_EPOCH)

    def build_training_data_loader(self):
        b = self.context.get_per_slot_batch_size()
        return DataLoader(SyntheticImages(max(4 * b, 64), (3, 64, 64), 10, 1, False), batch_size=b, shuffle=True, drop_last=True,
                          pin_memory=torch.cuda.is_available())

    def build_validation_data_loader(self):
        b = self.context.get_per_slot_batch_size()
        return DataLoader(SyntheticImages(b, (3, 64, 64), 10, 2, False), batch_size=b, drop_last=True)

    def train_batch(self, batch, epoch_idx, batch_idx):
        imgs, _ = batch
        n = imgs.shape[0]
        self.generator.requires_grad_(True)
        self.discriminator.requires_grad_(False)
        with _autocast():
            fake = self.generator(torch.randn(n, self.nz, 1, 1, device=imgs.device))
            g_loss = _bce_logits(self.discriminator(fake).reshape(n, 1), 1.0)
        self.context.backward(g_loss)
        self.context.step_optimizer(self.opt_g)
        self.generator.requires_grad_(False)
        self.discriminator.requires_grad_(True)
        with _autocast():
            real_loss = _bce_logits(self.discriminator(imgs).reshape(n, 1), 1.0)
            fake_loss = _bce_logits(self.discriminator(fake.detach()).reshape(n, 1), 0.0)
            d_loss = (real_loss + fake_loss) / 2
        self.context.backward(d_loss)
        self.context.step_optimizer(self.opt_d)
        return {"loss": d_loss, "g_loss": g_loss}

    def evaluate_batch(self, batch):
        imgs, _ = batch
        with _autocast():
            return {"loss": _bce_logits(self.discriminator(imgs).reshape(imgs.shape[0], 1), 1.0)}


MODELS = {
    # name: (trial class, per-slot batch, aggregation_frequency, samples unit)
    "bert": (RefBertTrial, 48, 4),
    "resnet50": (RefResNet50Trial, 128, 1),
    "dcgan": (RefDCGANTrial, 128, 1),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="bert", choices=sorted(MODELS))
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=0)
    args = ap.parse_args()
    W, K = max(args.warmup, 3), args.steps
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    trial_cls, per_slot, agg = MODELS[args.model]
    if args.batch:
        per_slot = args.batch
    if TINY:
        per_slot = 8
    config = det.ExperimentConfig({
        "description": f"reference_arm_{args.model}",
        "hyperparameters": {"global_batch_size": per_slot * world},
        "searcher": {"name": "single", "metric": "acc", "smaller_is_better": False, "max_length": {"batches": 400}},
        "scheduling_unit": constants.DEFAULT_SCHEDULING_UNIT,
        "resources": {"slots_per_trial": world, "native_parallel": False},
        "optimizations": {"aggregation_frequency": agg, "average_aggregated_gradients": True,
                          "average_training_metrics": False, "gradient_compression": False, "mixed_precision": "O0",
                          "tensor_fusion_threshold": 64, "tensor_fusion_cycle_time": 5, "auto_tune_tensor_fusion": False},
        "reproducibility": {"experiment_seed": 1234},
    })
    use_gpu, container_gpus, slot_ids = _execution._get_gpus(world)
    env = det.EnvContext(
        master_addr="", master_port=0, use_tls=False, master_cert_file=None, master_cert_name=None, container_id="",
        experiment_config=config, hparams=config["hyperparameters"],
        initial_workload=workload.train_workload(1, 1, 1, config.scheduling_unit()), latest_checkpoint=None,
        use_gpu=use_gpu, container_gpus=container_gpus, slot_ids=slot_ids, debug=False, workload_manager_type="",
        det_rendezvous_ports=f"{constants.LOCAL_RENDEZVOUS_PORT},{constants.LOCAL_RENDEZVOUS_PORT + 1}",
        det_trial_unique_port_offset=int(os.environ.get("REF_PORT_OFFSET", "0")),
        det_trial_runner_network_interface=constants.AUTO_DETECT_TRIAL_RUNNER_NETWORK_INTERFACE,
        det_trial_id="", det_experiment_id="", det_cluster_id="", trial_seed=config.experiment_seed(), managed_training=True)
    ports = env.rendezvous_ports()
    rendezvous_info = det.RendezvousInfo(addrs=[f"127.0.0.1:{ports[0]}"], addrs2=[f"127.0.0.1:{ports[1]}"], rank=0)
    hvd_config = horovod.HorovodContext.from_configs(env.experiment_config, rendezvous_info, env.hparams)
    assert hvd_config.use == (world > 1) and hvd_config.aggregation_frequency == agg
    timing = {}
    cuda = torch.cuda.is_available()

    def sync():
        if cuda:
            torch.cuda.synchronize()

    def workloads():
        icpt = workload.WorkloadResponseInterceptor()
        total = 0
        for i in range(W):
            yield from icpt.send(workload.train_workload(i + 1, num_batches=agg, total_batches_processed=total), [])
            total += agg
        sync()
        if world > 1:
            torch.distributed.barrier()
        if cuda:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        t0 = time.perf_counter()
        losses = []
        for i in range(K):
            yield from icpt.send(workload.train_workload(W + i + 1, num_batches=agg, total_batches_processed=total), [])
            total += agg
            r = icpt.result()
            if isinstance(r, dict):
                losses.append(r["metrics"]["avg_metrics"]["loss"])
        if cuda:
            e1.record()
        sync()
        timing["wall_ms"] = (time.perf_counter() - t0) * 1e3
        timing["ms"] = e0.elapsed_time(e1) if cuda else timing["wall_ms"]
        timing["loss"] = float(losses[-1]) if losses else None
        yield from icpt.send(workload.terminate_workload(), [])

    controller = load.load_controller_from_trial(trial_cls, env, workloads(), None, rendezvous_info, hvd_config)
    controller.run()
    ms = timing["ms"]
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda" if cuda else "cpu")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = float(t.item())
    if rank == 0:
        samples = per_slot * agg * K * world
        print("REF_RESULT " + json.dumps({"model": args.model, "ms": ms, "wall_ms": timing["wall_ms"], "samples": samples,
                                          "per_slot_batch": per_slot, "aggregation_frequency": agg, "world": world,
                                          "hvd_use": bool(hvd_config.use), "loss": timing["loss"],
                                          "det_version": det.__version__, "torch": torch.__version__}), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
