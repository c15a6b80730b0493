"""Drive every model-zoo trial through the PyTorchTrialController on one GPU for a few workloads: losses must be
finite (and fall for the classifiers), validation must produce the searcher metric, a checkpoint must round-trip."""
import json
import os
import pathlib
import sys
import tempfile
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from determined_b200.models import vision_trials as VT  # noqa: E402
from tests.harness_utils import TrainAndValidate, make_controller  # noqa: E402

CASES = [
    ("mnist", VT.MNistTrial, {"global_batch_size": 64, "learning_rate": 1.0}, {"train_size": 2048, "validation_size": 256}, 6, 8),
    ("cifar", VT.CIFARTrial, {"global_batch_size": 64, "learning_rate": 1e-3}, {"train_size": 2048, "validation_size": 256}, 6, 8),
    ("gan_mnist", VT.GANTrial, {"global_batch_size": 64, "latent_dim": 128}, {"train_size": 1024, "validation_size": 128}, 3, 8),
    ("dcgan", VT.DCGANTrial, {"global_batch_size": 64, "latent_dim": 128}, {"train_size": 512, "validation_size": 128}, 3, 4),
    ("resnet50", VT.ResNet50Trial, {"global_batch_size": 32, "learning_rate": 0.002}, {"train_size": 256, "validation_size": 64}, 4, 4),
    ("resnet50_graphs", VT.ResNet50Trial, {"global_batch_size": 32, "learning_rate": 0.002, "use_cuda_graphs": True},
     {"train_size": 256, "validation_size": 64}, 6, 4),
]
bad = []
summary = {}
for name, cls, hp, data, steps, unit in CASES:
    if len(sys.argv) > 1 and name not in sys.argv[1:]:
        continue
    try:
        tv = TrainAndValidate()
        with tempfile.TemporaryDirectory() as td:
            t0 = time.time()
            c = make_controller(cls, hp, tv.send(steps=steps, validation_freq=steps, scheduling_unit=unit,
                                                 checkpoint_dir=pathlib.Path(td) / "ckpt"), exp_config={"data": data})
            c.run()
            dt = time.time() - t0
            ok_ckpt = (pathlib.Path(td) / "ckpt" / "state_dict.pth").exists()
        losses = [m["loss"] for m in tv.training_metrics]
        finite = all(l == l and abs(l) < 1e6 for l in losses)
        first, last = sum(losses[:unit]) / unit, sum(losses[-unit:]) / unit
        summary[name] = {"first": round(first, 4), "last": round(last, 4), "val": tv.validation_metrics[-1], "secs": round(dt, 1)}
        print(name, json.dumps(summary[name], default=float), flush=True)
        ok = finite and ok_ckpt
        if name in ("mnist", "cifar", "resnet50", "resnet50_graphs"):
            ok = ok and last < first
        if not ok:
            bad.append(name)
    except Exception as e:  # noqa
        import traceback
        traceback.print_exc()
        bad.append(name)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(summary, open(os.path.join(ROOT, "gpurun_out", "models_check.json"), "w"), indent=1, default=float)
print("MODELS:", "all ok" if not bad else f"FAILED {bad}")
sys.exit(1 if bad else 0)
