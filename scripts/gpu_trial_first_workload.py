"""Why is the first workload of a short trial slow?  Builds the cifar10 trial the ASHA benchmark runs (same hyperparameters),
times workloads of 100 batches each, and profiles the first one on the host (cProfile, top cumulative)."""
import cProfile
import io
import json
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

t_import = time.time()
from determined_b200.models.vision_trials import CIFARTrial  # noqa: E402
from tests.harness_utils import TrainAndValidate, make_controller  # noqa: E402

hp = {"learning_rate": 1e-4, "learning_rate_decay": 1e-6, "layer1_dropout": 0.25, "layer2_dropout": 0.25, "layer3_dropout": 0.5,
      "global_batch_size": 48}
tv = TrainAndValidate()
marks = []


def stream():
    for w in tv.send(steps=4, scheduling_unit=100, validation_freq=2):
        t0 = time.time()
        yield w
        torch.cuda.synchronize()
        marks.append((w[0].kind.name, round(time.time() - t0, 3)))


t0 = time.time()
c = make_controller(CIFARTrial, hp, stream(), exp_config={"data": {"train_size": 4096, "validation_size": 1024}, "records_per_epoch": 4096,
                                                          "min_validation_period": {"epochs": 1}})
t_build = time.time() - t0
prof = cProfile.Profile()
prof.runcall(c.run)
print(json.dumps({"build_controller_s": round(t_build, 3), "workloads": marks}), flush=True)
buf = io.StringIO()
pstats.Stats(prof, stream=buf).sort_stats("cumulative").print_stats(45)
print(buf.getvalue()[:12000])
