"""``bench.py --impl reference``: run the unmodified reference harness on the benchmark config and report the
same JSON line with ``"impl": "reference"``.

Install outcome (recorded in DESIGN.md): ``pip install /root/reference`` fails — the repository root has neither
``setup.py`` nor ``pyproject.toml``. Its Python packages ``common/`` and ``harness/`` install with ``--no-deps``
into ``baseline/_ref``. The harness is written for Python 3.6 / torch 1.4 / Horovod / apex; here it runs on Python
3.12 / torch 2.11 with import stand-ins for absent third-party packages (``baseline/shims``) and no Horovod or apex:

* 1 GPU: runs (``ref_worker.py``), mixed precision through ``torch.autocast`` in the trial's ``train_batch``.
* N > 1: the reference's only data-parallel transport is Horovod. The package cannot be built offline, so
  ``baseline/shims/horovod`` provides the slice of ``horovod.torch`` the reference calls, over NCCL; the unmodified
  reference controller then runs its own ``hvd_config.use == True`` path, one process per GPU under
  ``torch.distributed.run``. The arm is labelled accordingly in ``config.transport``.

Besides the headline BERT block the line carries ``resnet50`` and ``dcgan`` blocks (BASELINE configs #2 and #4).
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import tempfile
from typing import Any, Dict

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
SHIMS = os.path.join(HERE, "shims")


def ensure_installed() -> str:
    """Install common/ and harness/ of the reference into baseline/_ref if missing. Returns '' or the reason it failed."""
    if os.path.isdir(os.path.join(REF, "determined")) and os.path.isdir(os.path.join(REF, "determined_common")):
        return ""
    src = "/root/reference"
    if not os.path.isdir(src):
        return "reference source tree /root/reference not present on this machine and baseline/_ref is empty"
    tmp = tempfile.mkdtemp(prefix="refcopy")
    subprocess.run(["cp", "-r", os.path.join(src, "common"), os.path.join(src, "harness"), tmp], check=True)
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--find-links",
           "/opt/wheelhouse", "--target", REF, os.path.join(tmp, "common"), os.path.join(tmp, "harness")]
    p = subprocess.run(cmd, capture_output=True, text=True)
    if p.returncode != 0:
        return "pip install of reference common/ + harness/ failed: " + p.stderr.strip().splitlines()[-1][:200]
    return ""


def _run_worker(model: str, args, workdir: str, env: Dict[str, str]) -> Dict[str, Any]:
    """One model through the reference worker on this rank; rank 0's dict carries the result line."""
    worker = os.path.join(workdir, "ref_worker.py")
    p = subprocess.run([sys.executable, worker, "--model", model, "--steps", str(args.steps), "--warmup", str(args.warmup)],
                       cwd=workdir, env=env, capture_output=True, text=True)
    line = next((l for l in p.stdout.splitlines() if l.startswith("REF_RESULT ")), None)
    if p.returncode != 0:
        tail = (p.stderr.strip().splitlines() or ["no output"])[-1][:300]
        return {"unavailable": f"reference harness failed on this stack (python {sys.version_info.major}."
                               f"{sys.version_info.minor}, no apex, Horovod stand-in over NCCL): {tail}"}
    return json.loads(line[len("REF_RESULT "):]) if line else {}


def run(args) -> Dict[str, Any]:
    """Every rank launched by torch.distributed.run calls this; each starts ITS rank of the reference job as a subprocess
    (same RANK / WORLD_SIZE / MASTER_* environment) and rank 0 assembles the JSON line."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        return {"impl": "reference", "n_gpus": args.gpus,
                "unavailable": f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run --nproc-per-node {args.gpus}"}
    why = ensure_installed()
    if why:
        return {"impl": "reference", "unavailable": why}
    env = dict(os.environ)
    env["PYTHONPATH"] = SHIMS + os.pathsep + REF
    # start-up only: the image sets PYTHONDONTWRITEBYTECODE and ships no __pycache__, so each of the 3 x N worker processes
    # would compile torch + transformers from source again (minutes at N = 8); a shared byte-code cache changes nothing
    # about the code that runs
    env.pop("PYTHONDONTWRITEBYTECODE", None)
    env.setdefault("PYTHONPYCACHEPREFIX", os.path.join(tempfile.gettempdir(), "det-ref-pycache"))
    if world == 1:
        env.pop("RANK", None), env.pop("WORLD_SIZE", None), env.pop("LOCAL_RANK", None)
    workdir = tempfile.mkdtemp(prefix="ref_arm_")  # outside the repo so `determined` resolves to baseline/_ref
    with open(os.path.join(HERE, "ref_worker.py")) as f, open(os.path.join(workdir, "ref_worker.py"), "w") as g:
        g.write(f.read())
    models = [m for m in getattr(args, "models", "bert,resnet50,dcgan").split(",") if m]
    results = {m: _run_worker(m, args, workdir, env) for m in models}
    if rank != 0:
        return {}
    transport = ("single process (hvd_config.use = False)" if world == 1 else
                 "reference harness hvd path (hvd.DistributedOptimizer / broadcast_parameters / synchronize) on a NCCL "
                 "stand-in for the Horovod package, which cannot be built offline (baseline/shims/horovod)")

    def block(m: str, unit: str, h2d: int, model_desc: str) -> Dict[str, Any]:
        r = results.get(m) or {}
        if "ms" not in r:
            return {"unavailable": r.get("unavailable", "no result line from the reference worker")}
        value = r["samples"] / (r["ms"] / 1e3)
        return {"value": round(value, 2), "unit": unit, "ms_per_step": round(r["ms"] / args.steps, 4),
                "e2e": {"value": round(value, 2), "unit": unit, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4 * r["aggregation_frequency"],
                        "path": "reference PyTorchTrialController.run(): the device-timed number IS the public-API path"},
                "config": {"model": model_desc, "global_batch": r["per_slot_batch"] * world, "per_slot_batch": r["per_slot_batch"],
                           "aggregation_frequency": r["aggregation_frequency"], "parallelism": f"dp{world}", "transport": transport,
                           "mixed_precision": "torch.autocast(bf16) inside train_batch (apex unavailable offline)"},
                "final_loss": r.get("loss")}

    bert = block("bert", "samples/s", 4 * (3 * 48 * 128 * 8 + 48 * 8), "bert-base HF BertForSequenceClassification") if "bert" in models else {}
    if "bert" in models and "value" not in bert:
        return {"impl": "reference", "n_gpus": world, "unavailable": bert["unavailable"]}
    r = results.get("bert") or next(iter(results.values()))
    out: Dict[str, Any] = {
        "metric": "samples/sec per trial (BERT-base GLUE seq128, mixed precision, aggregation_frequency=4)",
        "value": bert.get("value"), "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": bert.get("ms_per_step"), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic (random token ids of the GLUE/MRPC shape, random-init BERT-base weights)",
        "impl": "reference",
        "config": dict(bert.get("config", {}), seq_len=128, samples_per_step=192 * world,
                       harness=f"determined {r.get('det_version')} PyTorchTrialController (unmodified, baseline/_ref)",
                       torch=r.get("torch"), step_definition="one optimizer update = 4 batches"),
        "e2e": bert.get("e2e"), "final_loss": bert.get("final_loss"),
    }
    if "resnet50" in models:
        out["resnet50"] = block("resnet50", "images/s", 128 * (224 * 224 * 3 + 8), "torchvision ResNet-50 (cuDNN), channels_last, SGD momentum")
    if "dcgan" in models:
        out["dcgan"] = block("dcgan", "images/s", 128 * (3 * 64 * 64 * 4 + 8), "DCGAN 64x64 generator + discriminator (torch.nn, cuDNN), two Adam optimizers")
    return out
