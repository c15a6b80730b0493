"""``bench.py --impl reference``: run the unmodified reference harness on the benchmark config and report the
same JSON line with ``"impl": "reference"``.

Install outcome (recorded in DESIGN.md): ``pip install /root/reference`` fails — the repository root has neither
``setup.py`` nor ``pyproject.toml``. Its Python packages ``common/`` and ``harness/`` install with ``--no-deps``
into ``baseline/_ref``. The harness is written for Python 3.6 / torch 1.4 / Horovod / apex; here it runs on Python
3.12 / torch 2.11 with import stand-ins for absent third-party packages (``baseline/shims``) and no Horovod or apex:

* 1 GPU: runs (``ref_worker.py``), mixed precision through ``torch.autocast`` in the trial's ``train_batch``.
* N > 1: the reference's only data-parallel transport is Horovod (``horovodrun`` + NCCL), which cannot be built
  offline → ``unavailable``.
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


def run(args) -> Dict[str, Any]:
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 or world > 1:
        return {"impl": "reference", "n_gpus": args.gpus,
                "unavailable": "reference multi-GPU training requires Horovod (horovodrun + NCCL), which is not "
                               "installable offline; its PyTorchTrialContext has no other data-parallel transport"}
    why = ensure_installed()
    if why:
        return {"impl": "reference", "unavailable": why}
    env = dict(os.environ)
    env["PYTHONPATH"] = SHIMS + os.pathsep + REF
    env.pop("RANK", None), env.pop("WORLD_SIZE", None), env.pop("LOCAL_RANK", None)
    workdir = tempfile.mkdtemp(prefix="ref_arm_")  # outside the repo so `determined` resolves to baseline/_ref
    worker = os.path.join(workdir, "ref_worker.py")
    with open(os.path.join(HERE, "ref_worker.py")) as f, open(worker, "w") as g:
        g.write(f.read())
    p = subprocess.run([sys.executable, worker, "--steps", str(args.steps), "--warmup", str(args.warmup)],
                       cwd=workdir, env=env, capture_output=True, text=True)
    line = next((l for l in p.stdout.splitlines() if l.startswith("REF_RESULT ")), None)
    if p.returncode != 0 or line is None:
        tail = (p.stderr.strip().splitlines() or ["no output"])[-1][:300]
        return {"impl": "reference", "unavailable": f"reference harness failed on this stack (python {sys.version_info.major}."
                                                     f"{sys.version_info.minor}, no apex/Horovod): {tail}"}
    r = json.loads(line[len("REF_RESULT "):])
    value = r["samples"] / (r["ms"] / 1e3)
    return {
        "metric": "samples/sec per trial (BERT-base GLUE seq128, mixed precision, aggregation_frequency=4)",
        "value": round(value, 2), "unit": "samples/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(r["ms"] / args.steps, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic (random token ids of the GLUE/MRPC shape, random-init BERT-base weights)",
        "impl": "reference",
        "config": {"model": "bert-base HF BertForSequenceClassification", "global_batch": 48, "per_slot_batch": 48,
                   "seq_len": 128, "aggregation_frequency": 4, "samples_per_step": 192, "parallelism": "dp1",
                   "mixed_precision": "torch.autocast(bf16) inside train_batch (apex unavailable offline)",
                   "harness": f"determined {r['det_version']} PyTorchTrialController (unmodified, baseline/_ref)",
                   "torch": r["torch"], "step_definition": "one optimizer update = 4 batches"},
        "e2e": {"value": round(value, 2), "unit": "samples/s", "h2d_bytes_per_step": 4 * (3 * 48 * 128 * 8 + 48 * 8),
                "d2h_bytes_per_step": 16, "path": "reference PyTorchTrialController.run(): the device-timed number IS the public-API path"},
        "final_loss": r["loss"],
    }
