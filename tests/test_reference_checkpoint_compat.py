"""Checkpoints interchange with the UNMODIFIED reference harness in both directions: a checkpoint written by the
reference's ``PyTorchTrialController._save`` resumes under ours, and one written by ours resumes under the reference —
same model definition file (it only imports ``determined.pytorch``), CPU, no GPU code involved. Skipped when the
reference packages are not installed under ``baseline/_ref`` (they are git-ignored; see DESIGN.md "Reference arm")."""
import json
import os
import pathlib
import subprocess
import sys

import pytest
import torch

from tests.harness_utils import TrainAndValidate, make_controller

ROOT = pathlib.Path(__file__).resolve().parent.parent
REF = ROOT / "baseline" / "_ref"
MODEL_DIR = ROOT / "tests" / "fixtures" / "xor_model_def"
HP = {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5}

pytestmark = pytest.mark.skipif(not (REF / "determined").is_dir(), reason="reference harness not installed in baseline/_ref")


def run_reference(mode, ckpt, out, tmp_path):
    env = dict(os.environ, PYTHONPATH=f"{ROOT / 'baseline' / 'shims'}:{REF}", CUDA_VISIBLE_DEVICES="")
    work = tmp_path / "work"          # the reference copies its working directory into <checkpoint>/code: keep them apart
    work.mkdir(exist_ok=True)
    p = subprocess.run([sys.executable, str(ROOT / "baseline" / "ref_checkpoint_worker.py"), mode, str(MODEL_DIR), str(ckpt), str(out)],
                       cwd=str(work), env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0 and "REF_CKPT_OK" in p.stdout, p.stdout[-2000:] + p.stderr[-4000:]
    return json.loads(out.read_text())


def our_trial_class():
    sys.path.insert(0, str(MODEL_DIR))
    try:
        sys.modules.pop("model_def", None)
        import model_def
        return model_def.XorTrial
    finally:
        sys.path.remove(str(MODEL_DIR))


def flat(controller):
    return torch.cat([p.detach().reshape(-1) for m in controller.context.models for p in m.parameters()]).tolist()


def test_reference_checkpoint_resumes_under_this_framework(tmp_path):
    ckpt = tmp_path / "ckpts" / "ref_ckpt"
    rep = run_reference("save", ckpt, tmp_path / "save.json", tmp_path)
    assert (ckpt / "state_dict.pth").exists() and (ckpt / "code").is_dir()
    tv = TrainAndValidate()
    c = make_controller(our_trial_class(), HP, tv.send(steps=1, validation_freq=1, scheduling_unit=10, initial_step_id=3, batches_done=20),
                        load_path=ckpt, initial_batches=20)
    assert flat(c) == pytest.approx(rep["final_weights"], abs=1e-7), "weights restored from the reference's state_dict.pth"
    c.run()
    assert len(tv.training_metrics) == 10 and tv.training_metrics[-1]["loss"] == tv.training_metrics[-1]["loss"]


def test_our_checkpoint_resumes_under_the_reference(tmp_path):
    ckpt = tmp_path / "ckpts" / "our_ckpt"
    tv = TrainAndValidate()
    c = make_controller(our_trial_class(), HP, tv.send(steps=2, validation_freq=2, scheduling_unit=10, checkpoint_dir=ckpt))
    c.run()
    ours = flat(c)
    assert (ckpt / "state_dict.pth").exists()
    rep = run_reference("load", ckpt, tmp_path / "load.json", tmp_path)
    assert rep["loaded_weights"] == pytest.approx(ours, abs=1e-7), "the reference restored our weights"
    assert len(rep["losses"]) == 1 and rep["losses"][0] == rep["losses"][0]
