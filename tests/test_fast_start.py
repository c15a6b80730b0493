"""Trial start latency helpers (utils/fast_start.py, exec/warm_worker.py): the dynamo-shim bypass, the CPU thread share, and
the warm worker's adopt / release of a trial's environment, working directory, import path and modules."""
import os
import subprocess
import sys
import textwrap

import pytest
import torch

from determined_b200.exec import warm_worker
from determined_b200.utils import fast_start

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_optimizer_construction_does_not_import_dynamo():
    """A fresh interpreter: build an optimizer, zero its grads, take its state dict — torch._dynamo must stay unloaded
    (it costs 5-7 s of imports at every trial start otherwise); without the bypass the same code loads it."""
    prog = textwrap.dedent('''
        import sys, torch
        if {bypass}:
            from determined_b200.utils.fast_start import skip_dynamo_shims
            assert skip_dynamo_shims() == 4
        p = torch.nn.Parameter(torch.zeros(3))
        opt = torch.optim.SGD([p], lr=0.1)
        p.grad = torch.ones(3)
        opt.zero_grad()
        opt.load_state_dict(opt.state_dict())
        print("dynamo" if "torch._dynamo" in sys.modules else "clean")
    ''')
    env = dict(os.environ, PYTHONPATH=ROOT)
    with_bypass = subprocess.run([sys.executable, "-c", prog.format(bypass=True)], capture_output=True, text=True, env=env, timeout=300)
    assert with_bypass.returncode == 0, with_bypass.stderr[-2000:]
    assert with_bypass.stdout.strip() == "clean"
    without = subprocess.run([sys.executable, "-c", prog.format(bypass=False)], capture_output=True, text=True, env=env, timeout=300)
    assert without.stdout.strip() == "dynamo", "torch no longer imports dynamo here: the bypass can go"


def test_cpu_thread_share(monkeypatch):
    before = torch.get_num_threads()
    try:
        monkeypatch.delenv("OMP_NUM_THREADS", raising=False)
        monkeypatch.setenv("DET_AGENT_SLOT_COUNT", "8")
        cores = len(os.sched_getaffinity(0))
        assert fast_start.cap_cpu_threads() == max(1, min(before, cores // 8))
        monkeypatch.setenv("OMP_NUM_THREADS", "5")            # an explicit setting wins
        torch.set_num_threads(3)
        assert fast_start.cap_cpu_threads() == 3
        monkeypatch.delenv("OMP_NUM_THREADS")
        monkeypatch.setenv("DET_AGENT_SLOT_COUNT", "1")       # a one-slot agent keeps torch's default
        assert fast_start.cap_cpu_threads() == 3
    finally:
        torch.set_num_threads(before)


def test_warm_worker_adopts_and_releases_a_trial(tmp_path, monkeypatch):
    exp_a, exp_b = tmp_path / "exp_a", tmp_path / "exp_b"
    for d, val in ((exp_a, 1), (exp_b, 2)):
        d.mkdir()
        (d / "model_def.py").write_text(f"VALUE = {val}\n")
    monkeypatch.setenv("PYTHONPATH", f"{tmp_path}/first_trial:{ROOT}")
    monkeypatch.setenv("LEFTOVER", "from-the-agent")
    base_env, base_path, base_modules = dict(os.environ), list(sys.path) + [f"{tmp_path}/first_trial"], set(sys.modules)
    cwd = os.getcwd()
    try:
        warm_worker.adopt({"env": {"DET_TRIAL_ID": "7", "PYTHONPATH": f"{exp_a}:{ROOT}"}, "cwd": str(exp_a)}, base_env, base_path)
        assert os.environ["DET_TRIAL_ID"] == "7" and os.environ["LEFTOVER"] == "from-the-agent"
        assert os.getcwd() == str(exp_a) and sys.path[0] == str(exp_a)
        assert f"{tmp_path}/first_trial" not in sys.path            # only the agent's first trial needed it
        import model_def

        assert model_def.VALUE == 1
        os.environ["SET_BY_TRIAL_A"] = "1"
        warm_worker.release({"cwd": str(exp_a)}, base_modules)
        assert "model_def" not in sys.modules                        # the next experiment brings its own
        warm_worker.adopt({"env": {"DET_TRIAL_ID": "8", "PYTHONPATH": f"{exp_b}:{ROOT}"}, "cwd": str(exp_b)}, base_env, base_path)
        assert os.environ["DET_TRIAL_ID"] == "8" and "SET_BY_TRIAL_A" not in os.environ
        assert str(exp_a) not in sys.path and sys.path[0] == str(exp_b)
        import model_def as again

        assert again.VALUE == 2
    finally:
        os.chdir(cwd)
        sys.modules.pop("model_def", None)
        sys.path[:] = [p for p in sys.path if not p.startswith(str(tmp_path))]
        os.environ.clear()
        os.environ.update(base_env)
