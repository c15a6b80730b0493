"""A trial that spans two agents (reference: multi-machine data parallelism — trial.go pushRendezvous :813-908,
layers/_worker_process.py:113-199): the master hands every container the chief's address, a rank offset and the world size,
the workers of both containers join ONE process group, only global rank 0 talks to the master."""
import os
import pathlib
import subprocess
import sys
import time

from determined_b200.exec import harness
from determined_b200.experimental.client import Determined

ROOT = pathlib.Path(__file__).resolve().parent.parent
LIB = ROOT / "determined_b200" / "_lib"
MODEL_DIR = str(ROOT / "tests" / "fixtures" / "xor_model_def")


def test_rendezvous_env():
    one = harness.rendezvous_env(4, {"DET_NUM_CONTAINERS": "1"})
    assert one["addr"] == "127.0.0.1" and one["offset"] == 0 and one["world"] == 4 and one["port"] > 0
    env = {"DET_NUM_CONTAINERS": "2", "DET_CHIEF_ADDR": "10.0.0.7", "DET_CHIEF_PORT": "23456", "DET_RANK_OFFSET": "8", "DET_WORLD_SLOTS": "16"}
    assert harness.rendezvous_env(8, env) == {"addr": "10.0.0.7", "port": 23456, "offset": 8, "world": 16}


def test_trial_spans_two_agents(tmp_path):
    from determined_b200 import _build

    _build.build_core()
    port_file = tmp_path / "port"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp_path / "cfg"))
    procs = [subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp_path / "data"), "--slot-type", "none", "--repo-root", str(ROOT),
                               "--python", sys.executable, "--print-port-file", str(port_file)],
                              stdout=open(tmp_path / "master.log", "w"), stderr=subprocess.STDOUT, env=env)]
    try:
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        port = int(port_file.read_text())
        for name in ("node-a", "node-b"):
            procs.append(subprocess.Popen([str(LIB / "det-agent"), "--master-host", "127.0.0.1", "--master-port", str(port), "--agent-id", name,
                                           "--artificial-slots", "1"], stdout=open(tmp_path / f"{name}.log", "w"), stderr=subprocess.STDOUT))
        os.environ["DET_CONFIG_DIR"] = str(tmp_path / "cfg")
        d = Determined(f"127.0.0.1:{port}")
        for _ in range(100):
            agents = {a["id"]: a for a in d._session.get("/api/v1/agents")}
            if {"node-a", "node-b"} <= set(agents):
                break
            time.sleep(0.1)
        assert agents["node-a"]["addresses"][0] == "127.0.0.1"        # where the registration came from, not a host name
        cfg = {"entrypoint": "model_def:XorTrial", "description": "two containers", "scheduling_unit": 10,
               "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
               "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 20}},
               "resources": {"slots_per_trial": 2}, "max_restarts": 0,
               "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpts")}}
        exp = d.create_experiment(cfg, MODEL_DIR)
        assert exp.wait(timeout=240) == "COMPLETED", (tmp_path / "master.log").read_text()[-2000:]
        t = exp.describe()["trials"][0]
        assert t["state"] == "COMPLETED" and t["total_batches_processed"] == 20
        logs = d.get_trial(t["id"]).logs()
        hello = sorted((l["rank"], l["agent_id"]) for l in logs if "worker rank" in l["message"])
        assert [r for r, _ in hello] == [0, 1] and {a for _, a in hello} == {"node-a", "node-b"}, hello
        assert any("worker rank 1 of 2" in l["message"] and "2 container(s)" in l["message"] for l in logs)
    finally:
        for p in reversed(procs):
            p.terminate()
        for p in procs:
            try:
                p.wait(timeout=5)
            except Exception:
                p.kill()
