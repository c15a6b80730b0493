"""Kubernetes container back-end (reference: master/internal/kubernetes/{spec,pods,pod,informer,request_queue}.go and
resourcemanagers/kubernetes_resource_manager.go): pod manifests, and the master driving experiments through pods against a
stand-in API server whose "kubelet" runs the pod command as a local process."""
import os
import pathlib
import subprocess
import sys
import time

import pytest

from determined_b200.experimental.client import Determined
from determined_b200.master import _core
from tests.fake_kube import FakeKube

ROOT = pathlib.Path(__file__).resolve().parent.parent
MASTER = ROOT / "determined_b200" / "_lib" / "det-master"
MODEL_DIR = str(ROOT / "tests" / "fixtures" / "xor_model_def")


def test_pod_spec_generation():
    container = {"container_id": "Trial-7.r0_c1", "task_id": "trial-7", "description": "exp 3 trial 7",
                 "argv": ["python", "-m", "determined_b200.exec.harness"], "cwd": "/run/determined",
                 "env": {"DET_MASTER_ADDR": "127.0.0.1", "DET_MASTER_PORT": "8080", "DET_TRIAL_ID": "7"},
                 "devices": [{"id": 0, "type": "gpu"}, {"id": 1, "type": "gpu"}]}
    cfg = {"type": "kubernetes", "namespace": "ml", "image": "registry/det:1", "master_service_name": "det-master.ml.svc",
           "shared_fs_host_path": "/mnt/ckpt", "service_account": "det", "pod_spec": {"nodeSelector": {"pool": "b200"}}}
    pod = _core.pod_spec(container, cfg, cluster_id="abc", node="node-a", host_network=True)
    assert pod["kind"] == "Pod" and pod["metadata"]["namespace"] == "ml"
    name = pod["metadata"]["name"]
    assert name == "det-trial-7-r0-c1" and len(name) <= 63          # RFC 1123 label
    assert pod["metadata"]["labels"] == {"determined": "abc", "determined-task": "trial-7"}
    assert pod["metadata"]["annotations"]["determined.ai/container-id"] == "Trial-7.r0_c1"
    spec = pod["spec"]
    assert spec["restartPolicy"] == "Never" and spec["nodeName"] == "node-a" and spec["hostNetwork"] is True
    assert spec["serviceAccountName"] == "det" and spec["nodeSelector"] == {"pool": "b200"}
    c = spec["containers"][0]
    assert c["image"] == "registry/det:1" and c["command"] == container["argv"] and c["workingDir"] == "/run/determined"
    env = {e["name"]: e["value"] for e in c["env"]}
    assert env["DET_MASTER_ADDR"] == "det-master.ml.svc" and env["DET_MASTER_PORT"] == "8080" and env["DET_TRIAL_ID"] == "7"
    assert env["DET_CONTAINER_ID"] == "Trial-7.r0_c1" and env["DET_SLOT_IDS"] == "[0,1]" and env["DET_USE_GPU"] == "true"
    assert c["resources"]["limits"] == {"nvidia.com/gpu": 2}
    mounts = {m["name"]: m["mountPath"] for m in c["volumeMounts"]}
    assert mounts == {"det-shared-fs": "/mnt/ckpt", "det-shm": "/dev/shm"}
    vols = {v["name"]: v for v in spec["volumes"]}
    assert vols["det-shared-fs"]["hostPath"]["path"] == "/mnt/ckpt" and vols["det-shm"]["emptyDir"]["medium"] == "Memory"
    # zero-slot task: no GPU limit, no host network
    pod0 = _core.pod_spec(dict(container, devices=[]), {"type": "kubernetes"}, cluster_id="abc")
    assert pod0["spec"]["containers"][0]["resources"] == {} and "hostNetwork" not in pod0["spec"] and "nodeName" not in pod0["spec"]


@pytest.fixture(scope="module")
def kube_master(tmp_path_factory):
    from determined_b200 import _build

    _build.build_core()
    kube = FakeKube({"node-a": 2, "node-b": 2})
    tmp = tmp_path_factory.mktemp("kmaster")
    port_file = tmp / "port"
    rm = '{"type": "kubernetes", "api_server": "127.0.0.1:%d", "namespace": "det", "image": "det:test", "poll_ms": 200, "slot_type": "cpu"}' % kube.port
    proc = subprocess.Popen([str(MASTER), "--port", "0", "--data-dir", str(tmp / "data"), "--repo-root", str(ROOT), "--python", sys.executable,
                             "--resource-manager", rm, "--print-port-file", str(port_file)],
                            stdout=open(tmp / "master.log", "w"), stderr=subprocess.STDOUT,
                            env=dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg")))
    for _ in range(100):
        if port_file.exists() and port_file.read_text().strip():
            break
        time.sleep(0.05)
    port = int(port_file.read_text())
    os.environ["DET_CONFIG_DIR"] = str(tmp / "cfg")
    yield {"url": f"127.0.0.1:{port}", "tmp": tmp, "kube": kube}
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()
    kube.close()


def _config(tmp, **extra):
    cfg = {"entrypoint": "model_def:XorTrial", "description": "k8s", "scheduling_unit": 10, "max_restarts": 0,
           "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
           "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 20}},
           "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp / "ckpts")}}
    cfg.update(extra)
    return cfg


def test_nodes_register_as_agents(kube_master):
    d = Determined(kube_master["url"])
    for _ in range(50):
        agents = d._session.get("/api/v1/agents")
        if len(agents) == 2:
            break
        time.sleep(0.1)
    assert sorted(a["id"] for a in agents) == ["k8s-node-a", "k8s-node-b"]
    assert all(a["num_slots"] == 2 for a in agents)


def test_experiment_runs_in_pods(kube_master):
    kube = kube_master["kube"]
    d = Determined(kube_master["url"])
    kube.fail_next_creates = 1      # the request queue retries a 503
    exp = d.create_experiment(_config(kube_master["tmp"]), MODEL_DIR)
    assert exp.wait(timeout=180) == "COMPLETED"
    t = exp.describe()["trials"][0]
    assert t["state"] == "COMPLETED" and t["total_batches_processed"] == 20
    trial_pods = [m for m in kube.created if "trial" in m["metadata"]["name"]]
    assert trial_pods, [m["metadata"]["name"] for m in kube.created]
    m = trial_pods[0]
    assert m["metadata"]["namespace"] == "det" and m["spec"]["containers"][0]["image"] == "det:test"
    assert m["spec"]["nodeName"] in ("node-a", "node-b")
    assert m["spec"]["containers"][0]["resources"]["limits"] == {"nvidia.com/gpu": 1}
    logs = d.get_trial(t["id"]).logs()
    assert any("validation" in l["message"].lower() or "checkpoint" in l["message"].lower() for l in logs), logs[-5:]
    for _ in range(50):             # finished pods are deleted once their exit status is recorded
        if m["metadata"]["name"] in kube.deleted:
            break
        time.sleep(0.1)
    assert m["metadata"]["name"] in kube.deleted


def test_kill_deletes_the_pod(kube_master):
    kube = kube_master["kube"]
    d = Determined(kube_master["url"])
    cfg = _config(kube_master["tmp"], searcher={"name": "single", "metric": "validation_loss", "max_length": {"batches": 100000}})
    exp = d.create_experiment(cfg, MODEL_DIR)
    for _ in range(200):
        pods = d._session.get("/api/v1/kubernetes/pods")
        running = [p for p in pods["pods"] if p["state"] == "RUNNING"]
        if running:
            break
        time.sleep(0.1)
    assert running and pods["enabled"]
    exp.kill()
    assert exp.wait(timeout=60) == "CANCELED"
    for _ in range(100):
        if running[0]["pod"] in kube.deleted:
            break
        time.sleep(0.1)
    assert running[0]["pod"] in kube.deleted
