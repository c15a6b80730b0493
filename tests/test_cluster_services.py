"""Cluster services around the master: remote agent container protocol, reader/writer leases, provisioner scale
decisions, det-deploy local, data-layer cache, and the small utility modules."""
import json
import os
import pathlib
import subprocess
import sys
import threading
import time

import pytest
import torch

from determined_b200 import check, data_layer, gpu
from determined_b200.experimental.client import Determined
from determined_b200.layers._master_client import MasterClient
from determined_b200.master import _core

ROOT = pathlib.Path(__file__).resolve().parent.parent
LIB = ROOT / "determined_b200" / "_lib"
MODEL_DIR = str(ROOT / "tests" / "fixtures" / "xor_model_def")


@pytest.fixture(scope="module")
def slotless_master(tmp_path_factory):
    """A master whose own agent has NO slots: every trial must run on a remote det-agent."""
    from determined_b200 import _build

    _build.build_core()
    tmp = tmp_path_factory.mktemp("m2")
    port_file = tmp / "port"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg"))
    proc = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp / "data"), "--slot-type", "none", "--repo-root", str(ROOT),
                             "--python", sys.executable, "--print-port-file", str(port_file)],
                            stdout=open(tmp / "master.log", "w"), stderr=subprocess.STDOUT, env=env)
    for _ in range(200):
        if port_file.exists() and port_file.read_text().strip():
            break
        time.sleep(0.05)
    port = int(port_file.read_text())
    os.environ["DET_CONFIG_DIR"] = str(tmp / "cfg")
    agent = subprocess.Popen([str(LIB / "det-agent"), "--master-host", "127.0.0.1", "--master-port", str(port), "--agent-id", "remote-1",
                              "--artificial-slots", "2"], stdout=open(tmp / "agent.log", "w"), stderr=subprocess.STDOUT)
    yield {"url": f"127.0.0.1:{port}", "port": port, "tmp": tmp}
    for p in (agent, proc):
        p.terminate()
        try:
            p.wait(timeout=5)
        except Exception:
            p.kill()


def test_trial_runs_on_remote_agent(slotless_master):
    d = Determined(slotless_master["url"])
    for _ in range(100):
        agents = d._session.get("/api/v1/agents")
        if any(a["id"] == "remote-1" for a in agents):
            break
        time.sleep(0.1)
    assert any(a["id"] == "remote-1" and a["num_slots"] == 2 for a in agents)
    cfg = {"entrypoint": "model_def:XorTrial", "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
           "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 20}}, "scheduling_unit": 10,
           "checkpoint_storage": {"type": "shared_fs", "host_path": str(slotless_master["tmp"] / "ckpts")}}
    exp = d.create_experiment(cfg, MODEL_DIR)
    assert exp.wait(timeout=180) == "COMPLETED"
    t = exp.describe()["trials"][0]
    assert t["total_batches_processed"] == 20
    logs = d.get_trial(t["id"]).logs()
    assert logs, "the remote agent's log lines reached the master"


def test_agent_configuration_file_and_proxy_environment(slotless_master, tmp_path):
    """det-agent --config-file: the reference's agent.yaml keys (as JSON); flags override the file; the proxy settings reach
    the tasks the agent runs (reference agent/internal/options.go); unknown keys are refused."""
    cfg = tmp_path / "agent.json"
    cfg.write_text(json.dumps({"master_host": "127.0.0.1", "master_port": slotless_master["port"], "agent_id": "from-file", "label": "file-label",
                               "artificial_slots": 1, "http_proxy": "http://proxy.example:3128", "no_proxy": "localhost,127.0.0.1",
                               "security": {"tls": {"enabled": False}}, "fluent": {"image": "ignored"}}))
    agent = subprocess.Popen([str(LIB / "det-agent"), "--config-file", str(cfg), "--label", "flag-label"], stdout=open(tmp_path / "agent.log", "w"),
                             stderr=subprocess.STDOUT)
    try:
        c = MasterClient("127.0.0.1", slotless_master["port"])
        for _ in range(100):
            agents = {a["id"]: a for a in c.request("GET", "/api/v1/agents")}
            if "from-file" in agents:
                break
            time.sleep(0.1)
        assert agents["from-file"]["num_slots"] == 1 and agents["from-file"]["label"] == "flag-label"
        # a task that asks for that label lands on this agent and sees the proxy settings
        cmd = c.request("POST", "/api/v1/commands", {"argv": ["sh", "-c", "echo $http_proxy $HTTP_PROXY $no_proxy"], "slots": 1,
                                                     "config": {"resources": {"agent_label": "flag-label"}}})
        for _ in range(150):
            cur = c.request("GET", f"/api/v1/commands/{cmd['id']}")
            if cur["state"] == "TERMINATED":
                break
            time.sleep(0.1)
        assert [l["message"] for l in cur["logs"]] == ["http://proxy.example:3128 http://proxy.example:3128 localhost,127.0.0.1"], cur
    finally:
        agent.terminate()
        agent.wait(timeout=10)
    cfg.write_text(json.dumps({"master_host": "127.0.0.1", "colour": "red"}))
    bad = subprocess.run([str(LIB / "det-agent"), "--config-file", str(cfg)], capture_output=True, text=True, timeout=10)
    assert bad.returncode == 2 and "unknown field in agent config: 'colour'" in bad.stderr


def test_rw_coordinator_leases(slotless_master):
    c = MasterClient("127.0.0.1", slotless_master["port"])
    r1 = c.request("POST", "/api/v1/rw_coordinator/ds1/read_lock", {"wait_ms": 1000})
    r2 = c.request("POST", "/api/v1/rw_coordinator/ds1/read_lock", {"wait_ms": 1000})
    assert r1["granted"] and r2["granted"] and r1["lease"] != r2["lease"]
    w = c.request("POST", "/api/v1/rw_coordinator/ds1/write_lock", {"wait_ms": 300})
    assert not w["granted"], "a writer must wait for the readers"
    got = {}

    def writer():
        got["w"] = MasterClient("127.0.0.1", slotless_master["port"]).request("POST", "/api/v1/rw_coordinator/ds1/write_lock", {"wait_ms": 5000})

    th = threading.Thread(target=writer)
    th.start()
    time.sleep(0.3)
    st = c.request("GET", "/api/v1/rw_coordinator/ds1/stats")
    assert st["readers"] == 2 and st["waiting"] == 1
    late_reader = c.request("POST", "/api/v1/rw_coordinator/ds1/read_lock", {"wait_ms": 200})
    assert not late_reader["granted"], "readers queue behind a waiting writer (no writer starvation)"
    for r in (r1, r2):
        assert c.request("POST", "/api/v1/rw_coordinator/ds1/unlock", {"lease": r["lease"]})["ok"]
    th.join(timeout=10)
    assert got["w"]["granted"]
    assert not c.request("POST", "/api/v1/rw_coordinator/ds1/read_lock", {"wait_ms": 200})["granted"]
    c.request("POST", "/api/v1/rw_coordinator/ds1/unlock", {"lease": got["w"]["lease"]})
    # an abandoned lease expires
    lease = c.request("POST", "/api/v1/rw_coordinator/ds2/write_lock", {"wait_ms": 200, "ttl_ms": 300})
    assert lease["granted"]
    time.sleep(0.5)
    assert c.request("POST", "/api/v1/rw_coordinator/ds2/write_lock", {"wait_ms": 500})["granted"]


def test_provisioner_scale_decider():
    cfg = {"max_instances": 3, "slots_per_instance": 8, "max_idle_agent_period": 60, "max_agent_starting_period": 120}
    # demand for 20 slots, nothing running: launch ceil(20/8) = 3
    d = _core.scale_decide(cfg, {"now": 1000, "instances": [], "agents": [], "demand": {"pending_slots": 20}})
    assert d["launch"] == 3 and d["terminate"] == []
    # one instance still starting covers part of the demand; cap still applies
    d = _core.scale_decide(cfg, {"now": 1000, "instances": [{"id": "i1", "state": "starting", "launch_time": 950, "agent": ""}], "agents": [],
                                 "demand": {"pending_slots": 20}})
    assert d["launch"] == 2
    snap = {"now": 1000, "demand": {"pending_slots": 0},
            "instances": [{"id": "idle", "state": "running", "launch_time": 0, "agent": "a1"},
                          {"id": "busy", "state": "running", "launch_time": 0, "agent": "a2"},
                          {"id": "lost", "state": "running", "launch_time": 100, "agent": ""},
                          {"id": "fresh", "state": "starting", "launch_time": 990, "agent": ""},
                          {"id": "dead", "state": "stopped", "launch_time": 0, "agent": ""}],
            "agents": [{"id": "a1", "idle_since": 900, "connected": True}, {"id": "a2", "idle_since": None, "connected": True}]}
    d = _core.scale_decide(cfg, snap)
    assert d["launch"] == 0 and sorted(d["terminate"]) == ["dead", "idle", "lost"]
    assert "idle" in d["reasons"]["idle"]
    # pool above the cap: shed idle first
    cfg2 = dict(cfg, max_instances=1, max_idle_agent_period=1e9)
    d = _core.scale_decide(cfg2, {"now": 10, "demand": {"pending_slots": 0},
                                  "instances": [{"id": "x", "state": "running", "launch_time": 0, "agent": "a1"},
                                                {"id": "y", "state": "running", "launch_time": 1, "agent": "a2"}],
                                  "agents": [{"id": "a1", "idle_since": None, "connected": True}, {"id": "a2", "idle_since": 5, "connected": True}]})
    assert d["terminate"] == ["y"]


def test_det_deploy_local_up_down(tmp_path):
    env = dict(os.environ, DET_DEPLOY_DIR=str(tmp_path), DET_CONFIG_DIR=str(tmp_path / "cfg"))
    up = subprocess.run([sys.executable, str(ROOT / "det-deploy"), "local", "cluster-up", "--cluster-name", "t", "--master-port", "0",
                         "--artificial-slots", "2", "--agents", "1"], capture_output=True, text=True, env=env, timeout=120)
    assert up.returncode == 0, up.stderr
    info = json.loads(up.stdout.strip().splitlines()[-1])
    try:
        c = MasterClient(*info["master"].split(":")[:1], int(info["master"].split(":")[1]))
        for _ in range(100):
            agents = c.request("GET", "/api/v1/agents")
            if len(agents) == 2:
                break
            time.sleep(0.1)
        assert {a["id"] for a in agents} == {"agent-local", "agent-1"}
        import urllib.request
        page = urllib.request.urlopen(f"http://{info['master']}/ui", timeout=5).read().decode()
        assert "<title>determined_b200</title>" in page and "#/experiments" in page     # the WebUI, not the embedded fallback
        root = urllib.request.urlopen(f"http://{info['master']}/", timeout=5).read().decode()
        assert "/ui" in root
    finally:
        down = subprocess.run([sys.executable, str(ROOT / "det-deploy"), "local", "cluster-down", "--cluster-name", "t"], capture_output=True,
                              text=True, env=env, timeout=60)
    assert down.returncode == 0 and len(json.loads(down.stdout)["stopped"]) == 2


def test_det_deploy_local_master_and_agents_separately(tmp_path):
    """master-up with a configuration file, agent-up / agent-down against it, master-down (reference deploy/test_local.py:
    test_custom_etc, test_master_up_down, test_agent_up_down)."""
    env = dict(os.environ, DET_DEPLOY_DIR=str(tmp_path), DET_CONFIG_DIR=str(tmp_path / "cfg"))
    run = lambda *a: subprocess.run([sys.executable, str(ROOT / "det-deploy"), "local", *a, "--cluster-name", "sep"], capture_output=True,  # noqa: E731
                                    text=True, env=env, timeout=120)
    etc = tmp_path / "master.yaml"
    etc.write_text("port: 0\ncluster_name: from-the-file\nscheduler: priority\nartificial_slots: 1\n")
    up = run("master-up", "--master-config-path", str(etc), "--master-port", "0")
    assert up.returncode == 0, up.stderr
    info = json.loads(up.stdout.strip().splitlines()[-1])
    host, port = info["master"].split(":")
    try:
        c = MasterClient(host, int(port))
        assert c.request("GET", "/info")["cluster_name"] == "from-the-file" and c.request("GET", "/info")["scheduler"] == "priority"
        assert run("master-up", "--master-port", "0").returncode != 0          # one master per cluster name
        for name in ("a1", "a2"):
            assert run("agent-up", "127.0.0.1", "--master-port", port, "--agent-name", name, "--artificial-slots", "1",
                       "--agent-label", "pool-x").returncode == 0
        for _ in range(100):
            agents = {a["id"]: a for a in c.request("GET", "/api/v1/agents")}
            if {"a1", "a2"} <= set(agents):
                break
            time.sleep(0.1)
        assert {"agent-local", "a1", "a2"} == set(agents) and agents["a1"]["label"] == "pool-x"
        assert run("agent-up", "127.0.0.1", "--master-port", port, "--agent-name", "a1").returncode != 0     # already running
        assert len(json.loads(run("agent-down", "--agent-name", "a1").stdout)["stopped"]) == 1
        for _ in range(100):
            if "a1" not in {a["id"] for a in c.request("GET", "/api/v1/agents")}:
                break
            time.sleep(0.1)
        assert {a["id"] for a in c.request("GET", "/api/v1/agents")} == {"agent-local", "a2"}
        assert len(json.loads(run("agent-down", "--all").stdout)["stopped"]) == 1
    finally:
        down = run("master-down")
        run("cluster-down")
    assert down.returncode == 0 and len(json.loads(down.stdout)["stopped"]) == 1


def test_data_layer_builds_once_and_serves_cached_copy(tmp_path):
    calls = []
    dl = data_layer.DataLayer({"type": "lfs", "container_storage_path": str(tmp_path)})

    @dl.cache_train_dataset("mnist-like", "v1", shuffle=True)
    def make():
        calls.append(1)
        return torch.utils.data.TensorDataset(torch.arange(12.0).view(6, 2), torch.arange(6))

    a = make()
    b = make()
    assert len(calls) == 1 and dl.builds == 1
    assert len(a) == len(b) == 6 and torch.equal(b[3][0], torch.tensor([6.0, 7.0])) and int(b[3][1]) == 3
    xs, ys = b.__getitems__([0, 5])
    assert xs.shape == (2, 2) and ys.tolist() == [0, 5]
    # a new version is a new cache entry; a second process-level DataLayer finds the old one on disk
    dl2 = data_layer.DataLayer({"type": "lfs", "container_storage_path": str(tmp_path)})
    again = dl2.cache_train_dataset("mnist-like", "v1")(lambda: pytest.fail("must not rebuild"))()
    assert len(again) == 6 and dl2.builds == 0
    with pytest.raises(ValueError):
        data_layer.DataLayer({"type": "tape"})


def test_data_layer_bucket_backed_cache(tmp_path, monkeypatch):
    """type: s3 / gcs — the first builder uploads the built data set; another box (empty local cache) downloads it
    instead of rebuilding."""
    from tests import fake_objstore

    def build():
        return torch.utils.data.TensorDataset(torch.arange(10.0).reshape(5, 2), torch.arange(5))

    srv, url, objects, stats = fake_objstore.make_s3()
    gsrv, gurl, gobjects = fake_objstore.make_gcs()
    monkeypatch.setenv("STORAGE_EMULATOR_HOST", gurl)
    try:
        for kind, store in (("s3", objects), ("gcs", gobjects)):
            cfg = {"type": kind, "bucket": "data", "bucket_directory_path": "cache/v1", "access_key": "AKIDEXAMPLE",
                   "secret_key": "s3cr3t", "endpoint_url": url}
            a = data_layer.DataLayer(dict(cfg, local_cache_container_path=str(tmp_path / kind / "box-a")))
            ds = a.cache_train_dataset("toy", "1")(build)()
            assert a.builds == 1 and a.downloads == 0 and len(ds) == 5
            assert list(store) == ["cache/v1/toy-1-train.pt"]
            b = data_layer.DataLayer(dict(cfg, local_cache_container_path=str(tmp_path / kind / "box-b")))
            ds_b = b.cache_train_dataset("toy", "1")(lambda: pytest.fail("must come from the bucket"))()
            assert b.builds == 0 and b.downloads == 1 and torch.equal(ds_b.columns[0], ds.columns[0])
            again = b.cache_train_dataset("toy", "1")(lambda: pytest.fail("must come from the local copy"))()
            assert b.downloads == 1 and len(again) == 5
        assert stats["bad_signature"] == 0
    finally:
        srv.shutdown()
        gsrv.shutdown()


def test_data_layer_stream_resumes_shards_and_reshuffles(tmp_path):
    """The positioned stream: per-epoch seeded shuffle, round-robin shards that drop the remainder while training, and a
    start offset that may lie epochs in (reference _data_layer.py:57-69, 186-204)."""
    def build():
        return torch.utils.data.TensorDataset(torch.arange(11))

    def layer(**pos):
        return data_layer.DataLayer({"type": "lfs", "container_storage_path": str(tmp_path)}, position=pos)

    full = layer(shuffle_seed=7).cache_train_dataset("s", "1", shuffle=True)(build)()
    e0 = [int(x) for x in full.stream(epochs=1)]
    e01 = [int(x) for x in full.stream(epochs=2)]
    assert sorted(e0) == list(range(11)) and e01[:11] == e0 and e01[11:] != e0 and sorted(e01[11:]) == list(range(11))
    # resume 15 records in == the tail of an uninterrupted read
    resumed = layer(shuffle_seed=7, start_offset=15).cache_train_dataset("s", "1", shuffle=True)(build)()
    it = resumed.stream()
    assert [int(next(it)) for _ in range(7)] == e01[15:22]
    # two shards: disjoint, equal length (remainder dropped), together one epoch minus the dropped record
    shards = [[int(x) for x in layer(shuffle_seed=7, shard_rank=r, num_shards=2).cache_train_dataset("s", "1", shuffle=True)(build)()
               .stream(epochs=1)] for r in range(2)]
    assert len(shards[0]) == len(shards[1]) == 5 and not set(shards[0]) & set(shards[1])
    assert sorted(shards[0] + shards[1]) == sorted(e0[:10])
    # validation keeps the remainder and starts at 0 whatever the training offset is
    val = layer(start_offset=15, shard_rank=1, num_shards=2).cache_validation_dataset("s", "1")(build)()
    assert [int(x) for x in val.stream(epochs=1)] == [1, 3, 5, 7, 9]
    pinned = layer(shuffle_seed=7).cache_train_dataset("s", "1", shuffle=True, skip_shuffle_at_epoch_end=True)(build)()
    two = [int(x) for x in pinned.stream(epochs=2)]
    assert two[:11] == two[11:]


def test_check_and_gpu_helpers():
    check.true(1)
    check.eq(2, 2)
    check.is_in("a", {"a"})
    check.len_eq([1, 2], 2)
    with pytest.raises(check.CheckFailedError, match="does not equal"):
        check.eq(1, 2, "numbers")
    with pytest.raises(check.CheckFailedError):
        check.is_instance(1, str)
    assert isinstance(gpu.get_gpus(), list)          # empty on a CPU box
    assert gpu.get_gpu_uuids_and_validate(False, []) == []


def test_provisioner_scales_up_on_demand_and_down_when_idle(tmp_path):
    """A master with no slots of its own and a provisioner: submitting an experiment makes it start a det-agent
    instance, the trial runs there, and the idle instance is terminated afterwards."""
    import json as _json

    port_file = tmp_path / "port"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp_path / "cfg"))
    prov = {"max_instances": 2, "slots_per_instance": 1, "max_idle_agent_period": 2, "max_agent_starting_period": 30, "tick_ms": 300}
    proc = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp_path / "data"), "--slot-type", "none", "--repo-root", str(ROOT),
                             "--python", sys.executable, "--print-port-file", str(port_file), "--provisioner", _json.dumps(prov)],
                            stdout=open(tmp_path / "master.log", "w"), stderr=subprocess.STDOUT, env=env, start_new_session=True)
    try:
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        port = int(port_file.read_text())
        os.environ["DET_CONFIG_DIR"] = str(tmp_path / "cfg")
        d = Determined(f"127.0.0.1:{port}")
        c = MasterClient("127.0.0.1", port)
        assert c.request("GET", "/api/v1/provisioner")["instances"] == []
        cfg = {"entrypoint": "model_def:XorTrial", "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
               "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 20}}, "scheduling_unit": 10,
               "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpts")}}
        exp = d.create_experiment(cfg, MODEL_DIR)
        assert exp.wait(timeout=180) == "COMPLETED"
        s = c.request("GET", "/api/v1/provisioner")
        assert s["launched_total"] >= 1
        for _ in range(100):            # idle for > 2 s → scaled down
            s = c.request("GET", "/api/v1/provisioner")
            if s["terminated_total"] >= 1 and not s["instances"]:
                break
            time.sleep(0.2)
        assert s["terminated_total"] >= 1 and s["instances"] == []
        assert all(a["id"] == "agent-local" for a in c.request("GET", "/api/v1/agents"))
    finally:
        import signal as _signal
        os.killpg(proc.pid, _signal.SIGTERM)
        try:
            proc.wait(timeout=5)
        except Exception:
            os.killpg(proc.pid, _signal.SIGKILL)


def test_actor_trace_and_debug_endpoints(slotless_master, tmp_path):
    c = MasterClient("127.0.0.1", slotless_master["port"])
    actors = c.request("GET", "/debug/actors")
    assert {"rm", "trial_logger", "experiments"} <= {a["address"] for a in actors}
    # the caller names a file; the master decides the directory (<data_dir>/traces) — a path is reduced to its last component
    where = c.request("POST", "/debug/trace", {"file": str(tmp_path / "actors.trace")})["file"]
    assert where.endswith("/traces/actors.trace") and not where.startswith(str(tmp_path))
    trace = pathlib.Path(where)
    c.request("GET", "/api/v1/agents")           # an ask to the resource manager
    c.request("POST", "/debug/trace", {"file": ""})
    events = [json.loads(l) for l in trace.read_text().splitlines() if l.strip()]
    assert any(e["tid"] == "rm" and e["name"] == "summary" and e["ph"] == "X" and e["dur"] >= 0 for e in events)


def test_service_task_readiness_and_proxy(slotless_master, tmp_path):
    """A command submitted with ``proxy: true`` gets a port ({port} in argv, DET_TASK_PORT), turns ``is_ready`` once it
    answers HTTP, and is reachable through ``/proxy/<id>/...`` on the master (reference: master/internal/command readiness
    checks + master/internal/proxy/proxy.go; notebooks and TensorBoards use the same path)."""
    import urllib.error
    import urllib.request

    (tmp_path / "hello.txt").write_text("served through the master\n")
    (tmp_path / "sub").mkdir()
    (tmp_path / "sub" / "data.json").write_text('{"x": 1}')
    c = MasterClient("127.0.0.1", slotless_master["port"])
    cmd = c.request("POST", "/api/v1/commands", {"argv": [sys.executable, "-m", "http.server", "{port}", "--bind", "127.0.0.1"], "cwd": str(tmp_path),
                                                "proxy": True, "description": "file server"})
    assert cmd["spec"]["proxy_port"] > 0
    try:
        for _ in range(150):
            cur = c.request("GET", f"/api/v1/commands/{cmd['id']}")
            if cur.get("is_ready"):
                break
            time.sleep(0.1)
        assert cur["is_ready"] and cur["state"] == "RUNNING" and cur["service_address"] == f"/proxy/{cmd['id']}/"
        base = f"http://127.0.0.1:{slotless_master['port']}/proxy/{cmd['id']}"
        assert urllib.request.urlopen(base + "/hello.txt", timeout=10).read() == b"served through the master\n"
        r = urllib.request.urlopen(base + "/sub/data.json", timeout=10)
        assert json.loads(r.read()) == {"x": 1} and "json" in r.headers["Content-Type"]
        assert b"hello.txt" in urllib.request.urlopen(base + "/", timeout=10).read()        # directory listing of the task's cwd
        with pytest.raises(urllib.error.HTTPError) as e:
            urllib.request.urlopen(base + "/missing", timeout=10)
        assert e.value.code == 404
    finally:
        c.request("POST", f"/api/v1/commands/{cmd['id']}/kill")
    for _ in range(100):
        cur = c.request("GET", f"/api/v1/commands/{cmd['id']}")
        if cur["state"] == "TERMINATED":
            break
        time.sleep(0.1)
    assert cur["state"] == "TERMINATED" and not cur["is_ready"]
    with pytest.raises(urllib.error.HTTPError) as e:
        urllib.request.urlopen(f"http://127.0.0.1:{slotless_master['port']}/proxy/{cmd['id']}/hello.txt", timeout=10)
    assert e.value.code == 502
    plain = c.request("POST", "/api/v1/commands", {"argv": ["/bin/true"]})
    assert "proxy_port" not in plain["spec"]


def test_proxy_tunnels_protocol_upgrades(slotless_master):
    """A request with an ``Upgrade:`` header (WebSocket) is not replayed as HTTP: the master forwards the handshake and
    then copies bytes both ways, so interactive services (Jupyter kernels, terminals) work behind ``/proxy/<id>/``."""
    import socket

    c = MasterClient("127.0.0.1", slotless_master["port"])
    script = str(ROOT / "tests" / "fixtures" / "ws_echo.py")
    cmd = c.request("POST", "/api/v1/commands", {"argv": [sys.executable, script, "{port}"], "proxy": True, "description": "ws echo"})
    try:
        for _ in range(150):
            cur = c.request("GET", f"/api/v1/commands/{cmd['id']}")
            if cur.get("is_ready"):
                break
            time.sleep(0.1)
        assert cur["is_ready"]
        s = socket.create_connection(("127.0.0.1", slotless_master["port"]), timeout=10)
        s.sendall((f"GET /proxy/{cmd['id']}/api/kernels/1/channels?session_id=7 HTTP/1.1\r\nHost: x\r\nUpgrade: websocket\r\n"
                   "Connection: Upgrade\r\nSec-WebSocket-Key: abc123==\r\nSec-WebSocket-Version: 13\r\n\r\n").encode())
        head = b""
        while b"\r\n\r\n" not in head:
            head += s.recv(4096)
        text = head.decode()
        assert text.startswith("HTTP/1.1 101") and "X-Path: /api/kernels/1/channels?session_id=7" in text and "X-Key: abc123==" in text
        for msg in (b"hello", b"second frame"):
            s.sendall(msg)
            got = b""
            while len(got) < len(msg):
                got += s.recv(4096)
            assert got == msg.upper()
        s.sendall(b"bye")
        assert s.recv(4096) == b""          # the task closed its side; the tunnel closes ours
        s.close()
    finally:
        c.request("POST", f"/api/v1/commands/{cmd['id']}/kill")


def test_telemetry_sink_is_opt_in(tmp_path):
    """Usage telemetry (reference: master/internal/telemetry, Segment): off unless configured; when on, one JSON line per
    event in a local file."""
    def run_master(extra):
        port_file = tmp_path / f"port-{len(extra)}"
        proc = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp_path / f"data{len(extra)}"), "--artificial-slots", "1",
                                 "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file)] + extra,
                                stdout=open(tmp_path / "master.log", "a"), stderr=subprocess.STDOUT, env=dict(os.environ, DET_CONFIG_DIR=str(tmp_path / "cfg")))
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        return proc, int(port_file.read_text())

    sink = tmp_path / "events.jsonl"
    proc, port = run_master(["--telemetry", json.dumps({"enabled": True, "sink": str(sink)})])
    try:
        os.environ["DET_CONFIG_DIR"] = str(tmp_path / "cfg")
        d = Determined(f"127.0.0.1:{port}")
        cfg = {"entrypoint": "model_def:XorTrial", "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
               "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 10}}, "scheduling_unit": 10,
               "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpts")}}
        exp = d.create_experiment(cfg, MODEL_DIR)
        assert exp.wait(timeout=120) == "COMPLETED"
        time.sleep(0.3)
    finally:
        proc.terminate()
        proc.wait(timeout=5)
    events = [json.loads(l) for l in sink.read_text().splitlines()]
    names = [e["event"] for e in events]
    assert names[0] == "master_started" and "experiment_created" in names and "experiment_state_changed" in names
    created = events[names.index("experiment_created")]
    assert created["properties"]["searcher"] == "single" and created["cluster_id"]
    assert events[names.index("experiment_state_changed")]["properties"]["state"] == "COMPLETED"
    proc, _ = run_master([])
    proc.terminate()
    proc.wait(timeout=5)
    assert not (tmp_path / "data0" / "telemetry.jsonl").exists()


def test_det_tunnel_carries_raw_tcp_to_a_service_task(slotless_master, tmp_path):
    """``det tunnel``: the master answers ``Upgrade: det-tcp`` with 101 and then copies bytes to the task's port, so a task
    that speaks a non-HTTP protocol (sshd for ``det shell``; here a line-echo server) is reachable through the master —
    once over stdin/stdout (ssh ProxyCommand style) and once through ``--listener`` (reference cli/determined_cli/tunnel.py,
    master/internal/proxy/tcp.go)."""
    import socket

    server = tmp_path / "echo.py"
    server.write_text(
        "import socket, sys\n"
        "s = socket.socket(); s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1); s.bind(('127.0.0.1', int(sys.argv[1]))); s.listen(8)\n"
        "while True:\n"
        "    c, _ = s.accept()\n"
        "    c.sendall(b'BANNER echo-1.0\\n')\n"
        "    f = c.makefile('rwb', buffering=0)\n"
        "    for line in f:\n"
        "        f.write(b'echo: ' + line)\n"
        "    c.close()\n")
    c = MasterClient("127.0.0.1", slotless_master["port"])
    cmd = c.request("POST", "/api/v1/commands", {"argv": [sys.executable, str(server), "{port}"], "proxy": True, "description": "echo"})
    try:
        port = cmd["spec"]["proxy_port"]
        for _ in range(100):     # no HTTP readiness for a raw service: wait until the port accepts
            try:
                socket.create_connection(("127.0.0.1", port), timeout=0.2).close()
                break
            except OSError:
                time.sleep(0.1)
        out = subprocess.run([sys.executable, "-m", "determined_b200.cli", "-m", slotless_master["url"], "tunnel", str(cmd["id"])],
                             input=b"hello\nworld\n", capture_output=True, cwd=str(ROOT), timeout=60)
        assert out.returncode == 0, out.stderr.decode()
        assert out.stdout == b"BANNER echo-1.0\necho: hello\necho: world\n"
        lst = subprocess.Popen([sys.executable, "-m", "determined_b200.cli", "-m", slotless_master["url"], "tunnel", str(cmd["id"]), "--listener", "0"],
                               stdout=subprocess.PIPE, stderr=subprocess.STDOUT, cwd=str(ROOT), text=True)
        try:
            line = lst.stdout.readline()
            local = int(line.strip().rsplit(":", 1)[1])
            s = socket.create_connection(("127.0.0.1", local), timeout=10)
            f = s.makefile("rwb", buffering=0)
            assert f.readline() == b"BANNER echo-1.0\n"
            f.write(b"via listener\n")
            assert f.readline() == b"echo: via listener\n"
            s.close()
        finally:
            lst.terminate()
    finally:
        c.request("POST", f"/api/v1/commands/{cmd['id']}/kill")


def test_trial_logs_in_elasticsearch(tmp_path):
    """``logging: {type: elastic}``: trial log rows are bulk-indexed into Elasticsearch and read back through the same REST
    route, position-addressed (reference master/internal/elastic/elastic_trial_logs.go). Runs against tests/fake_elastic.py."""
    import urllib.request

    es = subprocess.Popen([sys.executable, str(ROOT / "tests" / "fake_elastic.py"), "0"], stdout=subprocess.PIPE, text=True)
    es_port = int(es.stdout.readline())
    port_file = tmp_path / "port"
    master = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp_path / "data"), "--artificial-slots", "1",
                               "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file),
                               "--logging", json.dumps({"type": "elastic", "host": "127.0.0.1", "port": es_port, "index": "det-logs-test"})],
                              stdout=open(tmp_path / "master.log", "w"), stderr=subprocess.STDOUT, env=dict(os.environ, DET_CONFIG_DIR=str(tmp_path / "cfg")))
    try:
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        c = MasterClient("127.0.0.1", int(port_file.read_text()))
        rows = [{"trial_id": 7, "message": f"line {i}", "level": "INFO", "rank": i % 2, "stdtype": "stdout"} for i in range(25)]
        c.request("POST", "/api/v1/trial_logs", {"logs": rows[:10]})
        c.request("POST", "/api/v1/trial_logs", {"logs": rows[10:] + [{"trial_id": 8, "message": "other trial", "level": "INFO"}]})
        got = c.request("GET", "/api/v1/trials/7/logs")
        assert [l["message"] for l in got] == [f"line {i}" for i in range(25)]
        page = c.request("GET", "/api/v1/trials/7/logs?offset=20&limit=3")
        assert [l["message"] for l in page] == ["line 20", "line 21", "line 22"]
        assert [l["message"] for l in c.request("GET", "/api/v1/trials/8/logs")] == ["other trial"]
        counts = json.loads(urllib.request.urlopen(f"http://127.0.0.1:{es_port}/_count", timeout=5).read())
        assert counts["indices"] == {"det-logs-test": 26}
        assert not (tmp_path / "data" / "logs" / "trial-7.jsonl").exists()      # nothing was written to the file store
    finally:
        for p in (master, es):
            p.terminate()
            try:
                p.wait(timeout=5)
            except Exception:
                p.kill()


def test_warm_workers_take_over_trial_starts(tmp_path):
    """`--warm-workers eager`: a pre-started harness process per slot is handed the trial (environment + working
    directory) instead of a cold fork; a replacement is started for the next trial; the trial log carries the start-up
    timing line with the warm worker's preload costs."""
    from determined_b200 import _build

    _build.build_core()
    port_file = tmp_path / "port"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp_path / "cfg"))
    proc = subprocess.Popen([str(LIB / "det-master"), "--port", "0", "--data-dir", str(tmp_path / "data"), "--artificial-slots", "1",
                             "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file),
                             "--warm-workers", "eager"], stdout=open(tmp_path / "master.log", "w"), stderr=subprocess.STDOUT, env=env)
    try:
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        os.environ["DET_CONFIG_DIR"] = str(tmp_path / "cfg")
        d = Determined(f"127.0.0.1:{int(port_file.read_text())}")
        cfg = {"entrypoint": "model_def:XorTrial", "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
               "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 10}}, "scheduling_unit": 10,
               "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpts")}}
        timings = []
        for _ in range(2):       # the second trial runs in the same, re-pooled worker
            exp = d.create_experiment(cfg, MODEL_DIR)
            assert exp.wait(timeout=180) == "COMPLETED"
            t = exp.describe()["trials"][0]
            lines = [l["message"] if isinstance(l, dict) else str(l) for l in d.get_trial(t["id"]).logs()]
            line = next(l for l in lines if "startup_timing:" in l)
            timings.append(json.loads(line.split("startup_timing:", 1)[1]))
        for tm in timings:
            assert tm["warm_worker"] is not None and "import_torch" in tm["warm_worker"], tm
            assert tm["imports"] < 0.5, tm            # everything was imported before the trial arrived
        # the SAME process served both: after a clean trial the worker goes back to its slot's pool
        assert [tm["warm_worker"]["trials_served_before"] for tm in timings] == [0, 1], timings
        # an environment that must be in place before the interpreter starts goes down the cold path
        cold = dict(cfg, environment={"environment_variables": ["OMP_NUM_THREADS=1"]})
        exp = d.create_experiment(cold, MODEL_DIR)
        assert exp.wait(timeout=180) == "COMPLETED"
        t = exp.describe()["trials"][0]
        lines = [l["message"] if isinstance(l, dict) else str(l) for l in d.get_trial(t["id"]).logs()]
        tm = json.loads(next(l for l in lines if "startup_timing:" in l).split("startup_timing:", 1)[1])
        assert tm["warm_worker"] is None
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=5)
        except Exception:
            proc.kill()


def test_containers_that_never_become_ready_are_killed(tmp_path):
    """`all_ready_timeout_ms` (reference allReadyTimeout, trial.go:408-418): a trial whose container never brings up a working
    harness — here a start-up hook that hangs — is killed when the timeout passes and takes the restart path; with
    max_restarts 1 the experiment ends in ERROR instead of holding its slot forever."""
    import shutil

    from determined_b200 import _build

    _build.build_core()
    model_dir = tmp_path / "model"
    shutil.copytree(MODEL_DIR, model_dir)
    (model_dir / "startup-hook.sh").write_text("sleep 600\n")
    cfg_file = tmp_path / "master.json"
    cfg_file.write_text(json.dumps({"all_ready_timeout_ms": 1500, "terminate_timeout_ms": 2000}))
    port_file = tmp_path / "port"
    proc = subprocess.Popen([str(LIB / "det-master"), "--config", str(cfg_file), "--port", "0", "--data-dir", str(tmp_path / "data"),
                             "--artificial-slots", "1", "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file)],
                            stdout=open(tmp_path / "master.log", "w"), stderr=subprocess.STDOUT, env=dict(os.environ, DET_CONFIG_DIR=str(tmp_path / "cfg")))
    try:
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        os.environ["DET_CONFIG_DIR"] = str(tmp_path / "cfg")
        d = Determined(f"127.0.0.1:{int(port_file.read_text())}")
        cfg = {"entrypoint": "model_def:XorTrial", "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
               "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 10}}, "max_restarts": 1,
               "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp_path / "ckpts")}}
        exp = d.create_experiment(cfg, str(model_dir))
        t0 = time.time()
        assert exp.wait(timeout=60) == "ERROR"
        assert time.time() - t0 < 30
        log = (tmp_path / "master.log").read_text()
        assert log.count("were not ready within") == 2, log[-1500:]       # the first run and its one restart
    finally:
        proc.terminate()
        try:
            proc.wait(timeout=5)
        except Exception:
            proc.kill()
