"""System behaviours end to end against the C++ master (the docker-free part of the reference's e2e_tests/tests/test_system.py
and test_users.py): failures surfacing with their message, initial validation, graceful termination, NUL bytes in logs, the
trial-log filters, per-rank logs of a two-slot trial, disabling / enabling slots, delete / archive / labels, listing filters,
test mode, and — on a master that requires authentication — users, passwords, deactivation and per-user listings."""
import json
import os
import pathlib
import subprocess
import sys
import time

import pytest

from determined_b200.experimental.client import APIError, Determined, Session

ROOT = pathlib.Path(__file__).resolve().parent.parent
MASTER = ROOT / "determined_b200" / "_lib" / "det-master"
NOOP_DIR = str(ROOT / "tests" / "fixtures" / "no_op")
XOR_DIR = str(ROOT / "tests" / "fixtures" / "xor_model_def")


def start_master(tmp, *flags):
    from determined_b200 import _build

    _build.build_core()
    port_file = tmp / "port"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg"))
    env.pop("DET_USER", None)
    proc = subprocess.Popen([str(MASTER), "--port", "0", "--data-dir", str(tmp / "data"), "--artificial-slots", "2", "--repo-root", str(ROOT),
                             "--python", sys.executable, "--print-port-file", str(port_file), *flags],
                            stdout=open(tmp / "master.log", "w"), stderr=subprocess.STDOUT, env=env)
    for _ in range(100):
        if port_file.exists() and port_file.read_text().strip():
            break
        time.sleep(0.05)
    return proc, {"url": f"127.0.0.1:{int(port_file.read_text())}", "tmp": tmp, "env": env}


def stop_master(proc):
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()


@pytest.fixture(scope="module")
def master(tmp_path_factory):
    tmp = tmp_path_factory.mktemp("sysmaster")
    proc, info = start_master(tmp)
    old = os.environ.get("DET_CONFIG_DIR")
    os.environ["DET_CONFIG_DIR"] = info["env"]["DET_CONFIG_DIR"]
    yield info
    stop_master(proc)
    if old is None:
        os.environ.pop("DET_CONFIG_DIR", None)
    else:
        os.environ["DET_CONFIG_DIR"] = old


def det(master, *args, stdin=None, user=None, timeout=120):
    pre = ["-u", user] if user else []
    return subprocess.run([sys.executable, "-m", "determined_b200.cli", "-m", master["url"], *pre, *args], capture_output=True, text=True,
                          timeout=timeout, env=master["env"], cwd=str(ROOT), input=stdin)


def noop_config(tmp, batches=20, **hp):
    return {"entrypoint": "model_def:NoOpTrial", "description": "noop",
            "hyperparameters": dict({"global_batch_size": 32, "metrics_progression": "decreasing", "metrics_base": 0.9, "metrics_sigma": 0}, **hp),
            "searcher": {"name": "single", "metric": "validation_error", "max_length": {"batches": batches}},
            "scheduling_unit": 10, "checkpoint_storage": {"type": "shared_fs", "host_path": str(tmp / "ckpts")}, "max_restarts": 0}


def trial_messages(d, exp):
    out = []
    for t in exp.describe()["trials"]:
        out += [l["message"] for l in d.get_trial(t["id"]).logs()]
    return out


@pytest.mark.parametrize("hp,overrides", [
    ({"fail_on_first_validation": "failed on first validation"}, {}),
    ({"fail_on_chechpoint_save": "failed on checkpoint save"}, {}),
    # the checkpoint taken right before the trial closes fails, with a restart allowed
    ({"fail_on_chechpoint_save": "failed on checkpoint save"}, {"max_restarts": 1, "min_validation_period": {"batches": 1},
                                                               "searcher": {"name": "single", "metric": "validation_error", "max_length": {"batches": 1}}}),
])
def test_failures_surface_with_their_message(master, hp, overrides):
    d = Determined(master["url"])
    cfg = noop_config(master["tmp"], **hp)
    cfg.update(overrides)
    exp = d.create_experiment(cfg, NOOP_DIR)
    assert exp.wait(timeout=180) == "ERROR"
    assert any(list(hp.values())[0] in m for m in trial_messages(d, exp))


def test_perform_initial_validation(master):
    d = Determined(master["url"])
    cfg = noop_config(master["tmp"], batches=1)
    cfg["perform_initial_validation"] = True
    exp = d.create_experiment(cfg, NOOP_DIR)
    assert exp.wait(timeout=120) == "COMPLETED"
    vals = sorted(exp.describe(metrics=True)["trials"][0]["validations"], key=lambda v: v["total_batches"])
    assert vals[0]["total_batches"] == 0 and vals[0]["metrics"]["validation_metrics"]["validation_error"] == pytest.approx(1.0)


def test_graceful_trial_termination(master):
    """A grid over request_stop: the trial that asks to stop ends early — and cleanly — while the other trains to the end."""
    d = Determined(master["url"])
    cfg = noop_config(master["tmp"], batches=40, request_stop={"type": "categorical", "vals": [True, False]})
    cfg["searcher"]["name"] = "grid"
    exp = d.create_experiment(cfg, NOOP_DIR)
    assert exp.wait(timeout=180) == "COMPLETED"
    trials = exp.describe()["trials"]
    assert len(trials) == 2 and all(t["state"] == "COMPLETED" for t in trials)
    by_stop = {bool(t["hparams"]["request_stop"]): t["total_batches_processed"] for t in trials}
    assert by_stop[False] == 40 and 0 < by_stop[True] < 40


def test_null_bytes_in_logs_and_log_filters(master):
    d = Determined(master["url"])
    exp = d.create_experiment(noop_config(master["tmp"], batches=1, write_null=True), NOOP_DIR)
    assert exp.wait(timeout=120) == "COMPLETED"
    t = d.get_trial(exp.describe()["trials"][0]["id"])
    logs = t.logs()
    assert len(logs) > 3 and [l["id"] for l in logs] == list(range(len(logs)))
    nul = [l for l in logs if "null byte" in l["message"]]
    assert nul and "text after it" in nul[0]["message"]
    # head / tail / offset
    assert [l["id"] for l in t.logs(head=2)] == [0, 1]
    assert [l["id"] for l in t.logs(tail=2)] == [len(logs) - 2, len(logs) - 1]
    assert [l["id"] for l in t.logs(offset=len(logs) - 1)] == [len(logs) - 1]
    # search, level, stdtype, time window, unknown agent
    assert all("null byte" in l["message"] for l in t.logs(search="null byte")) and t.logs(search="null byte")
    assert t.logs(level="INFO") and not t.logs(level="CRITICAL")
    assert all(l["level"] in ("WARNING", "ERROR", "CRITICAL") for l in t.logs(level="WARNING"))
    assert {l["stdtype"] for l in t.logs(stdtypes=["stdout"])} <= {"stdout"}
    mid = logs[len(logs) // 2]["timestamp"]
    assert all(l["timestamp"] > mid for l in t.logs(timestamp_after=mid))
    assert all(l["timestamp"] < mid for l in t.logs(timestamp_before=mid))
    assert len(t.logs(timestamp_after=mid)) + len(t.logs(timestamp_before=mid)) <= len(logs)
    assert t.logs(agent_ids=["agent-local"]) and not t.logs(agent_ids=["no-such-agent"])
    with pytest.raises(APIError):
        t.logs(level="LOUD")
    # the CLI goes through the same filters
    r = det(master, "trial", "logs", str(t.id), "--head", "3")
    assert r.returncode == 0 and r.stdout.splitlines() == [l["message"] for l in logs[:3]]
    r = det(master, "trial", "logs", str(t.id), "--tail", "2")
    assert r.returncode == 0 and r.stdout.splitlines() == [l["message"] for l in logs[-2:]]
    r = det(master, "trial", "logs", str(t.id))
    assert r.returncode == 0 and len(r.stdout.splitlines()) == len(logs)
    assert det(master, "trial", "logs", str(t.id), "--head", "1", "--tail", "1").returncode != 0


def test_distributed_logging_carries_ranks(master):
    d = Determined(master["url"])
    cfg = {"entrypoint": "model_def:XorTrial", "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
           "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 2}}, "scheduling_unit": 1,
           "resources": {"slots_per_trial": 2}, "checkpoint_storage": {"type": "shared_fs", "host_path": str(master["tmp"] / "ckpts")},
           "max_restarts": 0}
    exp = d.create_experiment(cfg, XOR_DIR)
    assert exp.wait(timeout=240) == "COMPLETED", trial_messages(d, exp)[-20:]
    t = d.get_trial(exp.describe()["trials"][0]["id"])
    fields = d._session.get(f"/api/v1/trials/{t.id}/logs/fields")
    assert set(fields["rank_ids"]) >= {0, 1}
    for rank in (0, 1):
        lines = t.logs(rank_ids=[rank])
        assert lines and {l["rank"] for l in lines} == {rank}
    assert len(t.logs(rank_ids=[0, 1])) == len(t.logs(rank_ids=[0])) + len(t.logs(rank_ids=[1]))
    r = det(master, "trial", "logs", str(t.id), "--rank-id", "1")
    assert r.returncode == 0 and len(r.stdout.splitlines()) == len(t.logs(rank_ids=[1]))


def test_disable_and_enable_slots(master):
    d = Determined(master["url"])
    slots = json.loads(det(master, "slot", "list", "--json").stdout)
    assert len(slots) == 2
    for s in slots:
        assert det(master, "slot", "disable", s["agent_id"], str(s["slot_id"])).returncode == 0
    assert all(not s["enabled"] for s in json.loads(det(master, "slot", "list", "--json").stdout))
    exp = d.create_experiment(noop_config(master["tmp"], batches=10), NOOP_DIR)
    time.sleep(2.5)
    desc = exp.describe()
    assert desc["state"] == "ACTIVE" and all(t["total_batches_processed"] == 0 for t in desc["trials"])     # nowhere to run
    for s in slots:
        assert det(master, "slot", "enable", s["agent_id"], str(s["slot_id"])).returncode == 0
    assert exp.wait(timeout=120) == "COMPLETED"


def test_delete_archive_labels_and_listing_filters(master):
    d = Determined(master["url"])
    s = d._session
    cfg = noop_config(master["tmp"], batches=10)
    cfg["description"] = "listing-target"
    paused = d.create_experiment(cfg, NOOP_DIR, activate=False)
    # only finished experiments can be archived
    assert det(master, "experiment", "archive", str(paused.id)).returncode != 0
    assert not json.loads(det(master, "experiment", "describe", "--json", str(paused.id)).stdout)[0]["archived"]
    paused.activate()
    assert paused.wait(timeout=120) == "COMPLETED"
    assert det(master, "experiment", "archive", str(paused.id)).returncode == 0
    assert json.loads(det(master, "experiment", "describe", "--json", str(paused.id)).stdout)[0]["archived"]
    assert paused.id not in [e["id"] for e in s.get("/api/v1/experiments")]
    assert paused.id in [e["id"] for e in s.get("/api/v1/experiments?archived=true")]
    assert det(master, "experiment", "unarchive", str(paused.id)).returncode == 0
    assert paused.id in [e["id"] for e in s.get("/api/v1/experiments")]
    # labels
    label = "__det_test_dummy_label__"
    assert det(master, "experiment", "label", "add", str(paused.id), label).returncode == 0
    assert label in det(master, "experiment", "describe", str(paused.id)).stdout
    assert [e["id"] for e in s.get(f"/api/v1/experiments?labels={label}")] == [paused.id]
    assert det(master, "experiment", "label", "remove", str(paused.id), label).returncode == 0
    assert label not in det(master, "experiment", "describe", str(paused.id)).stdout
    assert s.get(f"/api/v1/experiments?labels={label}") == []
    # filters, sorting, paging
    assert [e["id"] for e in s.get("/api/v1/experiments?description=listing-target")] == [paused.id]
    assert all(e["state"] == "COMPLETED" for e in s.get("/api/v1/experiments?states=COMPLETED"))
    other = d.create_experiment(noop_config(master["tmp"], batches=10), NOOP_DIR, activate=False)       # a second row to sort and page over
    ids = [e["id"] for e in s.get("/api/v1/experiments?sort_by=id&order_by=desc")]
    assert ids == sorted(ids, reverse=True) and len(ids) >= 2
    assert [e["id"] for e in s.get("/api/v1/experiments?sort_by=id&order_by=desc&limit=1&offset=1")] == ids[1:2]
    assert s.get("/api/v1/experiments?users=nobody") == []
    with pytest.raises(APIError):
        s.get("/api/v1/experiments?sort_by=colour")
    other.cancel()
    # delete: gone from the API, and the CLI says so
    assert det(master, "experiment", "delete", str(paused.id), "--yes").returncode == 0
    assert det(master, "experiment", "describe", str(paused.id)).returncode != 0
    with pytest.raises(APIError):
        s.get(f"/api/v1/experiments/{paused.id}")


def test_create_in_test_mode(master, tmp_path):
    cfg = tmp_path / "noop.yaml"
    cfg.write_text(json.dumps(noop_config(master["tmp"], batches=500)))
    r = det(master, "experiment", "create", "--test-mode", str(cfg), NOOP_DIR, timeout=240)
    assert r.returncode == 0 and "Model definition test succeeded" in r.stdout, r.stdout + r.stderr
    bad = tmp_path / "bad.yaml"
    bad.write_text(json.dumps(noop_config(master["tmp"], batches=500, fail_on_first_validation="broken on purpose")))
    r = det(master, "experiment", "create", "--test-mode", str(bad), NOOP_DIR, timeout=240)
    assert r.returncode != 0


# ---- users (a master that requires authentication) -------------------------------------------------------------------
@pytest.fixture(scope="module")
def auth_master(tmp_path_factory):
    tmp = tmp_path_factory.mktemp("authmaster")
    proc, info = start_master(tmp, "--auth-required")
    yield info
    stop_master(proc)


def test_users_passwords_deactivation_and_per_user_listings(auth_master):
    m = auth_master
    # the default accounts have blank passwords: the CLI signs in by itself
    r = det(m, "user", "whoami")
    assert r.returncode == 0 and "determined" in r.stdout, r.stdout + r.stderr
    # only an administrator manages users
    assert det(m, "user", "create", "alice").returncode != 0
    assert det(m, "user", "create", "alice", user="admin").returncode == 0
    assert det(m, "user", "create", "bob", user="admin").returncode == 0
    assert det(m, "user", "deactivate", "bob").returncode != 0
    # alice sets her own password; nobody but her or an administrator can
    assert det(m, "user", "login", "alice", stdin="\n").returncode == 0
    assert det(m, "user", "change-password", stdin="s3cret\n", user="alice").returncode == 0
    assert det(m, "user", "change-password", "alice", stdin="hijack\n", user="determined").returncode != 0
    assert det(m, "user", "logout", user="alice").returncode == 0
    assert det(m, "user", "login", "alice", stdin="wrong\n").returncode != 0
    assert det(m, "user", "whoami", user="alice").returncode != 0          # no session, and the blank password no longer works
    assert det(m, "user", "login", "alice", stdin="s3cret\n").returncode == 0
    r = det(m, "user", "whoami", user="alice")
    assert r.returncode == 0 and "alice" in r.stdout
    # experiments and tasks are listed per user unless --all
    cfg = pathlib.Path(m["tmp"]) / "noop.yaml"
    cfg.write_text(json.dumps(noop_config(m["tmp"], batches=10)))
    r = det(m, "experiment", "create", str(cfg), NOOP_DIR, user="alice")
    assert r.returncode == 0, r.stdout + r.stderr
    r = det(m, "cmd", "run", "--detach", "echo", "alices-command", user="alice")
    assert r.returncode == 0, r.stdout + r.stderr
    assert "noop" in det(m, "experiment", "list", user="alice").stdout
    assert "noop" not in det(m, "experiment", "list", user="determined").stdout
    assert "noop" in det(m, "experiment", "list", "--all", user="determined").stdout
    assert "alices-command" in det(m, "cmd", "list", user="alice").stdout
    assert "alices-command" not in det(m, "cmd", "list", user="determined").stdout
    assert "alices-command" in det(m, "cmd", "list", "--all", user="determined").stdout
    # deactivation ends alice's session and refuses her password; activation brings both back
    assert det(m, "user", "deactivate", "alice", user="admin").returncode == 0
    assert det(m, "experiment", "list", user="alice").returncode != 0
    assert det(m, "user", "login", "alice", stdin="s3cret\n").returncode != 0
    assert det(m, "user", "activate", "alice", user="admin").returncode == 0
    assert det(m, "user", "login", "alice", stdin="s3cret\n").returncode == 0
    assert det(m, "experiment", "list", user="alice").returncode == 0
    # an administrator can reset anyone's password
    assert det(m, "user", "change-password", "alice", stdin="reset\n", user="admin").returncode == 0
    assert det(m, "user", "login", "alice", stdin="reset\n").returncode == 0
    listed = det(m, "user", "list", user="admin").stdout
    assert "alice" in listed and "bob" in listed


@pytest.mark.skipif(os.geteuid() != 0, reason="switching users needs an agent that runs as root")
def test_tasks_run_as_the_linked_agent_user(master):
    """`det user link-with-agent-user`: the owner's tasks run under that UID / GID on the agents (reference
    e2e_tests/tests/test_users.py test_link_with_agent_user / test_non_root_shell)."""
    s = Session(master["url"])
    r = det(master, "cmd", "run", "id")
    assert r.returncode == 0 and "uid=0" in r.stdout
    r = det(master, "user", "link-with-agent-user", "determined", "--agent-uid", "65534", "--agent-gid", "65534", "--agent-user", "nobody",
            "--agent-group", "nogroup")
    assert r.returncode == 0, r.stdout + r.stderr
    try:
        assert s.get("/api/v1/users/determined")["agent_user_group"]["uid"] == 65534
        r = det(master, "cmd", "run", "sh", "-c", "id; echo user=$USER")
        assert r.returncode == 0 and "uid=65534" in r.stdout and "gid=65534" in r.stdout and "user=nobody" in r.stdout, r.stdout + r.stderr
        assert "groups=65534" in r.stdout and "0(root)" not in r.stdout           # root's supplementary groups are gone too
    finally:
        s.request("PATCH", "/api/v1/users/determined", {"agent_user_group": None})
    r = det(master, "cmd", "run", "id")
    assert r.returncode == 0 and "uid=0" in r.stdout


# ---- master configuration file ------------------------------------------------------------------------------------------
def run_master_with_config(tmp, cfg):
    """Start det-master from a configuration file; returns (process, url) or (None, error text) when it refuses the file."""
    from determined_b200 import _build

    _build.build_core()
    tmp.mkdir(parents=True, exist_ok=True)
    path, port_file = tmp / "master.json", tmp / "port"
    path.write_text(json.dumps(dict({"port": 0, "data_dir": str(tmp / "data"), "artificial_slots": 1, "repo_root": str(ROOT), "python": sys.executable}, **cfg)))
    log = tmp / "master.log"
    proc = subprocess.Popen([str(MASTER), "--config", str(path), "--print-port-file", str(port_file)], stdout=open(log, "w"), stderr=subprocess.STDOUT,
                            env=dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg")))
    for _ in range(100):
        if port_file.exists() and port_file.read_text().strip():
            return proc, f"127.0.0.1:{int(port_file.read_text())}"
        if proc.poll() is not None:
            return None, log.read_text()
        time.sleep(0.05)
    proc.kill()
    return None, "the master did not start: " + log.read_text()


@pytest.mark.parametrize("bad,why", [
    ({"task_container_defaults": {"nccl_port_range": "12-15"}}, "MIN:MAX"),
    ({"task_container_defaults": {"gloo_port_range": "200:100"}}, "minimum exceeds"),
    ({"task_container_defaults": {"shm_size_bytes": 0}}, "shm_size_bytes"),
    ({"task_container_defaults": {"colour": "red"}}, "task_container_defaults.colour"),
    ({"log": {"level": "chatty"}}, "log.level"),
    ({"tensorboard_timeout": -1}, "tensorboard_timeout"),
    ({"security": {"default_task": {"user": "nobody"}}}, "default_task"),
])
def test_master_refuses_bad_configuration(tmp_path, bad, why):
    proc, text = run_master_with_config(tmp_path / "m", bad)
    if proc is not None:
        stop_master(proc)
    assert proc is None and why in text, text


def test_master_configuration_keys_of_the_reference_are_honoured(tmp_path):
    """A master.yaml in the reference's shape: log level, CORS, task container defaults reaching the trial's environment,
    a short tensorboard_timeout reaping an idle TensorBoard;
    db / root / harness_path name things this master does not have and are accepted."""
    import urllib.request

    cfg = {"log": {"level": "error", "color": False}, "enable_cors": True, "tensorboard_timeout": 2, "cluster_name": "cfg-test",
           "db": {"user": "postgres", "host": "db", "port": "5432"}, "root": "/usr/share/determined/master", "harness_path": "/opt/determined",
           "scheduler": {"type": "priority", "fit": "worst"},
           "task_container_defaults": {"dtrain_network_interface": "lo", "nccl_port_range": "12000:12100", "gloo_port_range": "13000:13100",
                                       "shm_size_bytes": 134217728}}
    proc, url = run_master_with_config(tmp_path / "m", cfg)
    assert proc is not None, url
    try:
        os.environ["DET_CONFIG_DIR"] = str(tmp_path / "m" / "cfg")
        s = Session(url)
        info = s.get("/info")
        assert info["cluster_name"] == "cfg-test" and info["scheduler"] == "priority"
        shown = s.get("/api/v1/master/config")
        assert shown["fitting_policy"] == "worst" and shown["tensorboard_timeout"] == 2 and shown["log"]["level"] == "error"
        # log.level error: the usual start-up chatter is not in the master's log
        assert "master listening" not in (tmp_path / "m" / "master.log").read_text()
        # CORS: the caller's origin is allowed, pre-flights are answered
        req = urllib.request.Request(f"http://{url}/info", headers={"Origin": "http://ui.example"})
        with urllib.request.urlopen(req, timeout=5) as resp:
            assert resp.headers["Access-Control-Allow-Origin"] == "http://ui.example" and "Authorization" in resp.headers["Access-Control-Allow-Headers"]
        req = urllib.request.Request(f"http://{url}/api/v1/experiments", method="OPTIONS", headers={"Origin": "http://ui.example"})
        with urllib.request.urlopen(req, timeout=5) as resp:
            assert resp.status == 204 and "PATCH" in resp.headers["Access-Control-Allow-Methods"]
        # the defaults reach a trial's environment (a command prints what a trial of the same master would get is not enough:
        # run a trial whose model definition checks them)
        model = tmp_path / "model"
        model.mkdir()
        (model / "model_def.py").write_text((pathlib.Path(NOOP_DIR) / "model_def.py").read_text() + '''

import os as _os

assert _os.environ["DET_TRIAL_RUNNER_NETWORK_INTERFACE"] == "lo" and _os.environ["NCCL_SOCKET_IFNAME"] == "lo" and _os.environ["GLOO_SOCKET_IFNAME"] == "lo"
assert _os.environ["NCCL_PORT_RANGE"] == "12000:12100" and _os.environ["GLOO_PORT_RANGE"] == "13000:13100"
assert int(_os.environ["DET_SHM_SIZE"]) == 134217728
''')
        d = Determined(url)
        exp = d.create_experiment(noop_config(tmp_path, batches=1), str(model))
        assert exp.wait(timeout=120) == "COMPLETED", trial_messages(d, exp)[-10:]
        # an idle TensorBoard is reaped after tensorboard_timeout seconds
        done = exp
        tb = s.post("/api/v1/tensorboards", {"argv": ["sleep", "600"], "experiment_ids": [done.id], "proxy": False})
        t0 = time.time()
        while time.time() - t0 < 20 and s.get(f"/api/v1/tensorboards/{tb['id']}")["state"] != "TERMINATED":
            time.sleep(0.25)
        assert s.get(f"/api/v1/tensorboards/{tb['id']}")["state"] == "TERMINATED" and time.time() - t0 >= 1.5
    finally:
        stop_master(proc)


@pytest.mark.skipif(os.geteuid() != 0, reason="switching users needs an agent that runs as root")
def test_security_default_task_is_the_identity_of_unlinked_users_tasks(tmp_path):
    proc, url = run_master_with_config(tmp_path / "m", {"security": {"default_task": {"uid": 65534, "gid": 65534, "user": "nobody", "group": "nogroup"}}})
    assert proc is not None, url
    try:
        os.environ["DET_CONFIG_DIR"] = str(tmp_path / "m" / "cfg")
        s = Session(url)
        cmd = s.post("/api/v1/commands", {"argv": ["id", "-u"]})
        for _ in range(100):
            cur = s.get(f"/api/v1/commands/{cmd['id']}")
            if cur["state"] == "TERMINATED":
                break
            time.sleep(0.1)
        assert [l["message"] for l in cur["logs"]] == ["65534"]
        # a linked user keeps their own identity
        s.request("PATCH", "/api/v1/users/determined", {"agent_user_group": {"uid": 0, "gid": 0, "user": "root", "group": "root"}})
        cmd = s.post("/api/v1/commands", {"argv": ["id", "-u"]})
        for _ in range(100):
            cur = s.get(f"/api/v1/commands/{cmd['id']}")
            if cur["state"] == "TERMINATED":
                break
            time.sleep(0.1)
        assert [l["message"] for l in cur["logs"]] == ["0"]
    finally:
        stop_master(proc)


def test_master_logs_follow_and_agent_toggles(master):
    s = Session(master["url"])
    lines = s.get("/api/v1/master/logs?limit=5")
    assert lines and [l["id"] for l in lines] == sorted(l["id"] for l in lines)
    newest = lines[-1]["id"]
    assert s.get(f"/api/v1/master/logs?after={newest}") == []
    d = Determined(master["url"])
    d.create_experiment(noop_config(master["tmp"], batches=1), NOOP_DIR).wait(timeout=120)        # makes the master log something
    newer = s.get(f"/api/v1/master/logs?after={newest}")
    assert newer and all(l["id"] > newest for l in newer) and any("created experiment" in l["message"] for l in newer)
    r = det(master, "master", "logs", "--tail", "3")
    assert r.returncode == 0 and len(r.stdout.splitlines()) == 3
    # --follow keeps printing until it is stopped
    p = subprocess.Popen([sys.executable, "-m", "determined_b200.cli", "-m", master["url"], "master", "logs", "--tail", "1", "-f"], stdout=subprocess.PIPE,
                         text=True, env=master["env"], cwd=str(ROOT))
    try:
        time.sleep(3.0)
        d.create_experiment(noop_config(master["tmp"], batches=1), NOOP_DIR, activate=False).cancel()
        time.sleep(3.0)
    finally:
        p.terminate()
    out = p.communicate(timeout=10)[0]
    assert len(out.splitlines()) >= 2 and "created experiment" in out
    # agent enable / disable: one agent or --all
    assert det(master, "agent", "disable").returncode != 0
    r = det(master, "agent", "disable", "--all")
    assert r.returncode == 0 and "Disabled agent agent-local" in r.stdout
    assert all(not a["enabled"] for a in json.loads(det(master, "agent", "list", "--json").stdout))
    r = det(master, "agent", "enable", "agent-local")
    assert r.returncode == 0 and all(a["enabled"] for a in json.loads(det(master, "agent", "list", "--json").stdout))
