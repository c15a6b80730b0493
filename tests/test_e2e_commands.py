"""Command tasks end to end through the CLI against the C++ master (the reference's e2e_tests/tests/command/test_run.py and
template/test_template.py cases that do not need docker): start / exit status, context upload with .detignore and the size
limit, task configs and their validation, templates, bind mounts, killing a command that never got resources."""
import json
import os
import pathlib
import subprocess
import sys
import time

import pytest

from determined_b200.experimental.client import Session

ROOT = pathlib.Path(__file__).resolve().parent.parent
MASTER = ROOT / "determined_b200" / "_lib" / "det-master"


@pytest.fixture(scope="module")
def master(tmp_path_factory):
    from determined_b200 import _build

    _build.build_core()
    tmp = tmp_path_factory.mktemp("cmdmaster")
    port_file = tmp / "port"
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg"))
    proc = subprocess.Popen([str(MASTER), "--port", "0", "--data-dir", str(tmp / "data"), "--artificial-slots", "2", "--repo-root", str(ROOT),
                             "--python", sys.executable, "--print-port-file", str(port_file)],
                            stdout=open(tmp / "master.log", "w"), stderr=subprocess.STDOUT, env=env)
    for _ in range(100):
        if port_file.exists() and port_file.read_text().strip():
            break
        time.sleep(0.05)
    port = int(port_file.read_text())
    yield {"url": f"127.0.0.1:{port}", "tmp": tmp, "env": env}
    proc.terminate()
    try:
        proc.wait(timeout=5)
    except Exception:
        proc.kill()


def det(master, *args, timeout=60):
    return subprocess.run([sys.executable, "-m", "determined_b200.cli", "-m", master["url"], *args], capture_output=True, text=True,
                          timeout=timeout, env=master["env"], cwd=str(ROOT))


def test_cold_and_warm_start_and_exit_status(master):
    for _ in range(3):
        r = det(master, "cmd", "run", "echo", "hello", "world")
        assert r.returncode == 0, r.stdout + r.stderr
        assert "hello world" in r.stdout and "exited successfully" in r.stdout
    r = det(master, "cmd", "run", "false")
    assert r.returncode == 1 and "task failed with exit code 1" in r.stdout
    r = det(master, "cmd", "run", "sleep 0.1; exit 7")          # one shell line
    assert r.returncode == 7 and "task failed with exit code 7" in r.stdout


def test_context_upload_detignore_and_size_limit(master, tmp_path):
    tree = tmp_path / "ctx"
    tree.mkdir()
    (tree / "hello.py").write_text("print('hello from the context')\n")
    (tree / "hello-link.py").symlink_to(tree / "hello.py")
    for script in ("hello.py", "hello-link.py"):
        r = det(master, "cmd", "run", "--context", str(tree), sys.executable, script)
        assert r.returncode == 0 and "hello from the context" in r.stdout, r.stdout + r.stderr
    r = det(master, "cmd", "run", "--context", str(tmp_path / "non-existent-path-here"), sys.executable, "hello.py")
    assert r.returncode != 0 and "not a directory" in r.stderr
    # a startup hook in the context runs in the task's shell before the command
    (tree / "startup-hook.sh").write_text("export HOOK_SAYS=ran\n")
    r = det(master, "cmd", "run", "--context", str(tree), "sh", "-c", "echo hook:$HOOK_SAYS")
    assert r.returncode == 0 and "hook:ran" in r.stdout, r.stdout + r.stderr
    # 120 MB (sparse) does not fit in a context ...
    big = tree / "large-file.bin"
    with open(big, "w") as f:
        f.seek(1024 * 1024 * 120)
        f.write("\0")
    r = det(master, "cmd", "run", "--context", str(tree), sys.executable, "hello.py")
    assert r.returncode != 0 and ".detignore" in r.stderr
    # ... unless .detignore leaves it out
    (tree / ".detignore").write_text("*.bin\n")
    r = det(master, "cmd", "run", "--context", str(tree), sys.executable, "hello.py")
    assert r.returncode == 0 and "hello from the context" in r.stdout, r.stdout + r.stderr


def test_task_configs(master, tmp_path):
    cfg = tmp_path / "config.yaml"
    cfg.write_text("resources:\n  slots: 1\nenvironment:\n  environment_variables:\n   - TEST=TEST\n")
    check = "import os; assert os.environ['TEST'] == 'TEST'; print('slots', os.environ.get('DET_SLOT_IDS'))"
    r = det(master, "cmd", "run", "--config-file", str(cfg), sys.executable, "-c", check)
    assert r.returncode == 0, r.stdout + r.stderr
    # --config overrides on top of the file; the description reaches the listing
    r = det(master, "cmd", "run", "--config-file", str(cfg), "--config", "description=from-override", "--config",
            "environment.environment_variables=[TEST=OTHER]", sys.executable, "-c", "import os; print('T=' + os.environ['TEST'])")
    assert r.returncode == 0 and "T=OTHER" in r.stdout, r.stdout + r.stderr
    assert "from-override" in det(master, "cmd", "list").stdout
    # the entrypoint may come from the config alone
    cfg.write_text("entrypoint: [echo, from, the, config]\n")
    r = det(master, "cmd", "run", "--config-file", str(cfg))
    assert r.returncode == 0 and "from the config" in r.stdout, r.stdout + r.stderr
    # mistakes are refused at submission, not at run time
    for bad, why in (("nonsense: 1\n", "unknown field"), ("bind_mounts:\n - host_path: relative\n   container_path: x\n", "absolute"),
                     ("resources:\n  slots: -1\n", "slots")):
        cfg.write_text(bad)
        r = det(master, "cmd", "run", "--config-file", str(cfg), "true")
        assert r.returncode != 0 and why in (r.stdout + r.stderr), (bad, r.stdout, r.stderr)
    r = det(master, "cmd", "run")
    assert r.returncode != 0 and "argv" in (r.stdout + r.stderr)


def test_bind_mounts_absolute_and_relative(master, tmp_path):
    host = tmp_path / "host_dir"
    host.mkdir()
    (host / "payload.txt").write_text("mounted\n")
    r = det(master, "cmd", "run", "--volume", f"{host}:data:ro", "cat", "data/payload.txt")        # relative: under the working directory
    assert r.returncode == 0 and "mounted" in r.stdout, r.stdout + r.stderr
    target = tmp_path / "abs_mount_point"
    r = det(master, "cmd", "run", "--volume", f"{host}:{target}", "cat", f"{target}/payload.txt")
    assert r.returncode == 0 and "mounted" in r.stdout, r.stdout + r.stderr
    r = det(master, "cmd", "run", "--volume", "nonsense", "true")
    assert r.returncode != 0 and "HOST_PATH:CONTAINER_PATH" in r.stderr


def test_templates_apply_to_commands_shells_and_notebooks(master, tmp_path):
    tpl = tmp_path / "tpl.yaml"
    tpl.write_text("description: templated\nenvironment:\n  environment_variables:\n    FROM_TEMPLATE: tpl\n    SHARED: template\n")
    assert det(master, "template", "set", "cmd-tpl", str(tpl)).returncode == 0
    assert "cmd-tpl" in det(master, "template", "list").stdout
    r = det(master, "cmd", "run", "--template", "cmd-tpl", "--config", "environment.environment_variables.SHARED=mine",
            "sh", "-c", "echo $FROM_TEMPLATE/$SHARED")
    assert r.returncode == 0 and "tpl/mine" in r.stdout, r.stdout + r.stderr       # the submitted config wins over the template
    r = det(master, "cmd", "run", "--template", "no-such-template", "true")
    assert r.returncode != 0 and "not found" in (r.stdout + r.stderr)
    s = Session(master["url"])
    for kind, cli in (("shells", "shell"), ("notebooks", "notebook")):
        r = det(master, cli, "start", "--template", "cmd-tpl", "--detach")
        assert r.returncode == 0, r.stdout + r.stderr
        task_id = int(r.stdout.split()[2])
        row = s.get(f"/api/v1/{kind}/{task_id}")
        assert row["spec"]["config"]["environment"]["environment_variables"]["FROM_TEMPLATE"] == "tpl"
        assert json.loads(det(master, cli, "config", str(task_id)).stdout)["description"] == "templated"
        assert det(master, cli, "kill", str(task_id)).returncode == 0
    assert det(master, "template", "remove", "cmd-tpl").returncode == 0


def test_killed_pending_command_terminates(master):
    s = Session(master["url"])
    # far more slots than the cluster has: it can never be scheduled
    r = det(master, "cmd", "run", "--detach", "--config", "resources.slots=1048576", "sleep", "infinity")
    assert r.returncode == 0, r.stdout + r.stderr
    task_id = int(r.stdout.split()[2])
    for _ in range(4):
        assert s.get(f"/api/v1/commands/{task_id}")["state"] == "PENDING"
        time.sleep(0.5)
    assert det(master, "cmd", "kill", str(task_id)).returncode == 0
    for _ in range(20):
        if s.get(f"/api/v1/commands/{task_id}")["state"] == "TERMINATED":
            break
        time.sleep(0.25)
    assert s.get(f"/api/v1/commands/{task_id}")["state"] == "TERMINATED"
    # and the queue is clean: a command that fits still runs
    assert det(master, "cmd", "run", "true").returncode == 0


def test_tensorboard_for_a_shared_fs_experiment(master):
    """`det tensorboard start <experiment>` serves the experiment's synced event files behind /proxy/<id>/ (reference
    e2e_tests/tests/command/test_tensorboard.py): the master resolves where the files live from the experiment's storage
    section, TensorBoard is told its URL prefix, and the scalars of the trial are there."""
    import urllib.request

    from determined_b200.experimental.client import Determined

    pytest.importorskip("tensorboard")
    d = Determined(master["url"])
    cfg = {"entrypoint": "model_def:XorTrial", "hyperparameters": {"global_batch_size": 8, "hidden_size": 4, "learning_rate": 0.5},
           "searcher": {"name": "single", "metric": "validation_loss", "max_length": {"batches": 20}}, "scheduling_unit": 10,
           "checkpoint_storage": {"type": "shared_fs", "host_path": str(master["tmp"] / "ckpts")}, "max_restarts": 0}
    exp = d.create_experiment(cfg, str(ROOT / "tests" / "fixtures" / "xor_model_def"))
    assert exp.wait(timeout=180) == "COMPLETED"
    trial_id = exp.describe()["trials"][0]["id"]
    assert det(master, "tensorboard", "start", "--detach").returncode != 0                 # nothing to serve
    assert det(master, "tensorboard", "start", "--detach", "424242").returncode != 0       # no such experiment
    r = det(master, "tensorboard", "start", "--detach", str(exp.id))
    assert r.returncode == 0, r.stdout + r.stderr
    task_id = int(r.stdout.split()[2])
    s = Session(master["url"])
    try:
        sources = __import__("json").loads(s.get(f"/api/v1/tensorboards/{task_id}")["spec"]["env"]["DET_TENSORBOARD_SOURCES"])
        assert sources[0]["experiment_id"] == exp.id and sources[0]["prefix"].endswith(f"/tensorboard/experiment/{exp.id}")
        deadline = time.time() + 120
        while time.time() < deadline and not s.get(f"/api/v1/tensorboards/{task_id}").get("is_ready"):
            assert s.get(f"/api/v1/tensorboards/{task_id}")["state"] != "TERMINATED", s.get(f"/api/v1/tensorboards/{task_id}")["logs"][-5:]
            time.sleep(0.5)
        assert s.get(f"/api/v1/tensorboards/{task_id}")["is_ready"]
        base = f"http://{master['url']}/proxy/{task_id}"
        tags = {}
        while time.time() < deadline and not tags:                    # TensorBoard loads runs in the background
            with urllib.request.urlopen(base + "/data/plugin/scalars/tags", timeout=10) as resp:
                tags = __import__("json").loads(resp.read())
            time.sleep(0.5)
        run = [k for k in tags if k.startswith(f"experiment-{exp.id}") and k.endswith(f"trial/{trial_id}")]
        assert run and "Determined/loss" in tags[run[0]], tags
        with urllib.request.urlopen(base + "/", timeout=10) as resp:
            assert b"TensorBoard" in resp.read()
    finally:
        det(master, "tensorboard", "kill", str(task_id))
    # one trial only
    r = det(master, "tensorboard", "start", "--detach", "--trial-ids", str(trial_id))
    assert r.returncode == 0, r.stdout + r.stderr
    tid2 = int(r.stdout.split()[2])
    src = __import__("json").loads(s.get(f"/api/v1/tensorboards/{tid2}")["spec"]["env"]["DET_TENSORBOARD_SOURCES"])
    assert src[0]["trial_id"] == trial_id and src[0]["prefix"].endswith(f"/trial/{trial_id}")
    det(master, "tensorboard", "kill", str(tid2))
