"""``det`` command line (``python -m determined_b200.cli`` or the ``det`` script at the repository root).

Verbs follow the reference CLI (``cli/determined_cli/cli.py:40-283``, ``experiment.py:630-924``):
``experiment {list,describe,config,create,activate,pause,cancel,kill,archive,unarchive,delete,wait,label,set,
list-trials,list-checkpoints,download,download-model-def}``, ``trial {describe,logs,kill,download}``,
``checkpoint {describe,download}``, ``task list``, ``preview-search``, ``slot``/``agent``, ``master {config,logs,info}``,
``user``, ``template``, ``model``, ``cmd``/``notebook``/``shell``/``tensorboard``, plus ``master up`` / ``agent detect``
that wrap the native daemons.
"""
from __future__ import annotations

import argparse
import json
import os
import pathlib
import subprocess
import sys
import time
from typing import Any, Dict, List, Optional

import yaml

from ..experimental.client import APIError, Determined, Session, read_context

LIB = pathlib.Path(__file__).resolve().parent.parent / "_lib"
ROOT = pathlib.Path(__file__).resolve().parent.parent.parent


OUTPUT = {"csv": False}     # `det --csv <list command>`: rows as CSV instead of a table (reference render.tabulate_or_csv)


def _table(rows: List[List[Any]], headers: List[str]) -> str:
    if OUTPUT["csv"]:
        import csv
        import io

        buf = io.StringIO()
        w = csv.writer(buf, lineterminator="\n")
        w.writerow(headers)
        w.writerows(rows)
        return buf.getvalue().rstrip("\n")
    try:
        from tabulate import tabulate

        return tabulate(rows, headers, tablefmt="presto")
    except Exception:
        widths = [max(len(str(x)) for x in col) for col in zip(headers, *rows)] if rows else [len(h) for h in headers]
        line = lambda r: " | ".join(str(x).ljust(w) for x, w in zip(r, widths))  # noqa: E731
        return "\n".join([line(headers), "-+-".join("-" * w for w in widths)] + [line(r) for r in rows])


def _load_config(path: str, overrides: Optional[List[str]] = None) -> Dict[str, Any]:
    return _apply_overrides(yaml.safe_load(pathlib.Path(path).read_text()) or {}, overrides)


def _apply_overrides(cfg: Dict[str, Any], overrides: Optional[List[str]] = None) -> Dict[str, Any]:
    for ov in overrides or []:  # --config key.sub=value
        key, _, val = ov.partition("=")
        cur = cfg
        parts = key.split(".")
        for p in parts[:-1]:
            cur = cur.setdefault(p, {})
        cur[parts[-1]] = yaml.safe_load(val)
    return cfg


def _follow_logs(d: Determined, exp_id: int) -> str:
    seen: Dict[int, int] = {}
    while True:
        desc = d.get_experiment(exp_id).describe()
        for t in desc["trials"]:
            off = seen.get(t["id"], 0)
            for l in d._session.get(f"/api/v1/trials/{t['id']}/logs?offset={off}"):
                prefix = f"[trial {t['id']}" + (f" rank={l['rank']}]" if l.get("rank") is not None else "]")
                print(prefix, l["message"], flush=True)
                off += 1
            seen[t["id"]] = off
        if desc["state"] in ("COMPLETED", "CANCELED", "ERROR"):
            return desc["state"]
        time.sleep(1.0)


def git_metadata(model_def: str) -> Dict[str, str]:
    """``experiment create --git`` (reference cli/determined_cli/experiment.py:124-186): record where the model definition
    came from. The directory must be a git work tree with a remote, nothing uncommitted, and a HEAD that has been pushed."""
    import subprocess

    def git(*args: str) -> str:
        p = subprocess.run(["git", "-C", model_def, *args], capture_output=True, text=True)
        if p.returncode != 0:
            raise SystemExit(f"--git: `git {' '.join(args)}` failed in {model_def}: {p.stderr.strip() or p.stdout.strip()}")
        return p.stdout.strip()

    if git("rev-parse", "--is-inside-work-tree") != "true":
        raise SystemExit(f"--git: {model_def} is not inside a git work tree")
    if git("status", "--porcelain", "--", "."):
        raise SystemExit("--git: the model definition has uncommitted changes; commit them (or drop --git)")
    remote = git("config", "--get", "remote.origin.url")
    commit = git("rev-parse", "HEAD")
    on_remote = subprocess.run(["git", "-C", model_def, "branch", "-r", "--contains", commit], capture_output=True, text=True).stdout.strip()
    if not on_remote:
        raise SystemExit("--git: HEAD has not been pushed to the remote; push it (or drop --git)")
    return {"git_remote": remote, "git_commit": commit, "git_committer": git("log", "-1", "--format=%cn <%ce>"),
            "git_commit_date": git("log", "-1", "--format=%cI")}


# ---- experiment ------------------------------------------------------------------------------------------------------
def experiment_create(a: argparse.Namespace) -> int:
    cfg = _load_config(a.config_file, a.config)
    if a.local:
        if not a.test_mode:
            raise SystemExit("local training is only supported in test mode: add --test-mode")
        from .. import _execution, load
        from ..experimental import test_one_batch

        with _execution.local_execution_manager(pathlib.Path(a.model_def)):
            trial_class = load.trial_class_from_entrypoint(cfg["entrypoint"])
            test_one_batch(trial_class=trial_class, config=cfg)
        print("Model definition test succeeded! 🎉")
        return 0
    d = Determined(a.master, a.user)
    if a.test_mode:
        from .._execution import make_test_experiment_config

        cfg = make_test_experiment_config(cfg)
    exp = d.create_experiment(cfg, a.model_def, activate=not a.paused, template=a.template,
                              git=git_metadata(a.model_def) if a.git else None)
    print(f"Created experiment {exp.id}")
    if a.follow or a.test_mode:
        state = _follow_logs(d, exp.id)
        print(f"Experiment {exp.id} finished: {state}")
        if a.test_mode:
            print("Model definition test succeeded! 🎉" if state == "COMPLETED" else "Model definition test failed", flush=True)
        return 0 if state == "COMPLETED" else 1
    return 0


def experiment_list(a: argparse.Namespace) -> int:
    d = Determined(a.master, a.user)
    # like the reference: your own, unarchived experiments unless --all
    exps = d._session.get("/api/v1/experiments" + ("?all=true" if a.all else f"?users={d._session.user}"))
    rows = [[e["id"], e.get("owner"), (e.get("description") or "")[:40], e["state"], f"{100 * (e.get('progress') or 0):.1f}%",
             e["start_time"][:19], (e.get("end_time") or "")[:19], e.get("num_trials")] for e in exps]
    print(_table(rows, ["ID", "Owner", "Description", "State", "Progress", "Start Time", "End Time", "Trials"]))
    return 0


def experiment_describe(a: argparse.Namespace) -> int:
    d = Determined(a.master, a.user)
    out = [d.get_experiment(int(i)).describe(metrics=a.metrics) for i in a.experiment_ids.split(",")]
    if a.json:
        print(json.dumps(out, indent=2))
        return 0
    for e in out:
        print(_table([[e["id"], e["state"], f"{100 * (e.get('progress') or 0):.1f}%", e["start_time"][:19], (e.get("end_time") or "")[:19],
                       e.get("description"), e.get("archived"), ",".join(e.get("labels") or [])]],
                     ["Experiment ID", "State", "Progress", "Start", "End", "Description", "Archived", "Labels"]))
        rows = [[t["id"], t["state"], json.dumps(t["hparams"]), t["total_batches_processed"], t.get("restarts"),
                 (t.get("latest_validation") or {}).get("metrics", {}).get("validation_metrics")] for t in e["trials"]]
        print(_table(rows, ["Trial ID", "State", "H-Params", "Batches", "Restarts", "Latest Validation"]))
    return 0


def _simple_exp(verb: str):
    def run(a: argparse.Namespace) -> int:
        if verb == "delete" and not a.yes:
            # reference cli/experiment.py delete_experiment: irreversible, so it asks unless --yes
            answer = input(f"Deleting experiment {a.experiment_id} removes its metadata, logs and checkpoints for good. Proceed? [y/N] ") \
                if sys.stdin.isatty() else sys.stdin.readline()
            if answer.strip().lower() not in ("y", "yes"):
                print("Aborted: the experiment was not deleted", file=sys.stderr)
                return 1
        exp = Determined(a.master, a.user).get_experiment(a.experiment_id)
        getattr(exp, verb)()
        print(f"{verb.capitalize()}d experiment {a.experiment_id}".replace("eed", "ed"))
        return 0
    return run


def experiment_wait(a: argparse.Namespace) -> int:
    state = Determined(a.master, a.user).get_experiment(a.experiment_id).wait(interval=a.polling_interval)
    print(f"Experiment {a.experiment_id} is {state}")
    return 0 if state == "COMPLETED" else 1


def experiment_config(a: argparse.Namespace) -> int:
    print(yaml.safe_dump(Determined(a.master, a.user).get_experiment(a.experiment_id).describe()["config"], default_flow_style=False))
    return 0


def experiment_set(a: argparse.Namespace) -> int:
    s = Session(a.master, a.user)
    body: Dict[str, Any] = {}
    if a.what == "description":
        body["description"] = a.value
    elif a.what in ("max-slots", "weight", "priority"):
        body["resources"] = {a.what.replace("-", "_"): yaml.safe_load(a.value)}
    elif a.what == "gc-policy":
        raise SystemExit("set gc-policy: pass --save-experiment-best/--save-trial-best/--save-trial-latest at create time")
    s.request("PATCH", f"/api/v1/experiments/{a.experiment_id}", body)
    print(f"Set {a.what} of experiment {a.experiment_id} to {a.value}")
    return 0


def experiment_label(a: argparse.Namespace) -> int:
    s = Session(a.master, a.user)
    labels = set(s.get(f"/api/v1/experiments/{a.experiment_id}").get("labels") or [])
    (labels.add if a.action == "add" else labels.discard)(a.label)
    s.request("PATCH", f"/api/v1/experiments/{a.experiment_id}", {"labels": sorted(labels)})
    return 0


def experiment_list_trials(a: argparse.Namespace) -> int:
    e = Determined(a.master, a.user).get_experiment(a.experiment_id).describe()
    print(_table([[t["id"], t["state"], json.dumps(t["hparams"]), t["start_time"][:19], (t.get("end_time") or "")[:19], t["total_batches_processed"]]
                  for t in e["trials"]], ["Trial ID", "State", "H-Params", "Start Time", "End Time", "# of Batches"]))
    return 0


def experiment_list_checkpoints(a: argparse.Namespace) -> int:
    s = Session(a.master, a.user)
    cks = s.get(f"/api/v1/experiments/{a.experiment_id}/checkpoints")
    print(_table([[c["trial_id"], c["total_batches"], c["state"], c["uuid"], (c.get("end_time") or "")[:19]] for c in cks],
                 ["Trial ID", "# of Batches", "State", "UUID", "End Time"]))
    return 0


def experiment_download_model_def(a: argparse.Namespace) -> int:
    files = Session(a.master, a.user).get(f"/api/v1/experiments/{a.experiment_id}/model_def")
    out = pathlib.Path(a.output_dir or f"experiment_{a.experiment_id}_model_def")
    for f in files:
        p = out / f["path"]
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_text(f["content"])
    print(f"Wrote {len(files)} files to {out}")
    return 0


# ---- trial / checkpoint -------------------------------------------------------------------------------------------------
def trial_describe(a: argparse.Namespace) -> int:
    t = Determined(a.master, a.user).get_trial(a.trial_id).describe()
    if a.json:
        print(json.dumps(t, indent=2))
        return 0
    print(_table([[t["id"], t["experiment_id"], t["state"], json.dumps(t["hparams"]), t["start_time"][:19], (t.get("end_time") or "")[:19]]],
                 ["Trial ID", "Exp ID", "State", "H-Params", "Start Time", "End Time"]))
    rows = []
    vals = {v["total_batches"]: v for v in t.get("validations", [])}
    cks = {c["total_batches"]: c for c in t.get("checkpoints", [])}
    for s in t.get("steps", []):
        v, c = vals.get(s["total_batches"]), cks.get(s["total_batches"])
        rows.append([s["step_id"], s["total_batches"], s["state"], json.dumps(s["metrics"].get("avg_metrics")) if a.metrics else "",
                     json.dumps(v["metrics"]["validation_metrics"]) if v else "", c["uuid"] if c else ""])
    print(_table(rows, ["Step", "# of Batches", "State", "Avg Metrics", "Validation Metrics", "Checkpoint UUID"]))
    return 0


def trial_logs(a: argparse.Namespace) -> int:
    """Reference ``det trial logs``: --head / --tail / --follow and the agent / container / rank / level / stdtype / source /
    timestamp filters; the filters are applied by the master, lines come back with their position so following resumes there."""
    t = Determined(a.master, a.user).get_trial(a.trial_id)
    ranks = list(a.rank_ids or []) + ([a.rank] if a.rank is not None else [])
    filters: Dict[str, Any] = {"agent_ids": a.agent_ids, "container_ids": a.container_ids, "rank_ids": ranks, "stdtypes": a.stdtypes,
                               "sources": a.sources, "level": a.level, "timestamp_before": a.timestamp_before,
                               "timestamp_after": a.timestamp_after}
    if a.head is not None and a.tail is not None:
        print("--head and --tail are mutually exclusive", file=sys.stderr)
        return 1
    off = 0
    if a.head is not None or a.tail is not None:
        lines = t.logs(head=a.head, tail=a.tail, limit=max(a.head or 0, a.tail or 0), **filters)
        for l in lines:
            print(l["message"], flush=True)
        if not a.follow or a.head is not None:
            return 0
        off = lines[-1]["id"] + 1 if lines else 0
    while True:
        lines = t.logs(offset=off, **filters)
        for l in lines:
            print(l["message"], flush=True)
        if lines:
            off = lines[-1]["id"] + 1
            continue        # more may be waiting behind a full page
        if not a.follow or t.describe()["state"] != "ACTIVE":
            return 0
        time.sleep(1.0)


def trial_kill(a: argparse.Namespace) -> int:
    Determined(a.master, a.user).get_trial(a.trial_id).kill()
    print(f"Killed trial {a.trial_id}")
    return 0


def experiment_download(a: argparse.Namespace) -> int:
    """Best ``--top-n`` checkpoints of the experiment, one directory per checkpoint uuid (reference:
    ``cli/determined_cli/experiment.py:384-398``)."""
    exp = Determined(a.master, a.user).get_experiment(a.experiment_id)
    smaller = None if a.smaller_is_better is None else a.smaller_is_better.lower() in ("1", "true", "yes")
    top = pathlib.Path(a.output_dir)
    top.mkdir(parents=True, exist_ok=True)
    for ck in exp.top_n_checkpoints(a.top_n, sort_by=a.sort_by, smaller_is_better=smaller):
        path = ck.download(str(top / ck.uuid))
        print(path if a.quiet else f"Checkpoint {ck.uuid} (trial {ck.trial_id}, batch {ck.batch_number}): {path}")
    return 0


def trial_download(a: argparse.Namespace) -> int:
    t = Determined(a.master, a.user).get_trial(a.trial_id)
    ck = t.select_checkpoint(latest=a.latest, best=a.best, uuid=a.uuid) if (a.latest or a.best or a.uuid) else t.select_checkpoint(latest=True)
    print("Local checkpoint path:", ck.download(a.output_dir))
    return 0


def checkpoint_describe(a: argparse.Namespace) -> int:
    c = Session(a.master, a.user).get(f"/api/v1/checkpoints/{a.uuid}")
    c.pop("experiment_config", None)
    print(json.dumps(c, indent=2))
    return 0


def checkpoint_download(a: argparse.Namespace) -> int:
    print("Local checkpoint path:", Determined(a.master, a.user).get_checkpoint(a.uuid).download(a.output_dir))
    return 0


# ---- cluster --------------------------------------------------------------------------------------------------------------
def task_list(a: argparse.Namespace) -> int:
    tasks = Session(a.master, a.user).get("/api/v1/tasks")
    print(_table([[t["id"], t["name"], t["slots_needed"], t["state"], ",".join(c["agent"] for c in t["containers"])] for t in tasks],
                 ["Task ID", "Name", "Slots Needed", "State", "Agents"]))
    return 0


def preview_search(a: argparse.Namespace) -> int:
    cfg = _load_config(a.config_file)
    out = Session(a.master, a.user).post("/api/v1/searcher/preview", {"searcher": cfg["searcher"], "hyperparameters": cfg.get("hyperparameters", {}),
                                                                       "seed": a.seed})
    print(f"Using search configuration:\n{yaml.safe_dump(cfg['searcher'], default_flow_style=False)}")
    print(_table([[n, k] for k, n in sorted(out["results"].items(), key=lambda kv: -kv[1])], ["Trials", "Breakdown"]))
    return 0


def preview_search_local(a: argparse.Namespace) -> int:
    from ..master import _core

    cfg = _load_config(a.config_file)
    out = _core.simulate(cfg["searcher"], cfg.get("hyperparameters", {}), seed=a.seed, simulation_seed=a.seed, random_order=True)
    print(_table([[n, k] for k, n in sorted(out["results"].items(), key=lambda kv: -kv[1])], ["Trials", "Breakdown"]))
    return 0


def agent_list(a: argparse.Namespace) -> int:
    agents = Session(a.master, a.user).get("/api/v1/agents")
    if a.json:
        print(json.dumps(agents, indent=2))
        return 0
    print(_table([[x["id"], x.get("registered_time", "")[:19], x["num_slots"], x["num_used_slots"], x.get("label"), x.get("enabled")] for x in agents],
                 ["Agent ID", "Registered Time", "Slots", "Used", "Label", "Enabled"]))
    return 0


def slot_list(a: argparse.Namespace) -> int:
    slots = Session(a.master, a.user).get("/api/v1/slots")
    if a.json:
        print(json.dumps([dict(s, slot_id=s["id"]) for s in slots], indent=2))
        return 0
    print(_table([[s["agent_id"], s["id"], s.get("enabled"), s.get("container_id") or "FREE", (s.get("device") or {}).get("type"),
                   (s.get("device") or {}).get("brand")] for s in slots], ["Agent ID", "Slot ID", "Enabled", "Task", "Type", "Device"]))
    return 0


def slot_toggle(verb: str):
    def run(a: argparse.Namespace) -> int:
        Session(a.master, a.user).post(f"/api/v1/agents/{a.agent_id}/slots/{a.slot_id}/{verb}")
        print(f"{verb}d slot {a.slot_id} of agent {a.agent_id}")
        return 0
    return run


def agent_toggle(verb: str):
    def run(a: argparse.Namespace) -> int:
        s = Session(a.master, a.user)
        if bool(a.agent_id) == bool(a.all):
            print("give an agent id or --all", file=sys.stderr)
            return 1
        for agent_id in ([a.agent_id] if a.agent_id else sorted(x["id"] for x in s.get("/api/v1/agents"))):
            s.post(f"/api/v1/agents/{agent_id}/{verb}")
            print(f"{verb.capitalize()}d agent {agent_id}")
        return 0
    return run


def master_info(a: argparse.Namespace) -> int:
    print(json.dumps(Session(a.master, a.user).get("/api/v1/master"), indent=2))
    return 0


def master_config(a: argparse.Namespace) -> int:
    print(yaml.safe_dump(Session(a.master, a.user).get("/api/v1/master/config"), default_flow_style=False))
    return 0


def master_logs(a: argparse.Namespace) -> int:
    s = Session(a.master, a.user)
    lines = s.get(f"/api/v1/master/logs?limit={a.tail or 25000}")
    last = -1
    while True:
        for l in lines:
            print(l["time"], l["level"].upper(), l["message"], flush=True)
        if lines:
            last = lines[-1]["id"]
        if not a.follow:
            return 0
        time.sleep(1.0)
        lines = s.get(f"/api/v1/master/logs?after={last}&limit=5000")


def master_up(a: argparse.Namespace) -> int:
    from .. import _build

    _build.build_core()
    cmd = [str(LIB / "det-master"), "--port", str(a.port), "--data-dir", a.data_dir, "--repo-root", str(ROOT), "--python", sys.executable,
           "--scheduler", a.scheduler]
    if a.artificial_slots:
        cmd += ["--artificial-slots", str(a.artificial_slots)]
    os.execv(cmd[0], cmd)


def agent_detect(a: argparse.Namespace) -> int:
    from .. import _build

    _build.build_core()
    return subprocess.call([str(LIB / "det-agent"), "--detect-only"])


# ---- users / templates / models / commands ----------------------------------------------------------------------------------
def _read_password(prompt: str) -> str:
    """From the terminal without echo; from a pipe, one line of standard input (empty at end of input)."""
    if sys.stdin.isatty():
        import getpass

        return getpass.getpass(prompt)
    try:
        return sys.stdin.readline().rstrip("\n")
    except OSError:          # no usable standard input at all: the blank password
        return ""


def user_login(a: argparse.Namespace) -> int:
    s = Session(a.master, a.username)
    s.login(_read_password(f"Password for user '{a.username}': "))
    print(f"Logged in as {a.username}")
    return 0


def user_list(a: argparse.Namespace) -> int:
    print(_table([[u["username"], u.get("active"), u.get("admin")] for u in Session(a.master, a.user).get("/api/v1/users")], ["Username", "Active", "Admin"]))
    return 0


def user_create(a: argparse.Namespace) -> int:
    Session(a.master, a.user).post("/api/v1/users", {"username": a.username, "admin": a.admin})
    print(f"Created user {a.username}")
    return 0


def user_change_password(a: argparse.Namespace) -> int:
    s = Session(a.master, a.user)
    target = a.username or s.user
    s.request("PATCH", f"/api/v1/users/{target}", {"password": _read_password(f"New password for user '{target}': ")})
    print(f"Changed the password of {target}")
    return 0


def user_logout(a: argparse.Namespace) -> int:
    s = Session(a.master, a.user)
    try:
        s.post("/api/v1/auth/logout")
    except APIError:
        pass                                  # an expired token is as logged out as it gets
    s.forget_token()
    print(f"Logged out {s.user}")
    return 0


def user_whoami(a: argparse.Namespace) -> int:
    s = Session(a.master, a.user)
    try:
        print(f"You are logged in as user '{s.get('/api/v1/auth/user')['username']}'")
    except APIError:
        print(f"You are not logged in (user '{s.user}' has no valid session at {s.master})")
        return 1
    return 0


def user_rename(a: argparse.Namespace) -> int:
    Session(a.master, a.user).request("PATCH", f"/api/v1/users/{a.target_user}", {"username": a.new_username})
    print(f"Renamed {a.target_user} to {a.new_username}")
    return 0


def user_link_agent(a: argparse.Namespace) -> int:
    """Run this user's tasks under a given UID / GID on the agents (reference user.py:link_with_agent_user)."""
    body = {"agent_user_group": {"uid": a.agent_uid, "gid": a.agent_gid, "user": a.agent_user, "group": a.agent_group}}
    Session(a.master, a.user).request("PATCH", f"/api/v1/users/{a.det_username}", body)
    return 0


def show_version(a: argparse.Namespace) -> int:
    import determined_b200

    out = {"client": {"version": determined_b200.__version__}}
    try:
        info = Session(a.master, a.user).get("/api/v1/master")
        out["master"] = {"version": info.get("version"), "cluster_id": info.get("cluster_id"), "master_address": Session(a.master, a.user).master}
        if str(info.get("version")) != determined_b200.__version__:
            print(f"warning: CLI version {determined_b200.__version__} differs from master version {info.get('version')}", file=sys.stderr)
    except APIError as e:
        out["master"] = {"unreachable": str(e)}
    print(json.dumps(out, indent=2))
    return 0


def user_activate(active: bool):
    def run(a: argparse.Namespace) -> int:
        Session(a.master, a.user).request("PATCH", f"/api/v1/users/{a.username}", {"active": active})
        return 0
    return run


def template_list(a: argparse.Namespace) -> int:
    print(_table([[t["name"]] for t in Session(a.master, a.user).get("/api/v1/templates")], ["Name"]))
    return 0


def template_describe(a: argparse.Namespace) -> int:
    print(yaml.safe_dump(Session(a.master, a.user).get(f"/api/v1/templates/{a.name}")["config"], default_flow_style=False))
    return 0


def template_set(a: argparse.Namespace) -> int:
    Session(a.master, a.user).request("PUT", f"/api/v1/templates/{a.name}", {"config": _load_config(a.template_file)})
    print(f"Set template {a.name}")
    return 0


def template_remove(a: argparse.Namespace) -> int:
    Session(a.master, a.user).request("DELETE", f"/api/v1/templates/{a.name}")
    return 0


def model_list(a: argparse.Namespace) -> int:
    print(_table([[m.name, m.description, m.creation_time, m.last_updated_time] for m in Determined(a.master, a.user).get_models()],
                 ["Name", "Description", "Created", "Updated"]))
    return 0


def model_create(a: argparse.Namespace) -> int:
    Determined(a.master, a.user).create_model(a.name, a.description or "")
    print(f"Created model {a.name}")
    return 0


def model_describe(a: argparse.Namespace) -> int:
    m = Determined(a.master, a.user).get_model(a.name)
    print(json.dumps({"name": m.name, "description": m.description, "metadata": m.metadata,
                      "versions": [c.uuid for c in m.get_versions()]}, indent=2))
    return 0


def model_register_version(a: argparse.Namespace) -> int:
    Determined(a.master, a.user).get_model(a.name).register_version(a.uuid)
    return 0


def command_run(kind: str):
    def run(a: argparse.Namespace) -> int:
        s = Session(a.master, a.user)
        argv = list(a.command) if getattr(a, "command", None) else {
            "commands": [],     # the entrypoint then comes from --config-file / --template
            "notebooks": [sys.executable, "-m", "jupyter", "lab", "--ip=127.0.0.1", "--no-browser", "--port={port}", "--ServerApp.token=",
                          "--ServerApp.base_url=" + "{prefix}"],
            "shells": ["/bin/sh", "-c", "sleep infinity"],
            "tensorboards": [sys.executable, "-m", "determined_b200.exec.tensorboard"] + [str(x) for x in getattr(a, "experiment_ids", [])],
        }[kind]
        if argv and argv[0] == "--":
            argv = argv[1:]
        if len(argv) == 1 and any(ch.isspace() for ch in argv[0]):      # det cmd run "sleep 3; touch x": one shell line
            argv = ["/bin/sh", "-c", argv[0]]
        body: Dict[str, Any] = {"argv": argv} if argv else {}
        # task config: --config-file, then --config key=value overrides, then --volume host:container[:ro] bind mounts; the
        # master merges it over --template (reference: cli/command.py parse_config + "--template")
        config: Dict[str, Any] = {}
        if getattr(a, "config_file", None):
            config = _load_config(a.config_file, a.config)
        elif getattr(a, "config", None):
            config = _apply_overrides({}, a.config)
        for vol in getattr(a, "volume", None) or []:
            parts = vol.split(":")
            if len(parts) not in (2, 3) or (len(parts) == 3 and parts[2] not in ("ro", "rw")):
                print(f"--volume wants HOST_PATH:CONTAINER_PATH[:ro], not {vol!r}", file=sys.stderr)
                return 1
            config.setdefault("bind_mounts", []).append({"host_path": os.path.abspath(parts[0]), "container_path": parts[1],
                                                          "read_only": len(parts) == 3 and parts[2] == "ro"})
        if config:
            body["config"] = config
        if kind == "tensorboards":
            if not a.experiment_ids and not a.trial_ids:
                print("a TensorBoard needs at least one experiment id or --trial-ids", file=sys.stderr)
                return 1
            body["experiment_ids"], body["trial_ids"] = list(a.experiment_ids), list(a.trial_ids or [])
        if getattr(a, "template", None):
            body["template"] = a.template
        if getattr(a, "slots", None) is not None:
            body["slots"] = a.slots
        if not config.get("description") and argv:
            body["description"] = " ".join(argv)[:60]
        if getattr(a, "context", None):
            from ..experimental.client import read_context

            try:
                body["context"] = read_context(a.context)
            except ValueError as e:
                print(f"cannot package the context directory: {e}", file=sys.stderr)
                return 1
        out = s.post(f"/api/v1/{kind}", body)
        print(f"Started {kind[:-1]} {out['id']}")
        if out["spec"].get("proxy_port"):
            print(f"Service address (once ready): {s.master}/proxy/{out['id']}/")
        if getattr(a, "detach", False):
            return 0
        seen = 0
        while True:
            c = s.get(f"/api/v1/{kind}/{out['id']}")
            for l in c["logs"][seen:]:
                print(l["message"], flush=True)
            seen = len(c["logs"])
            if c["state"] == "TERMINATED":
                rc = int(c.get("exit_code") or 0)
                print(f"{kind[:-1]} {out['id']} exited successfully" if rc == 0 else f"finished {kind[:-1]} {out['id']}: task failed with exit code {rc}")
                return rc
            time.sleep(0.5)
    return run


def command_list(kind: str):
    def run(a: argparse.Namespace) -> int:
        s = Session(a.master, a.user)
        rows = s.get(f"/api/v1/{kind}" + ("" if a.all else f"?users={s.user}"))
        if getattr(a, "quiet", False):
            print("\n".join(str(c["id"]) for c in rows))
            return 0
        print(_table([[c["id"], c.get("owner", ""), c["state"], c["spec"]["description"], c["start_time"][:19]] for c in rows],
                     ["ID", "Owner", "State", "Description", "Start Time"]))
        return 0
    return run


def command_kill(kind: str):
    def run(a: argparse.Namespace) -> int:
        Session(a.master, a.user).post(f"/api/v1/{kind}/{a.id}/kill")
        return 0
    return run


def command_logs(kind: str):
    def run(a: argparse.Namespace) -> int:
        s = Session(a.master, a.user)
        seen = 0
        while True:
            for l in s.get(f"/api/v1/{kind}/{a.id}/logs?offset={seen}&limit=5000"):
                print(l.get("message", ""), flush=True)
                seen += 1
            if not a.follow or s.get(f"/api/v1/{kind}/{a.id}")["state"] == "TERMINATED":
                return 0
            time.sleep(0.5)
    return run


def command_config(kind: str):
    def run(a: argparse.Namespace) -> int:
        c = Session(a.master, a.user).get(f"/api/v1/{kind}/{a.id}")
        print(json.dumps(c.get("spec", {}).get("config") or c.get("spec", {}), indent=2))
        return 0
    return run


def command_open(kind: str):
    def run(a: argparse.Namespace) -> int:
        s = Session(a.master, a.user)
        c = s.get(f"/api/v1/{kind}/{a.id}")
        if c["state"] == "TERMINATED":
            print(f"{kind[:-1]} {a.id} has terminated", file=sys.stderr)
            return 1
        if kind == "shells":
            print(f"ssh -o ProxyCommand='det -m {s.master} tunnel {a.id}' <user>@{a.id}      # or: det -m {s.master} tunnel {a.id} --listener PORT")
            return 0
        url = f"{s.master}/proxy/{a.id}/"
        print(url)
        if not a.no_browser:
            import webbrowser

            webbrowser.open(url)
        return 0
    return run


# ---- parser ---------------------------------------------------------------------------------------------------------------
def tunnel(a: argparse.Namespace) -> int:
    from . import tunnel as T

    d = Determined(a.master, a.user)
    if a.listener is not None:
        return T.listen_tunnel(d._session.master, a.task_id, a.listener, d._session.token,
                               ready=lambda port: print(f"listening on 127.0.0.1:{port}", flush=True))
    return T.stdio_tunnel(d._session.master, a.task_id, d._session.token)


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="det", description="Determined (B200-native) command line")
    p.add_argument("-m", "--master", default=None, help="master address (default $DET_MASTER or localhost:8080)")
    p.add_argument("-u", "--user", default=None)
    p.add_argument("--csv", action="store_true", help="print list output as CSV")
    sub = p.add_subparsers(dest="group")

    def group(name: str, aliases: List[str] = []) -> Any:
        g = sub.add_parser(name, aliases=aliases)
        return g.add_subparsers(dest="verb")

    def cmd(g: Any, name: str, fn: Any, *args: Any, aliases: List[str] = []) -> argparse.ArgumentParser:
        c = g.add_parser(name, aliases=aliases)
        for names, kw in args:
            c.add_argument(*names, **kw)
        c.set_defaults(func=fn)
        return c

    A = lambda *n, **k: (n, k)  # noqa: E731
    e = group("experiment", ["e"])
    cmd(e, "create", experiment_create, A("config_file"), A("model_def"), A("--local", action="store_true"), A("--test-mode", "--test", action="store_true"),
        A("--paused", action="store_true"), A("--follow", "-f", action="store_true"), A("--template"), A("--config", action="append"),
        A("--git", action="store_true", help="record the git remote / commit of the model definition with the experiment"))
    cmd(e, "list", experiment_list, A("--all", "-a", action="store_true"), aliases=["ls"])
    cmd(e, "describe", experiment_describe, A("experiment_ids"), A("--metrics", action="store_true"), A("--json", action="store_true"))
    cmd(e, "config", experiment_config, A("experiment_id", type=int))
    for verb in ("activate", "pause", "cancel", "kill", "archive", "unarchive"):
        cmd(e, verb, _simple_exp(verb), A("experiment_id", type=int))
    cmd(e, "delete", _simple_exp("delete"), A("experiment_id", type=int), A("--yes", action="store_true", help="do not ask for confirmation"))
    cmd(e, "wait", experiment_wait, A("experiment_id", type=int), A("--polling-interval", type=float, default=1.0))
    cmd(e, "set", experiment_set, A("what", choices=["description", "max-slots", "weight", "priority", "gc-policy"]), A("experiment_id", type=int), A("value"))
    cmd(e, "label", experiment_label, A("action", choices=["add", "remove"]), A("experiment_id", type=int), A("label"))
    cmd(e, "list-trials", experiment_list_trials, A("experiment_id", type=int), aliases=["lt"])
    cmd(e, "list-checkpoints", experiment_list_checkpoints, A("experiment_id", type=int), aliases=["lc"])
    cmd(e, "download", experiment_download, A("experiment_id", type=int), A("--output-dir", "-o", default="checkpoints"), A("--top-n", type=int, default=1),
        A("--sort-by"), A("--smaller-is-better"), A("--quiet", "-q", action="store_true"))
    cmd(e, "download-model-def", experiment_download_model_def, A("experiment_id", type=int), A("--output-dir", "-o"))
    t = group("trial", ["t"])
    cmd(t, "describe", trial_describe, A("trial_id", type=int), A("--metrics", action="store_true"), A("--json", action="store_true"))
    cmd(t, "logs", trial_logs, A("trial_id", type=int), A("--follow", "-f", action="store_true"), A("--rank", type=int),
        A("--head", type=int, help="number of lines to show, counting from the beginning"),
        A("--tail", type=int, help="number of lines to show, counting from the end"),
        A("--agent-id", dest="agent_ids", action="append"), A("--container-id", dest="container_ids", action="append"),
        A("--rank-id", dest="rank_ids", type=int, action="append"), A("--timestamp-before"), A("--timestamp-after"),
        A("--level", help="show lines of this level or higher (TRACE, DEBUG, INFO, WARNING, ERROR, CRITICAL)"),
        A("--source", dest="sources", action="append"), A("--stdtype", dest="stdtypes", action="append"))
    cmd(t, "kill", trial_kill, A("trial_id", type=int))
    cmd(t, "download", trial_download, A("trial_id", type=int), A("--latest", action="store_true"), A("--best", action="store_true"), A("--uuid"), A("--output-dir", "-o"))
    c = group("checkpoint", ["c"])
    cmd(c, "describe", checkpoint_describe, A("uuid"))
    cmd(c, "download", checkpoint_download, A("uuid"), A("--output-dir", "-o"))
    tk = group("task")
    cmd(tk, "list", task_list, aliases=["ls"])
    tn = sub.add_parser("tunnel", help="byte tunnel to a service task through the master (ssh ProxyCommand, or --listener PORT)")
    tn.add_argument("task_id")
    tn.add_argument("--listener", type=int, default=None, help="listen on this local port instead of using stdin / stdout")
    tn.set_defaults(func=tunnel)
    ps = sub.add_parser("preview-search")
    ps.add_argument("config_file"), ps.add_argument("--seed", type=int, default=0), ps.add_argument("--offline", action="store_true")
    ps.set_defaults(func=lambda a: preview_search_local(a) if a.offline else preview_search(a))
    ag = group("agent", ["a"])
    cmd(ag, "list", agent_list, A("--json", action="store_true"), aliases=["ls"])
    cmd(ag, "enable", agent_toggle("enable"), A("agent_id", nargs="?"), A("--all", action="store_true", help="every agent"))
    cmd(ag, "disable", agent_toggle("disable"), A("agent_id", nargs="?"), A("--all", action="store_true", help="every agent"))
    cmd(ag, "detect", agent_detect)
    sl = group("slot", ["s"])
    cmd(sl, "list", slot_list, A("--json", action="store_true"), aliases=["ls"])
    cmd(sl, "enable", slot_toggle("enable"), A("agent_id"), A("slot_id"))
    cmd(sl, "disable", slot_toggle("disable"), A("agent_id"), A("slot_id"))
    ms = group("master")
    cmd(ms, "info", master_info)
    cmd(ms, "config", master_config)
    cmd(ms, "logs", master_logs, A("--tail", type=int), A("--follow", "-f", action="store_true"))
    cmd(ms, "up", master_up, A("--port", type=int, default=8080), A("--data-dir", default="/tmp/determined-master"),
        A("--scheduler", default="fair_share"), A("--artificial-slots", type=int, default=0))
    us = group("user", ["u"])
    cmd(us, "login", user_login, A("username"))
    cmd(us, "list", user_list, aliases=["ls"])
    cmd(us, "create", user_create, A("username"), A("--admin", action="store_true"))
    cmd(us, "change-password", user_change_password, A("username", nargs="?", help="default: the current user"))
    cmd(us, "logout", user_logout)
    cmd(us, "whoami", user_whoami)
    cmd(us, "rename", user_rename, A("target_user"), A("new_username"))
    cmd(us, "link-with-agent-user", user_link_agent, A("det_username"), A("--agent-uid", type=int, required=True),
        A("--agent-gid", type=int, required=True), A("--agent-user", required=True), A("--agent-group", required=True))
    cmd(us, "activate", user_activate(True), A("username"))
    cmd(us, "deactivate", user_activate(False), A("username"))
    tp = group("template", ["tpl"])
    cmd(tp, "list", template_list, aliases=["ls"])
    cmd(tp, "describe", template_describe, A("name"))
    cmd(tp, "set", template_set, A("name"), A("template_file"))
    cmd(tp, "remove", template_remove, A("name"), aliases=["rm"])
    mo = group("model")
    cmd(mo, "list", model_list, aliases=["ls"])
    cmd(mo, "create", model_create, A("name"), A("--description"))
    cmd(mo, "describe", model_describe, A("name"))
    cmd(mo, "register-version", model_register_version, A("name"), A("uuid"))
    for name, kind in (("cmd", "commands"), ("notebook", "notebooks"), ("shell", "shells"), ("tensorboard", "tensorboards")):
        g = group(name)
        extra = [A("command", nargs=argparse.REMAINDER)] if name == "cmd" else ([A("experiment_ids", nargs="*", type=int), A("--trial-ids", "-t", nargs="+", type=int, help="serve these trials only")] if name == "tensorboard" else [])
        cmd(g, "run" if name == "cmd" else "start", command_run(kind), *extra, A("--slots", type=int, default=None), A("--detach", "-d", action="store_true"),
            A("--config-file"), A("--config", action="append", help="key.sub=value override of the task config"), A("--template"),
            A("--context", "-c", help="directory to upload; becomes the task's working directory"),
            A("--volume", "-v", action="append", help="HOST_PATH:CONTAINER_PATH[:ro] bind mount"),
            A("--proxy", action="store_true", help="give the task a port ({port} in the command, $DET_TASK_PORT) reachable at /proxy/<id>/ on the master"))
        cmd(g, "list", command_list(kind), A("--all", "-a", action="store_true", help="tasks of every user, not only your own"),
            A("--quiet", "-q", action="store_true", help="only the ids"), aliases=["ls"])
        cmd(g, "kill", command_kill(kind), A("id"))
        cmd(g, "logs", command_logs(kind), A("id"), A("--follow", "-f", action="store_true"))
        cmd(g, "config", command_config(kind), A("id"))
        if name != "cmd":
            cmd(g, "open", command_open(kind), A("id"), A("--no-browser", action="store_true"))
    ver = sub.add_parser("version", help="client and master versions")
    ver.set_defaults(func=show_version)
    return p


def main(argv: Optional[List[str]] = None) -> int:
    parser = build_parser()
    args = parser.parse_args(argv)
    if not getattr(args, "func", None):
        parser.print_help()
        return 2
    OUTPUT["csv"] = bool(getattr(args, "csv", False))
    try:
        return int(args.func(args) or 0)
    except APIError as e:
        print(f"det: {e}", file=sys.stderr)
        return 1


if __name__ == "__main__":
    sys.exit(main())
