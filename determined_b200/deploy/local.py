"""``det-deploy local``: run a master (with its in-process agent) and optional extra agents as background processes on
this box, with pid / log files under a state directory.

Role of the reference's ``deploy/determined_deploy/local`` (docker-compose: Postgres + master + agent containers,
``cluster-up / cluster-down / master-up / master-down / agent-up / agent-down / logs``, ``--master-config-path``). There is no database or container runtime to start here:
the master keeps its own journal and launches harness processes directly.
"""
from __future__ import annotations

import argparse
import json
import os
import pathlib
import signal
import subprocess
import sys
import time
import urllib.request
from typing import List, Optional

ROOT = pathlib.Path(__file__).resolve().parent.parent.parent
LIB = ROOT / "determined_b200" / "_lib"


def _state_dir(name: str) -> pathlib.Path:
    d = pathlib.Path(os.environ.get("DET_DEPLOY_DIR", pathlib.Path.home() / ".local" / "share" / "determined_b200")) / name
    d.mkdir(parents=True, exist_ok=True)
    return d


def _alive(pid: int) -> bool:
    try:
        os.kill(pid, 0)
        return True
    except OSError:
        return False


def _read_pid(f: pathlib.Path) -> Optional[int]:
    try:
        pid = int(f.read_text().strip())
        return pid if _alive(pid) else None
    except Exception:
        return None


def _master_config_flag(d: pathlib.Path, config_path: Optional[str]) -> List[str]:
    """``--master-config-path``: a master.yaml (or JSON) in the keys of docs/cluster-config.md, handed to det-master as JSON
    (reference ``det-deploy local --master-config-path`` mounts the file as /etc/determined/master.yaml)."""
    if not config_path:
        return []
    import yaml

    cfg = yaml.safe_load(pathlib.Path(config_path).read_text()) or {}
    if not isinstance(cfg, dict):
        raise SystemExit(f"{config_path}: a master configuration is a mapping")
    out = d / "master-config.json"
    out.write_text(json.dumps(cfg))
    return ["--config", str(out)]


def master_up(name: str, port: int, slots: int = 0, scheduler: str = "fair_share", extra: Optional[List[str]] = None,
              config_path: Optional[str] = None) -> int:
    from determined_b200 import _build

    _build.build_core()
    d = _state_dir(name)
    if _read_pid(d / "master.pid"):
        raise SystemExit(f"cluster {name!r} already has a running master (pid {_read_pid(d / 'master.pid')})")
    port_file = d / "master.port"
    port_file.unlink(missing_ok=True)
    # flags win over the file inside det-master, so only what the caller chose (or the file leaves open) is passed as a flag
    file_flag = _master_config_flag(d, config_path)
    in_file = json.loads((d / "master-config.json").read_text()) if file_flag else {}
    cmd = [str(LIB / "det-master"), *file_flag, "--repo-root", str(ROOT), "--python", sys.executable, "--print-port-file", str(port_file)]
    if "port" not in in_file or port != 8080:
        cmd += ["--port", str(port)]
    if "data_dir" not in in_file:
        cmd += ["--data-dir", str(d / "data")]
    if "scheduler" not in in_file or scheduler != "fair_share":
        cmd += ["--scheduler", scheduler]
    if slots:
        cmd += ["--artificial-slots", str(slots)]
    cmd += extra or []
    proc = subprocess.Popen(cmd, stdout=open(d / "master.log", "a"), stderr=subprocess.STDOUT, start_new_session=True)
    (d / "master.pid").write_text(str(proc.pid))
    for _ in range(200):
        if port_file.exists() and port_file.read_text().strip():
            break
        if proc.poll() is not None:
            raise SystemExit(f"master exited with code {proc.returncode}; see {d / 'master.log'}")
        time.sleep(0.05)
    bound = int(port_file.read_text())
    for _ in range(100):
        try:
            urllib.request.urlopen(f"http://127.0.0.1:{bound}/info", timeout=1).read()
            break
        except Exception:
            time.sleep(0.05)
    return bound


def agent_up(name: str, master_port: int, agent_id: str, slots: int = 0, master_host: str = "127.0.0.1", label: str = "") -> int:
    from determined_b200 import _build

    _build.build_core()
    d = _state_dir(name)
    if _read_pid(d / f"agent-{agent_id}.pid"):
        raise SystemExit(f"agent {agent_id!r} of cluster {name!r} is already running (pid {_read_pid(d / f'agent-{agent_id}.pid')})")
    cmd = [str(LIB / "det-agent"), "--master-host", master_host, "--master-port", str(master_port), "--agent-id", agent_id,
           "--repo-root", str(ROOT), "--python", sys.executable]
    if slots:
        cmd += ["--artificial-slots", str(slots)]
    if label:
        cmd += ["--label", label]
    proc = subprocess.Popen(cmd, stdout=open(d / f"agent-{agent_id}.log", "a"), stderr=subprocess.STDOUT, start_new_session=True)
    (d / f"agent-{agent_id}.pid").write_text(str(proc.pid))
    return proc.pid


def grpc_up(name: str, master_port: int, grpc_port: int) -> int:
    """The gRPC front (determined.api.v1.Determined) as a sidecar of the master; stopped with the cluster."""
    d = _state_dir(name)
    pf = d / "grpc.port"
    pf.unlink(missing_ok=True)
    proc = subprocess.Popen([sys.executable, "-m", "determined_b200.master.grpc_gateway", "--master", f"127.0.0.1:{master_port}",
                             "--port", str(grpc_port), "--print-port-file", str(pf)], cwd=str(ROOT),
                            stdout=open(d / "grpc.log", "a"), stderr=subprocess.STDOUT, start_new_session=True)
    (d / "agent-grpc-gateway.pid").write_text(str(proc.pid))
    for _ in range(200):
        if pf.exists() and pf.read_text().strip():
            return int(pf.read_text())
        if proc.poll() is not None:
            raise SystemExit(f"grpc gateway exited with code {proc.returncode}; see {d / 'grpc.log'}")
        time.sleep(0.05)
    return grpc_port


def cluster_down(name: str, pattern: str = "*.pid") -> List[int]:
    """Stop the processes whose pid files match ``pattern`` (everything by default; ``master.pid`` / ``agent-<id>.pid`` for
    master-down / agent-down)."""
    d = _state_dir(name)
    stopped = []
    for f in sorted(d.glob(pattern), key=lambda p: p.name.startswith("master")):   # agents first, master last
        pid = _read_pid(f)
        if pid:
            try:
                os.killpg(pid, signal.SIGTERM)
            except OSError:
                try:
                    os.kill(pid, signal.SIGTERM)
                except OSError:
                    pass
            for _ in range(100):
                if not _alive(pid):
                    break
                time.sleep(0.05)
            stopped.append(pid)
        f.unlink(missing_ok=True)
    return stopped


def status(name: str) -> dict:
    d = _state_dir(name)
    out = {"name": name, "dir": str(d), "master_pid": _read_pid(d / "master.pid"), "agents": {}}
    try:
        out["port"] = int((d / "master.port").read_text())
    except Exception:
        out["port"] = None
    for f in d.glob("agent-*.pid"):
        out["agents"][f.stem[len("agent-"):]] = _read_pid(f)
    return out


def main(argv: Optional[List[str]] = None) -> None:
    ap = argparse.ArgumentParser(prog="det-deploy", description="bring a determined_b200 cluster up or down: local processes, AWS CloudFormation, GCP Terraform")
    sub = ap.add_subparsers(dest="flavour", required=True)
    local = sub.add_parser("local")
    lsub = local.add_subparsers(dest="verb", required=True)
    up = lsub.add_parser("cluster-up")
    up.add_argument("--cluster-name", default="determined")
    up.add_argument("--master-port", type=int, default=8080)
    up.add_argument("--agents", type=int, default=0, help="extra det-agent processes besides the master's in-process agent")
    up.add_argument("--artificial-slots", type=int, default=0, help="CPU slots instead of detected GPUs (testing)")
    up.add_argument("--scheduler", default="fair_share", choices=["fair_share", "priority", "round_robin"])
    up.add_argument("--grpc-port", type=int, default=-1, help="also start the gRPC gateway on this port (0 = any free port)")
    up.add_argument("--master-config-path", default=None, help="master configuration file (YAML or JSON, keys of docs/cluster-config.md)")
    down = lsub.add_parser("cluster-down")
    down.add_argument("--cluster-name", default="determined")
    mu = lsub.add_parser("master-up", help="only the master (with its in-process agent unless the configuration says otherwise)")
    mu.add_argument("--cluster-name", default="determined")
    mu.add_argument("--master-port", type=int, default=8080)
    mu.add_argument("--master-config-path", default=None)
    mu.add_argument("--artificial-slots", type=int, default=0)
    mu.add_argument("--scheduler", default="fair_share", choices=["fair_share", "priority", "round_robin"])
    md = lsub.add_parser("master-down")
    md.add_argument("--cluster-name", default="determined")
    au = lsub.add_parser("agent-up", help="one det-agent that registers with a master, here or elsewhere")
    au.add_argument("master_host")
    au.add_argument("--master-port", type=int, default=8080)
    au.add_argument("--agent-name", default="det-agent")
    au.add_argument("--agent-label", default="")
    au.add_argument("--artificial-slots", type=int, default=0)
    au.add_argument("--cluster-name", default="determined")
    ad = lsub.add_parser("agent-down")
    ad.add_argument("--agent-name", default="det-agent")
    ad.add_argument("--all", action="store_true", help="every agent of the cluster")
    ad.add_argument("--cluster-name", default="determined")
    st = lsub.add_parser("status")
    st.add_argument("--cluster-name", default="determined")
    lg = lsub.add_parser("logs")
    lg.add_argument("--cluster-name", default="determined")
    from . import cloud

    cloud.add_parsers(sub)
    args = ap.parse_args(argv)
    if args.flavour != "local":
        print(json.dumps(cloud.run(args), indent=1))
        return
    if args.verb == "cluster-up":
        port = master_up(args.cluster_name, args.master_port, args.artificial_slots, args.scheduler, config_path=args.master_config_path)
        for i in range(args.agents):
            agent_up(args.cluster_name, port, f"agent-{i + 1}", args.artificial_slots)
        extra = {"grpc": f"127.0.0.1:{grpc_up(args.cluster_name, port, args.grpc_port)}"} if args.grpc_port >= 0 else {}
        print(json.dumps({"master": f"127.0.0.1:{port}", "ui": f"http://127.0.0.1:{port}/ui", **extra, **status(args.cluster_name)}))
    elif args.verb == "cluster-down":
        print(json.dumps({"stopped": cluster_down(args.cluster_name)}))
    elif args.verb == "master-up":
        port = master_up(args.cluster_name, args.master_port, args.artificial_slots, args.scheduler, config_path=args.master_config_path)
        print(json.dumps({"master": f"127.0.0.1:{port}", "ui": f"http://127.0.0.1:{port}/ui", **status(args.cluster_name)}))
    elif args.verb == "master-down":
        print(json.dumps({"stopped": cluster_down(args.cluster_name, "master.pid")}))
    elif args.verb == "agent-up":
        pid = agent_up(args.cluster_name, args.master_port, args.agent_name, args.artificial_slots, args.master_host, args.agent_label)
        print(json.dumps({"agent": args.agent_name, "pid": pid}))
    elif args.verb == "agent-down":
        gone = cluster_down(args.cluster_name, "agent-*.pid" if args.all else f"agent-{args.agent_name}.pid")
        print(json.dumps({"stopped": gone}))
    elif args.verb == "status":
        print(json.dumps(status(args.cluster_name)))
    elif args.verb == "logs":
        f = _state_dir(args.cluster_name) / "master.log"
        print(f.read_text() if f.exists() else "")


if __name__ == "__main__":
    main()
