"""TensorBoard task entrypoint: serve the synced event files of the given experiments / trials.

The master hands over ``DET_TENSORBOARD_SOURCES``: one entry per experiment with its storage section and the key prefix of
its event files (``<cluster_id>/tensorboard/experiment/<id>[/trial/<t>]``). Shared-fs sources are served in place; S3 / GCS
sources are mirrored into a local directory by a fetch loop that keeps running beside TensorBoard, so curves of running
trials keep growing. Without the variable (a hand-started task) every experiment found under ``DET_TENSORBOARD_ROOT`` is
served.

Parity: reference ``harness/determined/exec/tensorboard.py:12-92`` (launch ``tensorboard --logdir``, wait until event files
are visible or the process dies) and the per-backend fetchers under ``harness/determined/tensorboard/fetchers``."""
import json
import os
import pathlib
import subprocess
import sys
import tempfile
import threading
import time
from typing import Any, Dict, List, Tuple


def _legacy_dirs(root: str, exp_ids: List[str]) -> List[str]:
    dirs = []
    for base, sub, _files in os.walk(root):
        if base.rstrip("/").endswith("/tensorboard/experiment"):
            dirs += [os.path.join(base, e) for e in sub if not exp_ids or e in exp_ids]
    return dirs


class Fetcher:
    """Mirrors the objects under ``prefix`` into ``local``; only new or grown objects are downloaded again."""

    def __init__(self, client: Any, prefix: str, local: pathlib.Path) -> None:
        self.client, self.prefix, self.local = client, prefix, local
        self._sizes: Dict[str, int] = {}

    def fetch_new(self) -> int:
        got = 0
        for key, size in self.client.list(self.prefix):
            if self._sizes.get(key) == size or key.endswith("/"):
                continue
            dst = self.local / key[len(self.prefix):].lstrip("/")
            dst.parent.mkdir(parents=True, exist_ok=True)
            tmp = dst.with_name(dst.name + ".part")
            self.client.get_file(key, str(tmp))
            os.replace(tmp, dst)                # TensorBoard never sees a half-written event file
            self._sizes[key] = size
            got += 1
        return got


def resolve_sources(sources: List[Dict[str, Any]], scratch: pathlib.Path) -> Tuple[List[Tuple[str, str]], List[Fetcher]]:
    """``[(run name, directory)]`` to serve and the fetchers that keep the remote ones current."""
    from determined_b200.tensorboard.build import object_store_for

    runs: List[Tuple[str, str]] = []
    fetchers: List[Fetcher] = []
    for src in sources:
        cfg, prefix = src.get("storage") or {}, src["prefix"].strip("/")
        name = "experiment-" + str(src.get("experiment_id")) + (f"-trial-{src['trial_id']}" if src.get("trial_id") else "")
        kind = cfg.get("type", "shared_fs")
        if kind == "shared_fs":
            root = pathlib.Path(cfg.get("container_path") or cfg.get("host_path") or "/tmp/determined-checkpoints")
            if cfg.get("storage_path"):
                root = root / cfg["storage_path"]
            runs.append((name, str(root / prefix)))
            continue
        client, key_prefix = object_store_for(cfg)
        if client is None:
            print(f"TensorBoard: storage type {kind!r} of {name} cannot be read from here; skipping it", flush=True)
            continue
        local = scratch / name
        local.mkdir(parents=True, exist_ok=True)
        fetchers.append(Fetcher(client, key_prefix + prefix + "/", local))
        runs.append((name, str(local)))
    return runs, fetchers


def main() -> int:
    port = os.environ.get("DET_TENSORBOARD_PORT", "6006")
    raw = os.environ.get("DET_TENSORBOARD_SOURCES", "")
    fetchers: List[Fetcher] = []
    if raw:
        scratch = pathlib.Path(tempfile.mkdtemp(prefix="det-tensorboard-"))
        runs, fetchers = resolve_sources(json.loads(raw), scratch)
        if not runs:
            print("TensorBoard: nothing to serve", flush=True)
            return 1
        for f in fetchers:
            f.fetch_new()
        dirs = [d for _n, d in runs]
        logdir_args = ["--logdir_spec", ",".join(f"{n}:{d}" for n, d in runs)]
    else:
        root = os.environ.get("DET_TENSORBOARD_ROOT", "/tmp/determined-checkpoints")
        dirs = _legacy_dirs(root, sys.argv[1:])
        logdir_args = ["--logdir", ",".join(dirs) if dirs else os.path.join(root, os.environ.get("DET_CLUSTER_ID", ""), "tensorboard")]
    prefix = os.environ.get("DET_PROXY_PREFIX", "")
    cmd = [sys.executable, "-m", "tensorboard.main", *logdir_args, "--port", port, "--host", "127.0.0.1", "--reload_interval", "5"]
    if prefix:
        cmd += ["--path_prefix", prefix.rstrip("/")]
    proc = subprocess.Popen(cmd)

    def keep_fetching() -> None:
        while proc.poll() is None:
            time.sleep(float(os.environ.get("DET_TENSORBOARD_FETCH_INTERVAL", "10")))
            for f in fetchers:
                try:
                    f.fetch_new()
                except Exception as e:  # a failed poll is retried on the next round
                    print(f"TensorBoard: fetch failed: {e}", flush=True)

    if fetchers:
        threading.Thread(target=keep_fetching, daemon=True).start()
    deadline = time.time() + 600
    while time.time() < deadline and proc.poll() is None:
        if any(any("tfevents" in f for f in fs) for d in dirs for _b, _s, fs in os.walk(d)):
            print(f"TensorBoard serving {' '.join(logdir_args)} on port {port}", flush=True)
            break
        time.sleep(1)
    return proc.wait()


if __name__ == "__main__":
    sys.exit(main())
