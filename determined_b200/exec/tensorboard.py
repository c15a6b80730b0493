"""TensorBoard task entrypoint: serve the synced event files of the given experiments.
Parity: reference ``harness/determined/exec/tensorboard.py:12-92`` (launch ``tensorboard --logdir``, wait until event
files are visible or the process dies)."""
import os
import subprocess
import sys
import time


def main() -> int:
    root = os.environ.get("DET_TENSORBOARD_ROOT", "/tmp/determined-checkpoints")
    cluster = os.environ.get("DET_CLUSTER_ID", "")
    exp_ids = sys.argv[1:]
    dirs = []
    for base, sub, _files in os.walk(root):
        if base.endswith("/tensorboard/experiment") or base.endswith("/tensorboard/experiment/"):
            for e in sub:
                if not exp_ids or e in exp_ids:
                    dirs.append(os.path.join(base, e))
    logdir = ",".join(dirs) if dirs else os.path.join(root, cluster, "tensorboard")
    port = os.environ.get("DET_TENSORBOARD_PORT", "6006")
    proc = subprocess.Popen([sys.executable, "-m", "tensorboard.main", "--logdir", logdir, "--port", port, "--host", "127.0.0.1"])
    deadline = time.time() + 600
    while time.time() < deadline and proc.poll() is None:
        if any(any("tfevents" in f for f in fs) for d in dirs for _b, _s, fs in os.walk(d)):
            print(f"TensorBoard serving {logdir} on port {port}", flush=True)
            break
        time.sleep(1)
    return proc.wait()


if __name__ == "__main__":
    sys.exit(main())
