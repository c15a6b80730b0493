#!/usr/bin/env python
"""BASELINE config #5 end to end: the C++ master on one box schedules a cifar10_pytorch `adaptive_asha` search
(1-slot trials, packed onto every GPU) TOGETHER with a 2-slot data-parallel experiment, real harness processes, real
GPUs. Reports wall-clock, trials run, slot occupancy over time and the best validation error.

    python scripts/gpu_asha_search.py [--max-trials 16] [--epochs 4] [--records 4096] [--slots N]

With no GPUs (`--slots N` given) it runs the same flow on artificial CPU slots — used to smoke-test the script here."""
from __future__ import annotations

import argparse
import json
import os
import pathlib
import subprocess
import sys
import tempfile
import time

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import yaml  # noqa: E402

from determined_b200.experimental.client import Determined  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-trials", type=int, default=16)
    ap.add_argument("--epochs", type=int, default=4)
    ap.add_argument("--records", type=int, default=4096)
    ap.add_argument("--slots", type=int, default=0, help="artificial CPU slots (0 = detect GPUs)")
    ap.add_argument("--timeout", type=int, default=1500)
    ap.add_argument("--warm-workers", default="eager", help="off | gpu | lazy | eager (pre-started harness process per slot)")
    ap.add_argument("--settle", type=float, default=20.0, help="seconds to let eager warm workers finish preloading before submitting")
    ap.add_argument("--tag", default="", help="suffix of the output file name")
    args = ap.parse_args()
    from determined_b200 import _build

    _build.build_core()
    tmp = pathlib.Path(tempfile.mkdtemp(prefix="det-asha-"))
    port_file = tmp / "port"
    cmd = [str(ROOT / "determined_b200" / "_lib" / "det-master"), "--port", "0", "--data-dir", str(tmp / "data"), "--repo-root", str(ROOT),
           "--python", sys.executable, "--print-port-file", str(port_file), "--scheduler", "fair_share", "--warm-workers", args.warm_workers]
    if args.slots:
        cmd += ["--artificial-slots", str(args.slots)]
    env = dict(os.environ, DET_CONFIG_DIR=str(tmp / "cfg"))
    os.environ["DET_CONFIG_DIR"] = str(tmp / "cfg")
    master = subprocess.Popen(cmd, stdout=open(tmp / "master.log", "w"), stderr=subprocess.STDOUT, env=env)
    try:
        for _ in range(200):
            if port_file.exists() and port_file.read_text().strip():
                break
            time.sleep(0.05)
        d = Determined(f"127.0.0.1:{int(port_file.read_text())}")
        slots = d._session.get("/api/v1/slots") if hasattr(d, "_session") else None
        ex_dir = ROOT / "examples" / "computer_vision" / "cifar10_pytorch"
        cfg = yaml.safe_load((ex_dir / "adaptive.yaml").read_text())
        cfg["data"] = {"train_size": args.records, "validation_size": 1024}
        cfg["records_per_epoch"] = args.records
        cfg["searcher"].update({"max_length": {"epochs": args.epochs}, "max_trials": args.max_trials})
        cfg["hyperparameters"]["global_batch_size"] = {"type": "int", "minval": 32, "maxval": 64}
        cfg["checkpoint_storage"] = {"type": "shared_fs", "host_path": str(tmp / "ckpts")}
        cfg["min_validation_period"] = {"epochs": 1}
        dist_cfg = yaml.safe_load((ex_dir / "distributed.yaml").read_text())
        dist_cfg["data"] = cfg["data"]
        dist_cfg["records_per_epoch"] = args.records
        dist_cfg["resources"] = {"slots_per_trial": 2}
        dist_cfg["hyperparameters"]["global_batch_size"] = 128
        dist_cfg["searcher"]["max_length"] = {"epochs": args.epochs}
        dist_cfg["checkpoint_storage"] = cfg["checkpoint_storage"]
        if args.warm_workers == "eager":
            time.sleep(args.settle)
        t0 = time.time()
        asha = d.create_experiment(cfg, str(ex_dir))
        two = d.create_experiment(dist_cfg, str(ex_dir))
        occupancy = []
        states = {}
        while time.time() - t0 < args.timeout:
            try:
                sl = d._session.get("/api/v1/slots")
                items = sl.get("slots", sl) if isinstance(sl, dict) else sl
                busy = sum(1 for s in items if s.get("container_id"))
                occupancy.append((round(time.time() - t0, 1), busy, len(items)))
            except Exception:
                pass
            states = {"asha": asha.describe()["state"], "two_slot": two.describe()["state"]}
            if all(s in ("COMPLETED", "CANCELED", "ERROR") for s in states.values()):
                break
            time.sleep(1.0)
        wall = time.time() - t0
        desc = asha.describe(metrics=True)
        trials = desc["trials"]
        best = None
        for t in trials:
            for v in t.get("validations", []):
                m = v["metrics"]["validation_metrics"].get("validation_error")
                if m is not None and (best is None or m < best):
                    best = m
        d2 = two.describe()
        # trial start latency: one startup_timing line per trial run (exec/harness.py::_report_startup)
        starts = []
        for t in trials + d2["trials"]:
            try:
                for l in d.get_trial(t["id"]).logs():
                    msg = l["message"] if isinstance(l, dict) else str(l)
                    if "startup_timing:" in msg:
                        starts.append(dict(json.loads(msg.split("startup_timing:", 1)[1]), trial_id=t["id"]))
            except Exception:
                pass
        def _mean(key):
            vals = [s_[key] for s_ in starts if isinstance(s_.get(key), (int, float))]
            return round(sum(vals) / len(vals), 3) if vals else None
        latency = {"runs": len(starts), "warm_runs": sum(1 for s_ in starts if s_.get("warm_worker")),
                   "mean_s": {k: _mean(k) for k in ("assigned_to_entry", "imports", "env_comm_storage", "build_controller", "controller_built_to_first_batch",
                                                    "assigned_to_first_batch", "first_workload", "assigned_to_first_workload_done")},
                   "max_assigned_to_first_batch_s": max((s_.get("assigned_to_first_batch", 0.0) for s_ in starts), default=None),
                   "max_assigned_to_first_workload_done_s": max((s_["assigned_to_first_workload_done"] for s_ in starts), default=None),
                   "note": "first_workload includes the workload's own batches (a full scheduling unit of training)"}
        out = {"config": "cifar10_pytorch adaptive_asha + one 2-slot experiment on one box", "states": states,
               "wall_clock_s": round(wall, 1), "asha_trials": len(trials), "asha_trials_completed": sum(t["state"] == "COMPLETED" for t in trials),
               "asha_total_batches": sum(t.get("total_batches_processed", 0) for t in trials),
               "two_slot_trial_batches": sum(t.get("total_batches_processed", 0) for t in d2["trials"]),
               "best_validation_error": best, "max_busy_slots": max((b for _, b, _ in occupancy), default=None),
               "total_slots": occupancy[-1][2] if occupancy else None,
               "mean_busy_slots": round(sum(b for _, b, _ in occupancy) / max(len(occupancy), 1), 2),
               "max_trials": args.max_trials, "epochs": args.epochs, "records_per_epoch": args.records,
               "warm_workers": args.warm_workers, "start_latency": latency}
        print(json.dumps(out), flush=True)
        os.makedirs(ROOT / "gpurun_out", exist_ok=True)
        (ROOT / "gpurun_out" / f"asha_search{args.tag}.json").write_text(json.dumps(dict(out, occupancy=occupancy, starts=starts), indent=1))
        ok = all(s == "COMPLETED" for s in states.values())
        if not ok:
            print((tmp / "master.log").read_text()[-4000:])
        sys.exit(0 if ok else 1)
    finally:
        master.terminate()
        try:
            master.wait(timeout=5)
        except Exception:
            master.kill()


if __name__ == "__main__":
    main()
