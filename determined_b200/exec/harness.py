"""Trial runner entrypoint (``python -m determined_b200.exec.harness``), launched by the master per allocation.

Reads the ``DET_*`` environment contract, validates checkpoint storage by a write/read/delete round trip, restores
the latest checkpoint, builds the layer pipeline (master channel → workload manager → rank broadcast → controller)
and runs it. With more than one slot it first re-executes itself once per GPU (``RANK`` / ``LOCAL_RANK`` /
``WORLD_SIZE`` set, output prefixed ``[rank=N]``) — no ssh, no horovodrun.
Parity: reference ``harness/determined/exec/harness.py:43-223``, ``exec/worker_process.py:19-59``,
``exec/worker_process_wrapper.py:15-54``, ``layers/_worker_process.py:75-339``.
"""
from __future__ import annotations

import contextlib
import faulthandler
import json
import logging
import os
import pathlib
import socket
import subprocess
import sys
import threading
from typing import Any, Dict, Iterator, List, Optional

ENVIRONMENT_VARIABLE_KEYS = {
    "DET_MASTER_ADDR", "DET_MASTER_PORT", "DET_CONTAINER_ID", "DET_SLOT_IDS", "DET_USE_GPU", "DET_EXPERIMENT_ID",
    "DET_TRIAL_ID", "DET_TRIAL_SEED", "DET_EXPERIMENT_CONFIG", "DET_HPARAMS", "DET_INITIAL_WORKLOAD",
    "DET_LATEST_CHECKPOINT", "DET_WORKLOAD_MANAGER_TYPE", "DET_RENDEZVOUS_PORTS", "DET_TRIAL_RUNNER_NETWORK_INTERFACE",
    "DET_CLUSTER_ID",
}


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def build_env():
    from determined_b200 import workload as wl
    from determined_b200._env import EnvContext

    missing = sorted(k for k in ENVIRONMENT_VARIABLE_KEYS if k not in os.environ)
    if missing:
        raise RuntimeError(f"harness environment is missing: {missing}")
    e = os.environ
    slot_ids = json.loads(e["DET_SLOT_IDS"])
    latest = json.loads(e["DET_LATEST_CHECKPOINT"])
    use_gpu = e["DET_USE_GPU"].lower() == "true"
    return EnvContext(
        master_addr=e["DET_MASTER_ADDR"], master_port=int(e["DET_MASTER_PORT"]), container_id=e["DET_CONTAINER_ID"],
        experiment_config=json.loads(e["DET_EXPERIMENT_CONFIG"]), hparams=json.loads(e["DET_HPARAMS"]),
        initial_workload=wl.Workload.from_json(json.loads(e["DET_INITIAL_WORKLOAD"])), latest_checkpoint=latest or None,
        use_gpu=use_gpu, container_gpus=[f"slot-{i}" for i in slot_ids] if use_gpu else [], slot_ids=slot_ids,
        debug=e.get("DET_DEBUG", "false").lower() == "true", workload_manager_type=e["DET_WORKLOAD_MANAGER_TYPE"],
        det_rendezvous_ports=e["DET_RENDEZVOUS_PORTS"],
        det_trial_runner_network_interface=e["DET_TRIAL_RUNNER_NETWORK_INTERFACE"], det_trial_id=e["DET_TRIAL_ID"],
        det_experiment_id=e["DET_EXPERIMENT_ID"], det_cluster_id=e["DET_CLUSTER_ID"], trial_seed=int(e["DET_TRIAL_SEED"]),
        use_tls=e.get("DET_USE_TLS", "false").lower() == "true", master_cert_file=e.get("DET_MASTER_CERT_FILE") or None,
        master_cert_name=e.get("DET_MASTER_CERT_NAME") or None,
        det_trial_unique_port_offset=int(e.get("DET_TRIAL_UNIQUE_PORT_OFFSET", "0")))


def _pump(stream, rank: int, out) -> None:
    for line in iter(stream.readline, b""):
        out.write(f"[rank={rank}] ".encode() + line)
        out.flush()


def rendezvous_env(n_local: int, environ=os.environ) -> dict:
    """Process-group coordinates of this container's workers. A trial inside one container meets on loopback at a free
    port; a trial that spans containers (several agents / pods — the reference's multi-machine Horovod job,
    ``layers/_worker_process.py:113-199``) meets at the chief container's address, and a worker's global rank is its local
    rank plus the slots of the containers ranked before its own (``DET_RANK_OFFSET``, set by the master)."""
    if int(environ.get("DET_NUM_CONTAINERS", "1")) > 1:
        return {"addr": environ.get("DET_CHIEF_ADDR", "127.0.0.1"), "port": int(environ["DET_CHIEF_PORT"]),
                "offset": int(environ.get("DET_RANK_OFFSET", "0")), "world": int(environ.get("DET_WORLD_SLOTS", str(n_local)))}
    return {"addr": "127.0.0.1", "port": _free_port(), "offset": 0, "world": n_local}


def launch_workers(n: int) -> int:
    """One worker process per slot on this machine; any worker failing takes the gang down."""
    rdv = rendezvous_env(n)
    procs: List[subprocess.Popen] = []
    threads = []
    for local in range(n):
        rank = rdv["offset"] + local
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(local), WORLD_SIZE=str(rdv["world"]), LOCAL_WORLD_SIZE=str(n),
                   MASTER_ADDR=rdv["addr"], MASTER_PORT=str(rdv["port"]), DET_WORKER_PROCESS="1")
        p = subprocess.Popen([sys.executable, "-m", "determined_b200.exec.harness"], env=env, stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT)
        procs.append(p)
        t = threading.Thread(target=_pump, args=(p.stdout, rank, sys.stdout.buffer), daemon=True)
        t.start()
        threads.append(t)
    rc = 0
    import time

    while procs:
        alive = []
        for p in procs:
            code = p.poll()
            if code is None:
                alive.append(p)
            elif code != 0 and rc == 0:
                rc = code
        if rc != 0:
            for p in alive:
                p.kill()
            break
        procs = alive
        time.sleep(0.2)
    for t in threads:
        t.join(timeout=2)
    return rc


def _report_startup(marks: Dict[str, float]) -> None:
    """One log line per trial run: where the time between the agent accepting the container and the end of the first
    workload went (trial start latency). ``assigned_to_entry`` is process start + interpreter + (cold start)
    nothing else, or (warm worker) just the hand-over."""
    try:
        t_assign = float(os.environ.get("DET_CONTAINER_START_TS", "0") or 0) or marks["entered"]
        out = {"assigned_to_entry": marks["entered"] - t_assign, "imports": marks["imports"] - marks["entered"],
               "env_comm_storage": marks.get("storage_ready", marks["imports"]) - marks["imports"],
               "build_controller": marks.get("controller_built", 0) - marks.get("storage_ready", 0),
               "first_workload": marks["first_workload_done"] - marks.get("controller_built", marks["first_workload_done"]),
               "assigned_to_first_workload_done": marks["first_workload_done"] - t_assign}
        from determined_b200.utils import fast_start

        if fast_start.FIRST_BATCH_TS is not None:
            out["controller_built_to_first_batch"] = fast_start.FIRST_BATCH_TS - marks.get("controller_built", fast_start.FIRST_BATCH_TS)
            out["assigned_to_first_batch"] = fast_start.FIRST_BATCH_TS - t_assign
        out = {k: round(v, 3) for k, v in out.items()}
        warm = os.environ.get("DET_WARM_WORKER")
        out["warm_worker"] = json.loads(warm) if warm else None
        print("startup_timing: " + json.dumps(out), flush=True)
    except Exception:
        pass


def run_worker() -> int:
    import time

    marks = {"entered": time.time()}
    import torch

    from determined_b200 import load, storage
    from determined_b200 import workload as wl
    from determined_b200._env import DistributedContext
    from determined_b200._execution import InvalidHP
    from determined_b200.layers import BroadcastWorkloads, HarnessProfiler, MasterWorkloadSource, build_workload_manager
    from determined_b200.parallel.comm import Communicator, make_communicator

    from determined_b200.utils.fast_start import cap_cpu_threads, skip_dynamo_shims

    skip_dynamo_shims()
    cap_cpu_threads()
    marks["imports"] = time.time()
    env = build_env()
    if env.debug:
        faulthandler.dump_traceback_later(30, repeat=True)
    logging.basicConfig(level=logging.DEBUG if env.debug else logging.INFO, format="%(levelname)s: %(message)s")
    dist = DistributedContext.from_environ()
    is_chief = dist.get_rank() == 0
    logging.info(f"worker rank {dist.get_rank()} of {dist.get_size()} (local rank {dist.get_local_rank()}, "
                 f"{dist.get_num_agents()} container(s)) on {os.environ.get('DET_AGENT_ID', 'this host')}")
    n_slots = max(len(env.slot_ids), 1)
    env.experiment_config.setdefault("resources", {})["slots_per_trial"] = max(dist.get_size(), 1) if dist.get_size() > 1 else env.experiment_config.get("resources", {}).get("slots_per_trial", 1)
    env._per_slot_batch_size, env._global_batch_size = env._calculate_batch_sizes()
    device = torch.device("cuda", dist.get_local_rank()) if (env.use_gpu and torch.cuda.is_available()) else torch.device("cpu")
    if device.type == "cuda":
        torch.cuda.set_device(device)
    comm = make_communicator(device) if dist.get_size() > 1 else Communicator()

    storage_cfg = env.experiment_config.get("checkpoint_storage") or {"type": "shared_fs", "host_path": "/tmp/determined-checkpoints"}
    if is_chief:
        storage.validate_config(storage_cfg)
    storage_mgr = storage.build(storage_cfg)

    source = MasterWorkloadSource(env) if is_chief else None
    profiler = HarnessProfiler(use_gpu=env.use_gpu) if (env.debug and is_chief) else None

    def profiled(stream: wl.Stream) -> wl.Stream:
        for w, a, r in stream:
            if profiler:
                profiler.start()
            yield w, a, r
            if profiler:
                profiler.stop()
                profiler.serialize(f"/tmp/step-{w.step_id}-{w.kind.name}")
            if "first_workload_done" not in marks:
                marks["first_workload_done"] = time.time()
                _report_startup(marks)

    chief_stream = iter(build_workload_manager(env, profiled(iter(source)), is_chief=True)) if is_chief else None
    if not is_chief:
        # non-chief ranks still need checkpoint args shaped like the chief's; the broadcast layer provides them
        pass
    workloads = iter(BroadcastWorkloads(comm, chief_stream))

    with contextlib.ExitStack() as stack:
        load_path: Optional[pathlib.Path] = None
        if env.latest_checkpoint:
            meta = storage.StorageMetadata.from_json(env.latest_checkpoint)
            load_path = pathlib.Path(stack.enter_context(storage_mgr.restore_path(meta)))
            logging.info(f"Restoring trial from checkpoint {meta.storage_id}")
        try:
            marks["storage_ready"] = time.time()
            if os.environ.get("DET_STARTUP_PROFILE") == "1":       # where does controller construction spend its time?
                import cProfile
                import io
                import pstats

                prof = cProfile.Profile()
                controller = prof.runcall(load.prepare_controller_with_comm, env, workloads, load_path, dist, comm)
                buf = io.StringIO()
                pstats.Stats(prof, stream=buf).sort_stats("cumulative").print_stats(45)
                print(buf.getvalue(), flush=True)
            else:
                controller = load.prepare_controller_with_comm(env, workloads, load_path, dist, comm)
            marks["controller_built"] = time.time()
            controller.run()
        except InvalidHP as e:
            logging.info(f"InvalidHP raised by the trial: {e}")
            if is_chief and source is not None:
                source.report_invalid_hp()
    if dist.get_size() > 1:
        try:
            comm.barrier()
        except Exception:
            pass
    return 0


def main() -> int:
    if os.environ.get("DET_WORKER_PROCESS") == "1" or int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return run_worker()
    slot_ids = json.loads(os.environ.get("DET_SLOT_IDS", "[]"))
    cfg = json.loads(os.environ.get("DET_EXPERIMENT_CONFIG", "{}"))
    native_parallel = cfg.get("resources", {}).get("native_parallel", False)
    n = len(slot_ids)
    if int(os.environ.get("DET_NUM_CONTAINERS", "1")) > 1:
        return launch_workers(max(n, 1))      # one of several containers of a trial: join the trial-wide process group
    if n > 1:
        if native_parallel:
            # one process that sees every slot of the container; wrap_model() puts nn.DataParallel around the model
            # (reference _pytorch_context.py:135-144). Multi-container trials cannot use it (checked by the master).
            print(f"native_parallel: one worker process over {n} slots", flush=True)
            os.environ["DET_NATIVE_PARALLEL_GPUS"] = str(n)
            return run_worker()
        return launch_workers(n)
    return run_worker()


if __name__ == "__main__":
    sys.exit(main())
