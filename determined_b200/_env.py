"""Per-trial environment: what the master tells a trial process about itself.

Parity: reference ``harness/determined/_env_context.py:10-117`` (``EnvContext`` and its batch-size
arithmetic), ``_experiment_config.py:4-43``, ``_rendezvous_info.py``, and the distributed identity the
reference gets from Horovod (``_train_context.py:186-220``). Here the identity comes from the launcher
(one process per GPU, ``RANK`` / ``LOCAL_RANK`` / ``WORLD_SIZE``) — there is no Horovod.
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional

from . import workload as wl


class ExperimentConfig(dict):
    """Dict with accessors for the few keys the harness itself reads."""

    def debug_enabled(self) -> bool:
        return bool(self.get("debug", False))

    def scheduling_unit(self) -> int:
        return int(self.get("scheduling_unit", 100))

    def native_parallel_enabled(self) -> bool:
        return bool(self.get("resources", {}).get("native_parallel", False))

    def native_enabled(self) -> bool:
        """The experiment was submitted through the native API: the harness runs ``internal.native.command``."""
        return bool((self.get("internal") or {}).get("native"))

    def experiment_seed(self) -> int:
        return int((self.get("reproducibility") or {}).get("experiment_seed", 0) or 0)

    def horovod_optional_args(self) -> list:
        """Extra launcher arguments smuggled in ``data.__det_dtrain_args`` (reference ``_experiment_config.py:8-9``); kept for
        config compatibility — there is no horovodrun here to pass them to."""
        return list((self.get("data") or {}).get("__det_dtrain_args", []))

    def slots_per_trial(self) -> int:
        return int(self.get("resources", {}).get("slots_per_trial", 1))

    def averaging_training_metrics_enabled(self) -> bool:
        return bool(self.get("optimizations", {}).get("average_training_metrics", False))

    def mixed_precision_enabled(self) -> bool:
        return self.get("optimizations", {}).get("mixed_precision", "O0") != "O0"

    def get_data_layer_type(self) -> str:
        return self.get("data_layer", {}).get("type", "shared_fs")

    def get_records_per_epoch(self) -> Optional[int]:
        rpe = self.get("records_per_epoch")
        return int(rpe) if rpe else None

    def get_min_validation_period(self) -> Dict[str, int]:
        return self.get("min_validation_period", {"batches": 0})

    def get_searcher_metric(self) -> str:
        return self["searcher"]["metric"]

    def get_optimizations_config(self) -> Dict[str, Any]:
        return self.get("optimizations", {})


class RendezvousInfo:
    """Addresses of every container of a trial plus this container's rank among them."""

    def __init__(self, addrs: List[str], rank: int, addrs2: Optional[List[str]] = None) -> None:
        if not 0 <= rank < max(len(addrs), 1):
            raise ValueError(f"rank {rank} out of range for {len(addrs)} addresses")
        self.addrs, self.addrs2, self.rank = list(addrs), list(addrs2 or addrs), rank

    def get_rank(self) -> int:
        return self.rank

    def get_size(self) -> int:
        return len(self.addrs)

    def get_addrs(self) -> List[str]:
        return self.addrs

    def get_ip_addresses(self) -> List[str]:
        return [a.rsplit(":", 1)[0] for a in self.addrs]

    def get_ports(self) -> List[int]:
        return [int(a.rsplit(":", 1)[1]) for a in self.addrs if ":" in a]


class DistributedContext:
    """Rank / size of this worker process within its trial."""

    def __init__(self, rank: int = 0, size: int = 1, local_rank: int = 0, local_size: int = 1,
                 cross_rank: int = 0, cross_size: int = 1) -> None:
        self._rank, self._size = rank, size
        self._local_rank, self._local_size = local_rank, local_size
        self._cross_rank, self._cross_size = cross_rank, cross_size

    @classmethod
    def from_environ(cls) -> "DistributedContext":
        e = os.environ
        size = int(e.get("WORLD_SIZE", "1"))
        local_size = int(e.get("LOCAL_WORLD_SIZE", str(size)))
        rank = int(e.get("RANK", "0"))
        return cls(rank, size, int(e.get("LOCAL_RANK", str(rank % max(local_size, 1)))), local_size,
                   rank // max(local_size, 1), max(size // max(local_size, 1), 1))

    def get_rank(self) -> int:
        return self._rank

    def get_local_rank(self) -> int:
        return self._local_rank

    def get_size(self) -> int:
        return self._size

    def get_num_agents(self) -> int:
        return self._cross_size

    @property
    def use(self) -> bool:
        return self._size > 1


class EnvContext:
    def __init__(self, master_addr: str = "", master_port: int = 0, container_id: str = "",
                 experiment_config: Optional[Dict[str, Any]] = None, hparams: Optional[Dict[str, Any]] = None,
                 initial_workload: Optional[wl.Workload] = None, latest_checkpoint: Optional[Dict[str, Any]] = None,
                 use_gpu: bool = False, container_gpus: Optional[List[str]] = None, slot_ids: Optional[List[int]] = None,
                 debug: bool = False, workload_manager_type: str = "TRIAL_WORKLOAD_MANAGER",
                 det_rendezvous_ports: str = "", det_trial_runner_network_interface: str = "",
                 det_trial_id: str = "1", det_experiment_id: str = "1", det_cluster_id: str = "local",
                 trial_seed: int = 0, managed_training: bool = True, use_tls: bool = False,
                 master_cert_file: Optional[str] = None, master_cert_name: Optional[str] = None,
                 det_trial_unique_port_offset: int = 0) -> None:
        self.master_addr, self.master_port = master_addr, master_port
        self.use_tls, self.master_cert_file, self.master_cert_name = use_tls, master_cert_file, master_cert_name
        self.det_trial_unique_port_offset = det_trial_unique_port_offset
        self.container_id = container_id
        self.experiment_config = ExperimentConfig(experiment_config or {})
        self.hparams = dict(hparams or {})
        self.initial_workload = initial_workload or wl.train_workload(1)
        self.latest_checkpoint = latest_checkpoint
        self.use_gpu = use_gpu
        self.container_gpus = list(container_gpus or [])
        self.slot_ids = list(slot_ids or [])
        self.debug = debug
        self.workload_manager_type = workload_manager_type
        self.det_rendezvous_ports = det_rendezvous_ports
        self.det_trial_runner_network_interface = det_trial_runner_network_interface
        self.det_trial_id, self.det_experiment_id, self.det_cluster_id = det_trial_id, det_experiment_id, det_cluster_id
        self.trial_seed = trial_seed
        self.managed_training = managed_training
        self._per_slot_batch_size, self._global_batch_size = self._calculate_batch_sizes()

    def first_step(self) -> int:
        return self.initial_workload.step_id

    def rendezvous_ports(self) -> List[int]:
        return [int(x) for x in self.det_rendezvous_ports.split(",") if x]

    def _calculate_batch_sizes(self):
        """``global_batch_size`` must be an int hyperparameter. With multi-process data parallelism each
        slot gets ``global // slots`` and the effective global batch is rounded down to a multiple of the
        slot count; with ``native_parallel`` the one process sees the whole batch.
        (Reference: ``_env_context.py:72-109``.)"""
        if "global_batch_size" not in self.hparams:
            raise ValueError("hyperparameters must include `global_batch_size`: "
                             f"got {sorted(self.hparams)}")
        gbs = self.hparams["global_batch_size"]
        if isinstance(gbs, dict):
            raise ValueError("global_batch_size must resolve to an integer before the trial starts")
        gbs = int(gbs)
        slots = max(self.experiment_config.slots_per_trial(), 1)
        if self.experiment_config.native_parallel_enabled():
            return gbs, gbs
        if gbs < slots:
            raise ValueError(f"global_batch_size ({gbs}) must be >= slots_per_trial ({slots})")
        per_slot = gbs // slots
        return per_slot, per_slot * slots

    @property
    def per_slot_batch_size(self) -> int:
        return self._per_slot_batch_size

    @property
    def global_batch_size(self) -> int:
        return self._global_batch_size
