"""Defaults and fixed numbers shared by the harness and the local master.

Parity: reference ``harness/determined/constants.py:21-64`` (local-mode experiment-config defaults, port
bases) and ``master/pkg/model/defaults.go:33-128`` (experiment-config defaults).
"""
import copy
from typing import Any, Dict

MAX_SLOTS_PER_AGENT = 16

# Port layout kept for config compatibility; on one box the launcher uses a Unix rendezvous + 127.0.0.1.
LOCAL_RENDEZVOUS_PORT = 1734
LOCAL_RENDEZVOUS_PORT_OFFSET = 16
DTRAIN_SSH_PORT = 12350
HOROVOD_GLOO_RENDEZVOUS_PORT = 12355  # unused here (no Horovod); reserved
INTER_TRAIN_PROCESS_COMM_PORT_1 = 12360
INTER_TRAIN_PROCESS_COMM_PORT_2 = INTER_TRAIN_PROCESS_COMM_PORT_1 + MAX_SLOTS_PER_AGENT
TORCH_DIST_PORT_BASE = 29400

MAX_CONTEXT_SIZE = 96 * 1024 * 1024  # model-definition upload cap (common/determined_common/constants.py)

DEFAULT_SEARCHER_CFG = {"name": "single", "max_length": {"batches": 100}}
DEFAULT_RESOURCES_CFG = {"slots_per_trial": 1, "native_parallel": False, "weight": 1, "priority": 42}
DEFAULT_SCHEDULING_UNIT = 100
DEFAULT_OPTIMIZATIONS = {
    "aggregation_frequency": 1,
    "average_aggregated_gradients": True,
    "average_training_metrics": False,
    "gradient_compression": False,
    "mixed_precision": "O0",
    "tensor_fusion_threshold": 64,
    "tensor_fusion_cycle_time": 5,
    "auto_tune_tensor_fusion": False,
}
DEFAULT_CHECKPOINT_STORAGE = {"type": "shared_fs", "host_path": "/tmp/determined-checkpoints", "save_experiment_best": 0,
                              "save_trial_best": 1, "save_trial_latest": 1}

# What a config gets in local (masterless) mode before the user's keys are merged over it.
DEFAULT_EXP_CFG: Dict[str, Any] = {
    "searcher": DEFAULT_SEARCHER_CFG,
    "scheduling_unit": DEFAULT_SCHEDULING_UNIT,
    "resources": DEFAULT_RESOURCES_CFG,
    "optimizations": DEFAULT_OPTIMIZATIONS,
    "checkpoint_storage": DEFAULT_CHECKPOINT_STORAGE,
    "max_restarts": 5,
    "min_validation_period": {"batches": 0},
    "min_checkpoint_period": {"batches": 0},
    "checkpoint_policy": "best",
    "perform_initial_validation": False,
    "debug": False,
}


def default_exp_cfg() -> Dict[str, Any]:
    return copy.deepcopy(DEFAULT_EXP_CFG)
