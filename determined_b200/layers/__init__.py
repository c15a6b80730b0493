"""Harness layers: master channel → workload manager → (rank broadcast) → trial controller.
Each layer is a workload ``Source`` wrapping the one above (reference: harness/determined/layers)."""
from ._master_client import MasterClient, MasterWorkloadSource
from ._workload_manager import WorkloadManager, build_workload_manager
from ._broadcast import BroadcastWorkloads
from ._profiler import HarnessProfiler

__all__ = ["MasterClient", "MasterWorkloadSource", "WorkloadManager", "build_workload_manager", "BroadcastWorkloads",
           "HarnessProfiler"]
