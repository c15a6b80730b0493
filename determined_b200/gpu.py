"""GPU discovery through ``nvidia-smi`` (no CUDA context is created, so the launcher can run it before forking the
per-GPU workers). Role of the reference's ``harness/determined/gpu.py:25-78``."""
from __future__ import annotations

import csv
import subprocess
from typing import List, NamedTuple, Optional, Tuple


class GPU(NamedTuple):
    id: int
    uuid: str
    load: float          # utilisation, 0..1
    memory_used_mb: float
    memory_total_mb: float
    name: str = ""


def _query(fields: str, ids: Optional[List[int]] = None) -> List[List[str]]:
    cmd = ["nvidia-smi", f"--query-gpu={fields}", "--format=csv,noheader,nounits"]
    if ids:
        cmd.append("--id=" + ",".join(str(i) for i in ids))
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=20)
    except (FileNotFoundError, subprocess.TimeoutExpired):
        return []
    if out.returncode != 0:
        return []
    return [[c.strip() for c in row] for row in csv.reader(out.stdout.strip().splitlines()) if row]


def get_gpus(ids: Optional[List[int]] = None) -> List[GPU]:
    gpus = []
    for row in _query("index,uuid,utilization.gpu,memory.used,memory.total,name", ids):
        try:
            def num(x: str) -> float:
                return float(x) if x not in ("[N/A]", "N/A", "") else 0.0
            gpus.append(GPU(int(row[0]), row[1], num(row[2]) / 100.0, num(row[3]), num(row[4]), row[5] if len(row) > 5 else ""))
        except (ValueError, IndexError):
            continue
    return gpus


def get_gpu_ids_and_uuids() -> Tuple[List[int], List[str]]:
    gpus = get_gpus()
    return [g.id for g in gpus], [g.uuid for g in gpus]


def get_gpu_uuids_and_validate(use_gpu: bool, slot_ids: List[int]) -> List[str]:
    """The launcher's sanity check: the slots the master assigned must all be visible GPUs."""
    if not use_gpu:
        return []
    ids, uuids = get_gpu_ids_and_uuids()
    if not uuids:
        raise RuntimeError("Failed to find GPUs for a GPU trial")
    if len(slot_ids) > len(uuids):
        raise RuntimeError(f"Mismatched slots: {len(slot_ids)} assigned, {len(uuids)} GPUs visible")
    return uuids


def get_driver_version() -> Optional[str]:
    rows = _query("driver_version")
    return rows[0][0] if rows else None
