"""Debug profiler: a sampling thread per workload recording CPU %, RSS, disk/net IO rates and per-GPU load/memory,
dumped to ``/tmp/step-{id}-{kind}.json`` when ``debug: true``.
Parity: reference ``harness/determined/layers/_harness_profiler.py:55-182`` (the matplotlib PNG is replaced by the
JSON series; the same fields are sampled at the same 0.1 s period)."""
from __future__ import annotations

import json
import subprocess
import threading
import time
from typing import Any, Dict, List, Optional

import psutil


def _gpu_sample() -> List[Dict[str, float]]:
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=index,utilization.gpu,memory.used,memory.total",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=2).stdout
        rows = []
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            rows.append({"index": float(f[0]), "load": float(f[1]), "mem_used": float(f[2]), "mem_total": float(f[3])})
        return rows
    except Exception:
        return []


class HarnessProfiler:
    def __init__(self, use_gpu: bool = False, period_s: float = 0.1, gpu_period_s: float = 1.0) -> None:
        self.use_gpu, self.period_s, self.gpu_period_s = use_gpu, period_s, gpu_period_s
        self.samples: List[Dict[str, Any]] = []
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._proc = psutil.Process()

    def _run(self) -> None:
        last_gpu = 0.0
        gpus: List[Dict[str, float]] = []
        disk0, net0 = psutil.disk_io_counters(), psutil.net_io_counters()
        t0 = time.time()
        while not self._stop.is_set():
            now = time.time()
            if self.use_gpu and now - last_gpu >= self.gpu_period_s:
                gpus, last_gpu = _gpu_sample(), now
            disk, net = psutil.disk_io_counters(), psutil.net_io_counters()
            dt = max(now - t0, 1e-6)
            self.samples.append({
                "t": now, "cpu_percent": psutil.cpu_percent(interval=None), "rss": self._proc.memory_info().rss,
                "disk_read_bps": ((disk.read_bytes - disk0.read_bytes) / dt) if disk and disk0 else 0.0,
                "disk_write_bps": ((disk.write_bytes - disk0.write_bytes) / dt) if disk and disk0 else 0.0,
                "net_sent_bps": (net.bytes_sent - net0.bytes_sent) / dt, "net_recv_bps": (net.bytes_recv - net0.bytes_recv) / dt,
                "gpus": gpus,
            })
            disk0, net0, t0 = disk, net, now
            self._stop.wait(self.period_s)

    def start(self) -> None:
        self.samples, self._stop = [], threading.Event()
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def stop(self) -> None:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=3)

    def serialize(self, path_prefix: str) -> str:
        path = path_prefix + ".json"
        with open(path, "w") as f:
            json.dump(self.samples, f)
        return path
