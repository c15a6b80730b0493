"""Communicator: everything a trial's worker processes say to one another.

The reference splits this over three transports — Horovod/NCCL for gradients and parameter broadcast
(``_pytorch_context.py:192-198``, ``_pytorch_trial.py:71-73``), ZMQ + pickle for metrics and barriers
(``_trial_controller.py:269-315``, ``_pytorch_trial.py:292-335, 580-616``) and ``hvd.broadcast_object``.
Here one object owns them all:

* **control plane** — ``torch.distributed`` (NCCL on GPUs, Gloo on CPU) for rendezvous, barriers and small
  pickled objects;
* **data plane** — on CUDA the hand-written peer-memory kernels (``csrc/comm/fused_allreduce.cu``) over a
  symmetric arena: fused gradient all-reduce (sum + average + unscale + non-finite flag + grad-norm partial),
  parameter/optimizer-state broadcast, metric reduce. On CPU (tests, gloo) the same API falls back to
  ``torch.distributed`` collectives so the host-side logic is exercised without GPUs.
"""
from __future__ import annotations

import ctypes
import os
import pickle
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch

from ..ops import _native
from ..ops import functional as F


class Communicator:
    """Single-process communicator (size 1). Also the interface."""

    rank = 0
    size = 1
    local_rank = 0

    def barrier(self) -> None:
        pass

    def allgather_object(self, obj: Any) -> List[Any]:
        return [obj]

    def broadcast_object(self, obj: Any, root: int = 0) -> Any:
        return obj

    def alloc(self, n: int, dtype: torch.dtype, kind: str, device: torch.device) -> torch.Tensor:
        return torch.zeros(n, dtype=dtype, device=device)

    def allreduce_grad_bucket(self, flat_g: torch.Tensor, scale: float, scale_dev: Optional[torch.Tensor] = None,
                              sumsq_dev: Optional[torch.Tensor] = None,
                              found_inf_dev: Optional[torch.Tensor] = None, max_blocks: int = 0,
                              cta_threads: int = 0, algo: int = 0) -> bool:
        """Sum ``flat_g`` over ranks and multiply by ``scale`` in place. Returns True when the scale (and the
        optional sumsq / found_inf side outputs) were applied by the collective itself."""
        return False

    def pipeline_plan(self, flat_g: torch.Tensor) -> Optional[List[Tuple[int, int]]]:
        return None

    def allreduce_tensors(self, tensors: Sequence[torch.Tensor], scale: float) -> None:
        if scale != 1.0:
            for t in tensors:
                t.mul_(scale)

    def allreduce_max_int(self, flag: torch.Tensor) -> None:
        """In-place MAX over ranks of a small integer tensor (the fp16 overflow flag)."""

    def broadcast_tensors(self, tensors: Sequence[torch.Tensor], root: int = 0) -> None:
        pass

    def reduce_metrics(self, values: torch.Tensor, op: int = F.RED_AVG, weights: Optional[Sequence[float]] = None) -> torch.Tensor:
        return values

    def close(self) -> None:
        pass


class _SymmetricRegion:
    """One peer-mapped allocation: local tensor + every rank's address of the same region."""

    def __init__(self, local: torch.Tensor, peer_ptrs: List[int], keep: Any) -> None:
        self.local = local
        self.peer_ptrs = peer_ptrs
        self.keep = keep
        self.base = local.data_ptr()
        self.nbytes = local.numel() * local.element_size()

    def table(self, offset_bytes: int = 0):
        arr = (ctypes.c_void_p * len(self.peer_ptrs))(*[p + offset_bytes for p in self.peer_ptrs])
        return arr


class TorchDistCommunicator(Communicator):
    """Multi-process communicator over an initialised ``torch.distributed`` default group."""

    def __init__(self, device: torch.device, use_p2p: Optional[bool] = None) -> None:
        import torch.distributed as dist

        if not dist.is_initialized():
            raise RuntimeError("torch.distributed must be initialised before building a communicator")
        self.dist = dist
        self.rank, self.size = dist.get_rank(), dist.get_world_size()
        self.local_rank = int(os.environ.get("LOCAL_RANK", self.rank))
        self.compress_gradients = False    # set from optimizations.gradient_compression by the trial context
        self.device = device
        if use_p2p is None:
            use_p2p = os.environ.get("DET_DISABLE_P2P") != "1"
        # the peer-memory arena spans ONE node (NVLink / NVSwitch); a trial over several nodes reduces with NCCL
        single_node = int(os.environ.get("LOCAL_WORLD_SIZE", str(self.size))) == self.size
        self.p2p = bool(use_p2p and device.type == "cuda" and self.size <= 8 and single_node and F.native_available())
        self._regions: List[_SymmetricRegion] = []
        self._flags: Optional[_SymmetricRegion] = None
        self._epoch: Optional[torch.Tensor] = None
        self._metric_region: Optional[_SymmetricRegion] = None
        self._wire_regions: Dict[int, _SymmetricRegion] = {}      # bf16 staging of a gradient bucket, keyed by its address
        self._vmm: Any = None                                     # (FdExchange, multicast usable) | False | None = not probed
        self._backend = "none"
        if self.p2p:
            try:
                lib = _native.load()
                self._flags = self._symmetric_alloc(int(lib.det_comm_flag_bytes()) // 4, torch.int32)
                self._epoch = torch.zeros(int(lib.det_comm_epoch_bytes()) // 4, dtype=torch.int32, device=device)
                self._metric_region = self._symmetric_alloc(4096, torch.float32)
                self.barrier()
            except Exception as e:  # pragma: no cover - depends on the box
                if os.environ.get("DET_REQUIRE_P2P") == "1":
                    raise
                print(f"[det] peer-memory arena unavailable ({e!r}); falling back to NCCL collectives", flush=True)
                self.p2p = False

    # ---- control plane ---------------------------------------------------------------------------
    def barrier(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)      # a host barrier orders device work only if each rank drained its own
            if "gloo" in str(self.dist.get_backend()) or "cpu" in str(self.dist.get_backend()):
                self.dist.all_reduce(torch.zeros(1))                # CPU tensor: rides gloo, never touches NCCL
                return
            self.dist.barrier(device_ids=[self.device.index if self.device.index is not None else 0])
        else:
            self.dist.barrier()

    def allgather_object(self, obj: Any) -> List[Any]:
        out: List[Any] = [None] * self.size
        self.dist.all_gather_object(out, obj)
        return out

    def broadcast_object(self, obj: Any, root: int = 0) -> Any:
        box = [obj]
        self.dist.broadcast_object_list(box, src=root)
        return box[0]

    # ---- symmetric memory --------------------------------------------------------------------------
    def _symmetric_alloc(self, n: int, dtype: torch.dtype) -> _SymmetricRegion:
        """Allocate the same-sized region on every rank and map every peer's copy.

        Preferred: ``torch.distributed._symmetric_memory`` (cuMem + fabric/fd handle exchange; also gives
        an NVLS multicast address when the switch supports it). Fallback: cudaMalloc + CUDA IPC handles
        exchanged through the control plane."""
        elem = torch.empty((), dtype=dtype).element_size()
        backend = os.environ.get("DET_ARENA_BACKEND", "vmm")
        if backend == "vmm":
            # native arena (csrc/comm/vmm_arena.cu + parallel/vmm.py): cuMemCreate + fd exchange + cuMulticast*. Whether it
            # (and multicast) can be used is agreed by all ranks once; afterwards a failure is fatal, not a fallback.
            from . import vmm

            if self._vmm is None:
                idx = self.device.index if self.device.index is not None else 0
                torch.zeros(1, device=self.device)          # primary context up before the driver API is used
                ok, mc = vmm.supported(self, idx)
                self._vmm = (vmm.FdExchange(self), mc) if ok else False
                if not ok and self.rank == 0:
                    print("[det] native VMM arena unavailable on this box; using torch symmetric memory", flush=True)
            if self._vmm:
                exch, mc = self._vmm
                idx = self.device.index if self.device.index is not None else 0
                reg, err = None, ""
                try:
                    reg = vmm.allocate(self, exch, idx, n * elem, with_multicast=mc and os.environ.get("DET_DISABLE_NVLS") != "1")
                except Exception as e:      # noqa: BLE001 - reported below, decided together
                    err = repr(e)
                if not self._regions:
                    # the FIRST region decides the backend for the whole run, and all ranks decide together: one failing
                    # rank (socket permissions, a driver that refuses the export) sends everybody to torch symmetric memory
                    oks = self.allgather_object(reg is not None)
                    if not all(oks):
                        if self.rank == 0 or err:
                            print(f"[det] native VMM arena failed on rank(s) {[r for r, ok in enumerate(oks) if not ok]} "
                                  f"({err or 'peer failure'}); using torch symmetric memory", flush=True)
                        exch.close()
                        self._vmm = False
                        reg = None
                elif reg is None:
                    raise RuntimeError(f"native VMM arena allocation failed after the arena was in use: {err}")
            if self._vmm and reg is not None:
                local = _tensor_from_ptr(reg.ptrs[self.rank], n, dtype, self.device)
                local.zero_()
                torch.cuda.synchronize(self.device)
                region = _SymmetricRegion(local, reg.ptrs, reg)
                region.multicast_ptr = reg.multicast_ptr
                self._backend = "cuda_vmm"
                self._regions.append(region)
                self.barrier()                               # every copy is zeroed before anyone writes through a peer mapping
                return region
            backend = "symm"
        if backend == "symm":
            try:
                import torch.distributed._symmetric_memory as symm

                t = symm.empty(n, dtype=dtype, device=self.device)
                hdl = symm.rendezvous(t, self.dist.group.WORLD)
                t.zero_()
                ptrs = [int(p) for p in hdl.buffer_ptrs]
                region = _SymmetricRegion(t, ptrs, hdl)
                region.multicast_ptr = int(getattr(hdl, "multicast_ptr", 0) or 0)
                self._backend = "torch_symmetric_memory"
                self._regions.append(region)
                return region
            except Exception as e:  # pragma: no cover
                if os.environ.get("DET_ARENA_BACKEND") == "symm":
                    raise
                print(f"[det] torch symmetric memory unavailable ({e!r}); using CUDA IPC arena", flush=True)
        lib = _native.load()
        ptr = ctypes.c_void_p()
        _native.check(lib.det_arena_alloc(n * elem, ctypes.byref(ptr)))
        handle = ctypes.create_string_buffer(64)
        _native.check(lib.det_ipc_export(ptr, handle))
        handles = self.allgather_object(bytes(handle.raw))
        ptrs: List[int] = []
        for r, h in enumerate(handles):
            if r == self.rank:
                ptrs.append(int(ptr.value))
            else:
                out = ctypes.c_void_p()
                _native.check(lib.det_ipc_import(h, ctypes.byref(out)))
                ptrs.append(int(out.value))
        local = _tensor_from_ptr(int(ptr.value), n, dtype, self.device)
        region = _SymmetricRegion(local, ptrs, None)
        region.multicast_ptr = 0
        self._backend = "cuda_ipc"
        self._regions.append(region)
        return region

    def alloc(self, n: int, dtype: torch.dtype, kind: str, device: torch.device) -> torch.Tensor:
        if self.p2p and kind in ("grad", "param", "state"):
            region = self._symmetric_alloc(n, dtype)
            return region.local
        return torch.zeros(n, dtype=dtype, device=device)

    def _region_of(self, t: torch.Tensor) -> Optional[_SymmetricRegion]:
        p = t.data_ptr()
        for r in self._regions:
            if r.base <= p < r.base + r.nbytes:
                return r
        return None

    # ---- data plane ------------------------------------------------------------------------------------
    def allreduce_grad_bucket(self, flat_g, scale, scale_dev=None, sumsq_dev=None, found_inf_dev=None, max_blocks: int = 0,
                              cta_threads: int = 0, algo: int = 0) -> bool:
        region = self._region_of(flat_g) if self.p2p else None
        if region is None:
            if self.compress_gradients:
                # `optimizations.gradient_compression` (reference: hvd.Compression.fp16, _pytorch_context.py:191-198): halve
                # the bytes on the wire of the library collective — the path a job takes once it leaves the NVSwitch domain,
                # where bandwidth is the scarce thing. bf16 keeps fp32's range, so no overflow handling is needed; gloo
                # (CPU tests) has no bf16 sum and gets fp16 like the reference.
                wire = torch.bfloat16 if flat_g.is_cuda else torch.float16
                packed = flat_g.to(wire)
                self.dist.all_reduce(packed)
                flat_g.copy_(packed)
            else:
                self.dist.all_reduce(flat_g)
            return False
        off = flat_g.data_ptr() - region.base
        n = flat_g.numel()
        assert n % 4 == 0
        if self.compress_gradients and n % 8 == 0 and self.size > 1 and algo == 0:
            # `optimizations.gradient_compression`: bf16 on the wire inside the arena too (pack -> switch / peer reduction
            # with fp32 accumulation -> unpack); the staging buffer is one more symmetric region, made on first use by
            # every rank at the same call
            wire = self._wire_regions.get(flat_g.data_ptr())
            if wire is None:
                wire = self._wire_regions[flat_g.data_ptr()] = self._symmetric_alloc(n, torch.bfloat16)
                self.barrier()
            wmc = wire.multicast_ptr if (wire.multicast_ptr and os.environ.get("DET_DISABLE_NVLS") != "1") else None
            _native.call("det_fused_allreduce_bf16wire", F._p(flat_g), wire.table(), self._flags.table(),
                         ctypes.c_void_p(wmc) if wmc else None, ctypes.c_void_p(self._epoch.data_ptr()), n, self.rank, self.size,
                         float(scale), F._p(scale_dev), F._p(sumsq_dev), F._p(found_inf_dev), int(max_blocks), F._stream())
            return True
        mc = region.multicast_ptr + off if (region.multicast_ptr and os.environ.get("DET_DISABLE_NVLS") != "1") else None
        _native.call("det_fused_allreduce_cta", region.table(off), self._flags.table(), ctypes.c_void_p(mc) if mc else None,
                     ctypes.c_void_p(self._epoch.data_ptr()), n, self.rank, self.size, float(scale),
                     F._p(scale_dev), F._p(sumsq_dev), F._p(found_inf_dev), int(max_blocks), int(cta_threads), int(algo),
                     F._stream())
        # sumsq_dev / found_inf_dev now describe the WHOLE bucket on every rank: the kernel's last block exchanges the
        # per-rank partials through the flag region (no second launch)
        return True

    # ---- pipelined all-reduce: reduce-scatter now, all-gather chunk by chunk under the optimizer -----------------------
    PIPE_CHUNKS = 4

    def pipeline_plan(self, flat_g: torch.Tensor) -> Optional[List[Tuple[int, int]]]:
        """Chunk bounds (elements) when the bucket qualifies for the pipelined form, else None. It pays where the in-switch
        reduction is the all-reduce path (more than two ranks, multicast mapping) and the bucket is large."""
        # Opt-in (DET_AR_PIPELINE=1). Measured on 8 GPUs (profiles/comm_timing_w8_r2.json, bench A/B 19.83 ms fused vs 19.86-19.91
        # pipelined): split in two, the halves lose the full-duplex overlap they have inside the fused kernel (reduce-scatter
        # alone 0.69 ms, both halves 1.41 ms against 1.02 ms fused), which eats what hiding the all-gather under the 0.61 ms
        # optimizer saves. Two ranks: the TMA bulk all-reduce is already shorter than the reduce-scatter half.
        if not self.p2p or self.size < 2 or self.compress_gradients or os.environ.get("DET_AR_PIPELINE", "0") != "1":
            return None
        region = self._region_of(flat_g)
        n = flat_g.numel()
        if region is None or n % 4 != 0 or n < int(os.environ.get("DET_AR_PIPELINE_MIN", str(1 << 22))):
            return None
        if not (region.multicast_ptr and os.environ.get("DET_DISABLE_NVLS") != "1") and os.environ.get("DET_AR_PIPELINE") != "1":
            return None
        chunks = int(os.environ.get("DET_AR_PIPELINE_CHUNKS", str(self.PIPE_CHUNKS)))
        chunk = int(_native.load().det_pipe_chunk_floats(n, self.size, chunks))
        return [(min(c * chunk, n), min((c + 1) * chunk, n)) for c in range(chunks)]

    def _pipe_args(self, flat_g: torch.Tensor):
        region = self._region_of(flat_g)
        off = flat_g.data_ptr() - region.base
        mc = region.multicast_ptr + off if (region.multicast_ptr and os.environ.get("DET_DISABLE_NVLS") != "1") else None
        return region.table(off), self._flags.table(), (ctypes.c_void_p(mc) if mc else None), ctypes.c_void_p(self._epoch.data_ptr())

    def reduce_scatter_stats(self, flat_g: torch.Tensor, chunks: int, scale: float, scale_dev=None, sumsq_dev=None,
                             found_inf_dev=None) -> None:
        bufs, flags, mc, ep = self._pipe_args(flat_g)
        _native.call("det_reduce_scatter_stats", bufs, flags, mc, ep, flat_g.numel(), self.rank, self.size, int(chunks), float(scale),
                     F._p(scale_dev), F._p(sumsq_dev), F._p(found_inf_dev), 0, F._stream())

    def allgather_send(self, flat_g: torch.Tensor, chunks: int, c: int) -> None:
        bufs, flags, mc, ep = self._pipe_args(flat_g)
        _native.call("det_allgather_send", bufs, flags, mc, ep, flat_g.numel(), self.rank, self.size, int(chunks), int(c), 0, F._stream())

    def wait_chunk(self, c: int) -> None:
        _native.call("det_wait_chunk", ctypes.c_void_p(self._flags.local.data_ptr()), int(c), self.rank, self.size, F._stream())

    def allreduce_tensors(self, tensors, scale) -> None:
        for t in tensors:
            if self.compress_gradients and t.dtype == torch.float32:
                packed = t.to(torch.bfloat16 if t.is_cuda else torch.float16)     # see allreduce_grad_bucket
                self.dist.all_reduce(packed)
                t.copy_(packed)
            else:
                self.dist.all_reduce(t)
            if scale != 1.0:
                t.mul_(scale)

    def allreduce_max_int(self, flag: torch.Tensor) -> None:
        self.dist.all_reduce(flag, op=self.dist.ReduceOp.MAX)

    def broadcast_tensors(self, tensors, root: int = 0) -> None:
        for t in tensors:
            region = self._region_of(t) if self.p2p else None
            nbytes = t.numel() * t.element_size()
            if region is None or nbytes % 16 != 0:
                self.dist.broadcast(t, src=root)
                continue
            off = t.data_ptr() - region.base
            _native.call("det_p2p_broadcast", region.table(off), self._flags.table(),
                         ctypes.c_void_p(self._epoch.data_ptr()), nbytes, self.rank, self.size, root, F._stream())

    def reduce_metrics(self, values, op=F.RED_AVG, weights=None) -> torch.Tensor:
        """Reduce a small fp32 vector across ranks; result valid on every rank."""
        if self.p2p and values.is_cuda and values.numel() <= 4096:
            reg = self._metric_region
            n = values.numel()
            reg.local[:n].copy_(values.reshape(-1).float())
            out = torch.empty(n, dtype=torch.float32, device=values.device)
            w = None
            if weights is not None:
                w = torch.tensor(list(weights), dtype=torch.float32, device=values.device)
            _native.call("det_metric_allreduce", reg.table(), self._flags.table(),
                         ctypes.c_void_p(self._epoch.data_ptr()), F._p(out), F._p(w), n, self.rank, self.size, op, F._stream())
            return out.view_as(values)
        v = values.float().clone()
        if op == F.RED_AVG:
            wsum = float(self.size)
            if weights is not None:
                v = v * float(weights[self.rank])
                wsum = float(sum(weights))
            self.dist.all_reduce(v)
            return v / wsum
        ops = {F.RED_SUM: self.dist.ReduceOp.SUM, F.RED_MAX: self.dist.ReduceOp.MAX, F.RED_MIN: self.dist.ReduceOp.MIN}
        self.dist.all_reduce(v, op=ops[op])
        return v

    def describe(self) -> Dict[str, Any]:
        return {"size": self.size, "p2p": self.p2p, "arena_backend": self._backend,
                "nvls": bool(self._regions and getattr(self._regions[0], "multicast_ptr", 0))}


def _tensor_from_ptr(ptr: int, n: int, dtype: torch.dtype, device: torch.device) -> torch.Tensor:
    """Wrap raw device memory as a tensor through ``__cuda_array_interface__``."""
    typestr = {torch.float32: "<f4", torch.int32: "<i4", torch.bfloat16: "<u2", torch.float16: "<f2"}[dtype]

    class _Holder:
        pass

    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (n,), "typestr": typestr, "data": (ptr, False), "version": 3}
    t = torch.as_tensor(h, device=device)
    if dtype == torch.bfloat16:
        t = t.view(torch.bfloat16)
    return t


def init_process_group_from_env(device: torch.device) -> None:
    import torch.distributed as dist

    if dist.is_initialized():
        return
    # GPUs: gloo carries the control plane (barriers, small objects) and NCCL exists only as the lazy fallback for CUDA
    # collectives — inside one NVSwitch box the gradient path is the peer-memory arena, so a trial that never leaves it
    # never pays NCCL's communicator set-up (several seconds of every multi-GPU trial start). DET_PG_BACKEND=nccl restores
    # an eagerly initialised NCCL group.
    backend = os.environ.get("DET_PG_BACKEND", "cpu:gloo,cuda:nccl") if device.type == "cuda" else "gloo"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    kwargs: Dict[str, Any] = {}
    if device.type == "cuda" and backend == "nccl":
        kwargs["device_id"] = device
    dist.init_process_group(backend=backend, rank=int(os.environ.get("RANK", "0")),
                            world_size=int(os.environ.get("WORLD_SIZE", "1")), **kwargs)


def make_communicator(device: torch.device) -> Communicator:
    if int(os.environ.get("WORLD_SIZE", "1")) <= 1:
        return Communicator()
    init_process_group_from_env(device)
    return TorchDistCommunicator(device)
