"""Host half of the native symmetric arena (``csrc/comm/vmm_arena.cu``): every rank allocates its region with the CUDA
virtual-memory-management API, the file descriptors that name the allocations travel between the ranks' processes over
unix sockets (``SCM_RIGHTS``), every rank maps every peer's region, and — where the devices support it — an NVLS multicast
object is created by rank 0, joined and bound by everyone, and mapped.

The result is what the collectives in ``csrc/comm/fused_allreduce.cu`` take: this rank's region, every rank's address of it,
and the multicast address. ``torch.distributed._symmetric_memory`` (the first implementation's backend) stays as the fallback
(``parallel/comm.py::_symmetric_alloc``).
"""
from __future__ import annotations

import ctypes
import os
import socket
import struct
import uuid
from typing import Any, List, Optional, Tuple

from ..ops import _native


class FdExchange:
    """All-to-all / one-to-all hand-over of file descriptors between the processes of one trial on one node."""

    def __init__(self, comm: Any) -> None:
        self.comm, self.rank, self.size = comm, comm.rank, comm.size
        token = comm.broadcast_object(uuid.uuid4().hex if comm.rank == 0 else None, root=0)
        # abstract-namespace unix sockets: nothing to clean up in the file system
        self.names = [f"\0det-vmm-{os.getuid()}-{token}-{r}" for r in range(self.size)]
        self.sock = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        self.sock.bind(self.names[self.rank])
        self.sock.listen(max(8, 4 * self.size))
        self.sock.settimeout(60.0)     # a peer that failed before sending must not leave the others in accept() forever
        comm.barrier()      # every listener is up before anyone connects

    def _send(self, dst: int, fd: int) -> None:
        with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as s:
            s.settimeout(60.0)
            s.connect(self.names[dst])
            socket.send_fds(s, [struct.pack("i", self.rank)], [fd])

    def _recv(self) -> Tuple[int, int]:
        conn, _ = self.sock.accept()
        with conn:
            conn.settimeout(60.0)
            msg, fds, _, _ = socket.recv_fds(conn, 4, 1)
        return struct.unpack("i", msg)[0], fds[0]

    def all_to_all(self, fd: int) -> List[int]:
        """Every rank contributes one descriptor; returns every rank's (own slot: the one passed in)."""
        out: List[Optional[int]] = [None] * self.size
        out[self.rank] = fd
        for dst in range(self.size):
            if dst != self.rank:
                self._send(dst, fd)
        for _ in range(self.size - 1):
            src, got = self._recv()
            out[src] = got
        self.comm.barrier()     # nobody starts the next exchange while this one's connections are still queued somewhere
        return [int(x) for x in out]  # type: ignore[arg-type]

    def from_root(self, fd: Optional[int], root: int = 0) -> int:
        if self.rank == root:
            assert fd is not None
            for dst in range(self.size):
                if dst != root:
                    self._send(dst, fd)
            got = fd
        else:
            _, got = self._recv()
        self.comm.barrier()
        return int(got)

    def close(self) -> None:
        try:
            self.sock.close()
        except OSError:
            pass


class VmmRegion:
    """One symmetric region: ``ptrs[r]`` is this process's address of rank r's copy, ``multicast_ptr`` the NVLS address (0
    when multicast is not available)."""

    def __init__(self, nbytes: int, ptrs: List[int], multicast_ptr: int, handles: List[int], mc_handle: int) -> None:
        self.nbytes, self.ptrs, self.multicast_ptr = nbytes, ptrs, multicast_ptr
        self._handles, self._mc_handle = handles, mc_handle


def supported(comm: Any, device_index: int) -> Tuple[bool, bool]:
    """(arena usable, multicast usable) — decided TOGETHER: every rank must take the same path or the exchange deadlocks."""
    lib = _native.load()
    mine = [False, False]
    if hasattr(lib, "det_vmm_granularity"):
        g = ctypes.c_longlong()
        mine[0] = lib.det_vmm_granularity(device_index, comm.size, 0, ctypes.byref(g)) == 0
        mc = ctypes.c_int()
        mine[1] = mine[0] and lib.det_vmm_multicast_supported(device_index, ctypes.byref(mc)) == 0 and mc.value == 1
    everyone = comm.allgather_object(mine)
    return all(x[0] for x in everyone), all(x[1] for x in everyone)


def allocate(comm: Any, exch: FdExchange, device_index: int, nbytes: int, with_multicast: bool) -> VmmRegion:
    lib = _native.load()
    size, rank = comm.size, comm.rank
    gran = ctypes.c_longlong()
    _native.check(lib.det_vmm_granularity(device_index, size, int(with_multicast), ctypes.byref(gran)))
    g = int(gran.value)
    total = (int(nbytes) + g - 1) // g * g
    handle, fd = ctypes.c_ulonglong(), ctypes.c_int()
    _native.check(lib.det_vmm_create(device_index, total, ctypes.byref(handle), ctypes.byref(fd)))
    fds = exch.all_to_all(int(fd.value))
    handles: List[int] = []
    ptrs: List[int] = []
    for r in range(size):
        if r == rank:
            h = int(handle.value)
        else:
            hh = ctypes.c_ulonglong()
            _native.check(lib.det_vmm_import(fds[r], ctypes.byref(hh)))
            h = int(hh.value)
        p = ctypes.c_void_p()
        _native.check(lib.det_vmm_map(ctypes.c_ulonglong(h), total, g, device_index, ctypes.byref(p)))
        handles.append(h)
        ptrs.append(int(p.value))
    for f in fds:
        try:
            os.close(f)
        except OSError:
            pass
    mc_ptr, mc_handle = 0, 0
    if with_multicast:
        mch, mfd = ctypes.c_ulonglong(), ctypes.c_int(-1)
        if rank == 0:
            _native.check(lib.det_vmm_multicast_create(size, total, ctypes.byref(mch), ctypes.byref(mfd)))
        got = exch.from_root(int(mfd.value) if rank == 0 else None, root=0)
        if rank != 0:
            _native.check(lib.det_vmm_import(got, ctypes.byref(mch)))
        try:
            os.close(got)
        except OSError:
            pass
        mc_handle = int(mch.value)
        _native.check(lib.det_vmm_multicast_add_device(ctypes.c_ulonglong(mc_handle), device_index))
        comm.barrier()      # all devices have joined: binding may start
        _native.check(lib.det_vmm_multicast_bind(ctypes.c_ulonglong(mc_handle), ctypes.c_ulonglong(handles[rank]), total))
        comm.barrier()      # all allocations are bound: the object is complete
        p = ctypes.c_void_p()
        _native.check(lib.det_vmm_map(ctypes.c_ulonglong(mc_handle), total, g, device_index, ctypes.byref(p)))
        mc_ptr = int(p.value)
    return VmmRegion(total, ptrs, mc_ptr, handles, mc_handle)
