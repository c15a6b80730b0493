"""ctypes binding to ``libdet_kernels.so`` — the C ABI declared in ``include/det_kernels.h``.

The library is built in-tree by :mod:`determined_b200._build`. On a machine with a GPU the ops in
this package REQUIRE the library: a missing ``.so`` raises instead of silently falling back to
eager PyTorch, so "tests pass" always means "the sm_100a kernels ran".
"""
from __future__ import annotations

import ctypes
import os
from pathlib import Path
from typing import Optional

_LIB: Optional[ctypes.CDLL] = None
# DET_KERNELS_LIB points at an alternative build of the same ABI (A/B timing of two kernel versions on one box).
_LIB_PATH = Path(os.environ.get("DET_KERNELS_LIB") or Path(__file__).resolve().parent.parent / "_lib" / "libdet_kernels.so")

c_void_p, c_int, c_ll, c_float, c_uint = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_uint

_SIGNATURES = {
    "det_gemm_bf16": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_ll, c_ll, c_ll, c_int, c_int,
                      c_float, c_void_p, c_void_p, c_void_p, c_ll, c_float, c_uint, c_void_p, c_int, c_int, c_void_p],
    "det_attention_fwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_uint, c_void_p, c_void_p],
    "det_attention_bwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_uint, c_void_p, c_void_p],
    "det_attention_flash_fwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_uint, c_void_p, c_void_p],
    "det_attention_flash_bwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                c_float, c_uint, c_void_p, c_void_p],
    "det_layernorm_fwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_uint, c_void_p, c_void_p],
    "det_layernorm_bwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                          c_int, c_int, c_float, c_uint, c_float, c_uint, c_void_p, c_void_p],
    "det_colsum_bf16": [c_void_p, c_void_p, c_int, c_int, c_ll, c_void_p],
    "det_colsum_select_bf16": [c_void_p, c_void_p, c_ll, c_void_p, c_int, c_int, c_ll, c_void_p],
    "det_embed_gather": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "det_embed_scatter": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "det_softmax_ce": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p],
    "det_gemm_small": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_ll, c_ll, c_ll, c_ll, c_ll, c_ll,
                       c_int, c_int, c_int, c_void_p, c_int, c_float, c_int, c_void_p],
    "det_cast_f32_bf16": [c_void_p, c_void_p, c_ll, c_void_p],
    "det_conv2d_implicit": [c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                            c_int, c_int, c_void_p, c_int, c_int, c_int, c_void_p],
    "det_conv2d_fprop_stats": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                               c_void_p, c_void_p, c_void_p],
    "det_gemm_bf16_stats": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_ll, c_ll, c_ll, c_int, c_void_p, c_void_p, c_void_p],
    "det_bn_finalize": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_float, c_float, c_void_p],
    "det_dwconv2d_fwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_dwconv2d_dgrad": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_dwconv2d_wgrad": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_avgpool2d_fwd": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_avgpool2d_bwd": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_highway_gate_fwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "det_highway_gate_bwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "det_im2col": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_col2im": [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_bn_stats": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_float, c_float, c_void_p],
    "det_bn_apply": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_int, c_float,
                     c_void_p],
    "det_bn_bwd": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                   c_void_p, c_ll, c_int, c_int, c_float, c_void_p],
    "det_maxpool_fwd": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_maxpool_bwd": [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_avgpool_fwd": [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "det_avgpool_bwd": [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "det_act_fwd": [c_void_p, c_void_p, c_ll, c_int, c_float, c_void_p],
    "det_act_bwd": [c_void_p, c_void_p, c_void_p, c_ll, c_int, c_float, c_void_p],
    "det_dropout": [c_void_p, c_void_p, c_ll, c_float, c_uint, c_void_p, c_void_p],
    "det_nchw_to_nhwc_bf16": [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p],
    "det_u8_to_bf16_norm": [c_void_p, c_void_p, c_ll, c_int, c_void_p, c_void_p, c_void_p],
    "det_fused_optimizer_step": [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_void_p, c_void_p,
                                 c_void_p, c_int, c_void_p],
    "det_grad_sumsq_finite": [c_void_p, c_ll, c_void_p, c_void_p, c_void_p, c_void_p],
    "det_clip_coef": [c_void_p, c_float, c_void_p, c_void_p, c_void_p],
    "det_loss_scale_update": [c_void_p, c_void_p, c_float, c_float, c_int, c_float, c_void_p],
    "det_fused_allreduce": [c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_int, c_float, c_void_p, c_void_p,
                            c_void_p, c_int, c_void_p],
    "det_fused_allreduce_cta": [c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_int, c_float, c_void_p, c_void_p,
                                c_void_p, c_int, c_int, c_int, c_void_p],
    "det_fused_allreduce_bf16wire": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_int, c_float, c_void_p,
                                     c_void_p, c_void_p, c_int, c_void_p],
    "det_reduce_scatter_stats": [c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_int, c_int, c_float, c_void_p, c_void_p,
                                 c_void_p, c_int, c_void_p],
    "det_allgather_send": [c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_int, c_int, c_int, c_int, c_void_p],
    "det_wait_chunk": [c_void_p, c_int, c_int, c_int, c_void_p],
    "det_p2p_broadcast": [c_void_p, c_void_p, c_void_p, c_ll, c_int, c_int, c_int, c_void_p],
    "det_metric_allreduce": [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p],
    "det_vmm_multicast_supported": [c_int, ctypes.POINTER(c_int)],
    "det_vmm_granularity": [c_int, c_int, c_int, ctypes.POINTER(c_ll)],
    "det_vmm_create": [c_int, c_ll, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(c_int)],
    "det_vmm_import": [c_int, ctypes.POINTER(ctypes.c_ulonglong)],
    "det_vmm_map": [ctypes.c_ulonglong, c_ll, c_ll, c_int, ctypes.POINTER(c_void_p)],
    "det_vmm_unmap": [c_void_p, c_ll],
    "det_vmm_release": [ctypes.c_ulonglong],
    "det_vmm_multicast_create": [c_int, c_ll, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(c_int)],
    "det_vmm_multicast_add_device": [ctypes.c_ulonglong, c_int],
    "det_vmm_multicast_bind": [ctypes.c_ulonglong, ctypes.c_ulonglong, c_ll],
    "det_arena_alloc": [c_ll, ctypes.POINTER(c_void_p)],
    "det_arena_free": [c_void_p],
    "det_ipc_export": [c_void_p, ctypes.c_char_p],
    "det_ipc_import": [ctypes.c_char_p, ctypes.POINTER(c_void_p)],
    "det_ipc_close": [c_void_p],
    "det_set_sm_limit": [c_int],
}


class NativeLibraryMissing(RuntimeError):
    pass


def lib_path() -> Path:
    return _LIB_PATH


def available() -> bool:
    return _LIB_PATH.exists()


def load(build_if_missing: bool = True) -> ctypes.CDLL:
    """Load (building first when absent and a compiler is present) and type the kernel library."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not _LIB_PATH.exists() and build_if_missing and os.environ.get("DET_NO_AUTOBUILD") != "1":
        from determined_b200 import _build

        _build.build_kernels()
    if not _LIB_PATH.exists():
        raise NativeLibraryMissing(
            f"{_LIB_PATH} not found — run `python -m determined_b200._build` (needs nvcc; no GPU required)"
        )
    lib = ctypes.CDLL(str(_LIB_PATH))
    lib.det_last_error.restype = ctypes.c_char_p
    lib.det_kernels_abi_version.restype = c_int
    lib.det_comm_flag_bytes.restype = c_ll
    lib.det_comm_epoch_bytes.restype = c_ll
    if hasattr(lib, "det_pipe_chunk_floats"):
        lib.det_pipe_chunk_floats.restype = c_ll
        lib.det_pipe_chunk_floats.argtypes = [c_ll, c_int, c_int]
    for name, sig in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if os.environ.get("DET_KERNELS_LIB"):
                continue      # an older build loaded for comparison may lack newer entry points
            raise
        fn.argtypes = sig
        fn.restype = c_int
    _LIB = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        raise RuntimeError("det_kernels: " + load().det_last_error().decode(errors="replace"))


# Count of kernel launches issued through this binding (bench.py reports it as `gpu_launches`).
launch_count = 0


def call(name: str, *args) -> None:
    global launch_count
    launch_count += 1
    check(getattr(load(), name)(*args))
