"""The allocation-free gradient-norm entry point, called through the C ABI and checked against the regular
entry point that uses caller-owned scratch.

``det_grad_norm_noalloc`` exists for code that is handed gradient pointers but may not allocate device memory
itself — a monitoring hook inside a trainer whose memory all belongs to the framework's caching allocator, or a
C plugin called from inside the step loop where ``cudaMalloc`` would stall every stream. Its scratch is a fixed
array inside the library, shared by all calls on a device, so the test makes two calls in a row with different
sizes and contents (the second must not see the first one's partial sums or overflow flag) and checks the size
limit is enforced. The file name sorts last so a ``-x`` run reaches it only after every other test has run; it
also passes when run alone in a fresh process (it is then the first call on the device and performs the
one-time address set-up itself).
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from determined_b200.ops import _native

    lib = _native.load(build_if_missing=False)
    vp, ll = ctypes.c_void_p, ctypes.c_longlong
    lib.det_grad_norm_noalloc.argtypes = [vp, ll, vp, vp, vp, vp]
    lib.det_grad_norm_noalloc.restype = ctypes.c_int
    lib.det_grad_norm_noalloc_max_elems.restype = ll
    lib.det_grad_sumsq_finite.argtypes = [vp, ll, vp, vp, vp, vp]
    lib.det_grad_sumsq_finite.restype = ctypes.c_int
    return lib


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _both(lib, grad, scale):
    """(norm, found_inf) from the allocation-free entry point and from the existing one."""
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    out2 = torch.full((2,), -1.0, device="cuda")
    flag = torch.full((1,), 7, dtype=torch.int32, device="cuda")
    rc = lib.det_grad_norm_noalloc(_ptr(grad), grad.numel(), _ptr(scale), _ptr(out2), _ptr(flag), stream)
    assert rc == 0, lib.det_last_error().decode()
    sumsq = torch.zeros(1, device="cuda")          # caller-owned, caller-cleared scratch of the regular path
    flag_ref = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = lib.det_grad_sumsq_finite(_ptr(grad), grad.numel(), _ptr(scale), _ptr(sumsq), _ptr(flag_ref), stream)
    assert rc == 0, lib.det_last_error().decode()
    torch.cuda.synchronize()
    return (out2[0].item(), out2[1].item(), int(flag.item())), (sumsq.sqrt().item(), sumsq.item(), int(flag_ref.item()))


def test_noalloc_grad_norm_matches_regular_entry_point_on_consecutive_calls():
    lib = _lib()
    torch.manual_seed(0)
    # call 1: large, unaligned tail, scaled, contains a non-finite value -> flag must be raised by both
    g1 = torch.randn(3 * 16384 * 40 + 123, device="cuda")
    g1[1000] = float("inf")
    scale = torch.full((1,), 0.25, device="cuda")
    new, ref = _both(lib, g1, scale)
    assert new[2] == 1 and ref[2] == 1
    # call 2, immediately after: smaller, finite, no scale -> must see neither call 1's partials nor its flag
    g2 = torch.randn(16384 * 5 + 7, device="cuda") * 3.0
    new, ref = _both(lib, g2, None)
    assert new[2] == 0 and ref[2] == 0
    assert new[0] == pytest.approx(ref[0], rel=1e-4)
    assert new[1] == pytest.approx(ref[1], rel=1e-4)
    assert new[0] == pytest.approx(g2.double().norm().item(), rel=1e-4)
    # call 3: same scratch again with a scale
    g3 = torch.randn(1 << 20, device="cuda")
    new, ref = _both(lib, g3, scale)
    assert new[2] == 0
    assert new[0] == pytest.approx(ref[0], rel=1e-4)
    assert new[0] == pytest.approx(0.25 * g3.double().norm().item(), rel=1e-4)
    # call 4: a pointer that is only 4-byte aligned (a view one element into an allocation) takes the scalar-load branch
    g4 = torch.randn((1 << 20) + 1, device="cuda")[1:]
    assert g4.data_ptr() % 16 == 4
    new, ref = _both(lib, g4, None)
    assert new[2] == 0 and ref[2] == 0
    assert new[0] == pytest.approx(ref[0], rel=1e-4)
    assert new[0] == pytest.approx(g4.double().norm().item(), rel=1e-4)


def test_noalloc_grad_norm_rejects_oversize_without_launching():
    lib = _lib()
    limit = lib.det_grad_norm_noalloc_max_elems()
    assert limit == 1 << 30
    g = torch.ones(16, device="cuda")
    out2 = torch.zeros(2, device="cuda")
    rc = lib.det_grad_norm_noalloc(_ptr(g), limit + 1, None, _ptr(out2), None,
                                   ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.det_last_error().decode()
    assert "exceeds" in msg and str(limit) in msg
    torch.cuda.synchronize()
    assert out2.abs().sum().item() == 0.0
