"""Host-side logic of the backward / all-reduce overlap (no GPU needed): chunking of the flat bucket, launch on the
last pending notification, double uses, untracked parameters, two optimizers (GAN pattern)."""
import types

import torch

from determined_b200.ops import modules as M
from determined_b200.parallel.overlap import GradOverlap


class FakeComm:
    size = 2

    def __init__(self):
        self.calls = []

    def allreduce_grad_bucket(self, view, scale, scale_dev=None, found_inf_dev=None, sumsq_dev=None, max_blocks=0, cta_threads=0):
        self.calls.append((view.data_ptr(), view.numel(), scale))
        return True


def make_opt(sizes):
    params = [torch.nn.Parameter(torch.zeros(n)) for n in sizes]
    offsets, off = [], 0
    for n in sizes:
        offsets.append(off)
        off += (n + 63) // 64 * 64
    flat = torch.zeros(off)
    grp = types.SimpleNamespace(params=params, offsets=offsets, total=off, flat_g=flat)
    opt = types.SimpleNamespace(fused=True, groups=[grp], found_inf_dev=None, _scratch=torch.zeros(4))
    return opt, params, flat


CTX = types.SimpleNamespace(needs_input_grad=(True,))


def test_chunks_launch_in_backward_order_and_finish_flushes():
    M.reset_param_uses()
    comm = FakeComm()
    ov = GradOverlap(comm, chunk_bytes=4 * 128)          # 128 elements per chunk
    opt, (a, b, c, d), flat = make_opt([64, 64, 128, 64])
    ov.add_optimizer(opt)
    assert [(ch.start, ch.n) for ch in ov.chunks] == [(0, 128), (128, 128), (256, 64)]
    try:
        for p in (a, b, c, d):
            M.note_param_use(CTX, p)
        ov.arm(0.5, None, None)
        M.notify_grads_written(d)                         # backward runs last layer first
        assert ov.launch_log == [(256, 64)]
        M.notify_grads_written(c)
        assert ov.launch_log == [(256, 64), (128, 128)]
        M.notify_grads_written(b)
        assert len(ov.launch_log) == 2, "chunk (a, b) still waits for a"
        assert ov.armed_for(opt)
        ov.finish(opt)                                    # a never reported (e.g. unused branch): flushed here
        assert ov.launch_log == [(256, 64), (128, 128), (0, 128)]
        assert all(s == 0.5 for _, _, s in comm.calls) and not ov.open
        base = flat.data_ptr()
        assert [((p - base) // 4, n) for p, n, _ in comm.calls] == ov.launch_log
    finally:
        ov.close()
        M.reset_param_uses()


def test_double_use_waits_for_both_and_untracked_defers():
    M.reset_param_uses()
    comm = FakeComm()
    ov = GradOverlap(comm, chunk_bytes=4 * 64)
    opt, (a, b), _ = make_opt([64, 64])
    ov.add_optimizer(opt)
    try:
        M.note_param_use(CTX, a)
        M.note_param_use(CTX, a)                          # discriminator applied to real and fake batches
        # b is used by plain torch autograd: no note, no notification
        ov.arm(1.0, None, None)
        M.notify_grads_written(a)
        assert ov.launch_log == []
        M.notify_grads_written(a)
        assert ov.launch_log == [(0, 64)]
        ov.finish(opt)
        assert ov.launch_log == [(0, 64), (64, 64)]
        # no_grad forwards are not counted
        M.note_param_use(types.SimpleNamespace(needs_input_grad=(False, False)), a)
        assert M.pending_uses(a) == 0
    finally:
        ov.close()
        M.reset_param_uses()


def test_two_optimizers_gan_pattern():
    M.reset_param_uses()
    comm = FakeComm()
    ov = GradOverlap(comm, chunk_bytes=4 * 64)
    g_opt, (g,), g_flat = make_opt([64])
    d_opt, (d,), d_flat = make_opt([64])
    ov.add_optimizer(g_opt)
    ov.add_optimizer(d_opt)
    try:
        # generator step: D's parameters do not require grad, so they are never noted
        d.requires_grad_(False)
        M.note_param_use(CTX, g)
        M.note_param_use(CTX, d)
        ov.arm(1.0, None, None)
        M.notify_grads_written(d, g)
        assert [c[0] for c in comm.calls] == [g_flat.data_ptr()]
        ov.finish(g_opt)
        assert len(comm.calls) == 1 and not ov.open
        # discriminator step: D used twice, G detached
        d.requires_grad_(True)
        M.note_param_use(CTX, d)
        M.note_param_use(CTX, d)
        ov.arm(1.0, None, None)
        M.notify_grads_written(d)
        M.notify_grads_written(d)
        assert [c[0] for c in comm.calls] == [g_flat.data_ptr(), d_flat.data_ptr()]
        ov.finish(d_opt)
        assert len(comm.calls) == 2
    finally:
        ov.close()
        M.reset_param_uses()
