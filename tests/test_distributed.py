"""Data-parallel trials over real process groups on CPU (gloo, world size 2): replicas stay bit-identical, the
update equals the single-process update on the global batch, metrics are averaged across slots. Mirrors the intent of
the reference's e2e distributed tests (e2e_tests/tests/experiment/test_pytorch.py) without a cluster."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from tests.fixtures import pytorch_models as M

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_ranks(world: int, case: str, tmp_path, extra_env=None, timeout: int = 300):
    out = tmp_path / f"{case}.json"
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(world), LOCAL_WORLD_SIZE=str(world),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), PYTHONPATH=ROOT, **(extra_env or {}))
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "tests", "dist_worker.py"), str(out), case],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    logs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        logs.append(o)
    assert all(p.returncode == 0 for p in procs), "\n".join(l[-3000:] for l in logs)
    return json.loads(out.read_text())


def test_dp2_onevar_matches_global_batch_sgd(tmp_path):
    rep = run_ranks(2, "onevar", tmp_path, {"CUDA_VISIBLE_DEVICES": ""})
    assert rep["world"] == 2
    a, b = rep["params_by_rank"]
    assert a == b, "replicas diverged"
    # single-process reference on the global batch of 8 (each rank saw every other batch of 4)
    data = M._OneVarData()
    w = 0.0
    for i, m in enumerate(rep["training"]):
        xs = np.array([float(data[j][0]) for j in range(8 * i, 8 * i + 8)])
        grad = np.mean(2 * (w * xs - 2.0 * xs) * xs)
        assert m["w_before"] == pytest.approx(w, abs=1e-5)
        # with average_training_metrics the reported loss is the mean over both slots = loss on the global batch
        assert m["loss"] == pytest.approx(np.mean((w * xs - 2.0 * xs) ** 2), rel=1e-4, abs=1e-6)
        w -= 0.01 * grad
    assert a[0] == pytest.approx(w, abs=1e-5)
    assert len(rep["validation"]) == 1 and "val_loss" in rep["validation"][0]


def test_dp2_aggregation_frequency_keeps_replicas_identical(tmp_path):
    rep = run_ranks(2, "xor_agg", tmp_path, {"CUDA_VISIBLE_DEVICES": ""})
    a, b = rep["params_by_rank"]
    assert a == b
    assert len(rep["training"]) == 8 and all(np.isfinite(m["loss"]) for m in rep["training"])


def test_dp2_gradient_compression_half_precision_wire(tmp_path):
    """optimizations.gradient_compression (reference: hvd.Compression.fp16): gradients cross the wire in half precision;
    replicas stay identical and the trajectory stays within half-precision rounding of the uncompressed one."""
    (tmp_path / "a").mkdir()
    (tmp_path / "b").mkdir()
    comp = run_ranks(2, "xor_compressed", tmp_path / "a", {"CUDA_VISIBLE_DEVICES": ""})
    plain = run_ranks(2, "xor_plain", tmp_path / "b", {"CUDA_VISIBLE_DEVICES": ""})
    a, b = comp["params_by_rank"]
    assert a == b
    pa = np.array(a), np.array(plain["params_by_rank"][0])
    assert not np.array_equal(*pa), "compression left no trace: the flag did not reach the collective"
    assert np.abs(pa[0] - pa[1]).max() < 5e-3 * max(1.0, np.abs(pa[1]).max())


def test_dp2_fp16_overflow_skips_the_step_on_every_rank(tmp_path):
    """Injected inf on one rank (library all-reduce path): both ranks skip that update and back the loss scale off."""
    rep = run_ranks(2, "xor_fp16_inf", tmp_path, {"CUDA_VISIBLE_DEVICES": ""})
    a, b = rep["params_by_rank"]
    assert a == b and all(np.isfinite(a)), "replicas diverged or went non-finite"
    t = rep["training"]
    assert [m["moved"] for m in t] == [1.0, 0.0, 1.0, 1.0]
    assert [m["scale"] for m in t] == [1024.0, 512.0, 512.0, 512.0]


def test_dp2_validation_with_fewer_batches_than_ranks(tmp_path):
    """A rank with an empty validation shard joins the same collectives as the others (no deadlock) and the result is
    the reduction over the batches that exist."""
    rep = run_ranks(2, "xor_val_short", tmp_path, {"CUDA_VISIBLE_DEVICES": ""}, timeout=120)
    assert len(rep["validation"]) == 1
    v = rep["validation"][0]
    assert np.isfinite(v["loss"]) and 0.0 <= v["binary_error"] <= 1.0
