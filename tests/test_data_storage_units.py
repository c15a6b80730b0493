"""Unit tests of the small contracts the trial runtime is built on — batch-sampler adapters (repeat / shard / skip),
the prefetching loader, storage managers, reducers, metric utilities. Test ideas follow the reference's
``harness/tests/experiment/pytorch/test_pytorch_data.py``, ``common/tests/storage/test_shared_fs.py`` and
``harness/tests/test_util.py``."""
import os
import pathlib

import numpy as np
import pytest
import torch
from torch.utils.data import BatchSampler, SequentialSampler

from determined_b200 import storage
from determined_b200.pytorch import _data, _reducer
from determined_b200.utils import util


def batches(n, bs, drop_last=False):
    return BatchSampler(SequentialSampler(range(n)), bs, drop_last)


def test_skip_then_repeat_resumes_mid_epoch():
    s = _data.adapt_batch_sampler(batches(10, 3), repeat=True, skip=2)
    it = iter(s)
    got = [next(it) for _ in range(5)]
    # epoch = [0,1,2] [3,4,5] [6,7,8] [9]; two batches skipped, then wraps around
    assert got == [[6, 7, 8], [9], [0, 1, 2], [3, 4, 5], [6, 7, 8]]


def test_distributed_sampler_takes_every_nth_batch_and_covers_everything():
    world = 3
    per_rank = [list(_data.adapt_batch_sampler(batches(24, 2), num_replicas=world, rank=r)) for r in range(world)]
    assert per_rank[0][0] == [0, 1] and per_rank[1][0] == [2, 3] and per_rank[2][0] == [4, 5]
    assert per_rank[0][1] == [6, 7]
    seen = sorted(i for r in per_rank for b in r for i in b)
    assert seen == list(range(24))
    assert len({len(r) for r in per_rank}) == 1, "every rank gets the same number of batches"


def test_distributed_then_skip_is_per_rank():
    a = list(_data.adapt_batch_sampler(batches(16, 2), skip=1, num_replicas=2, rank=1))
    assert a[0] == [6, 7]      # rank 1 sees batches 1, 3, 5, ...; one skipped → batch 3


def test_dataloader_rejects_iterable_and_conflicting_args():
    class It(torch.utils.data.IterableDataset):
        def __iter__(self):
            return iter(range(3))
    with pytest.raises(TypeError):
        _data.DataLoader(It())
    ds = torch.utils.data.TensorDataset(torch.arange(8.0))
    with pytest.raises(ValueError):
        _data.DataLoader(ds, batch_sampler=batches(8, 2), shuffle=True)
    with pytest.raises(ValueError):
        _data.DataLoader(ds, batch_size=None)
    assert len(_data.DataLoader(ds, batch_size=3)) == 3


def test_prefetcher_preserves_order_and_forwards_errors():
    ds = torch.utils.data.TensorDataset(torch.arange(20.0))
    plain = [b[0].tolist() for b in _data.DataLoader(ds, batch_size=4).get_data_loader()]
    pre = [b[0].tolist() for b in _data.DataLoader(ds, batch_size=4, prefetch_batches=2).get_data_loader()]
    assert pre == plain and len(pre) == 5

    class Bad(torch.utils.data.Dataset):
        def __len__(self):
            return 4

        def __getitem__(self, i):
            if i == 2:
                raise RuntimeError("broken sample")
            return torch.tensor(float(i))
    it = iter(_data.DataLoader(Bad(), batch_size=1, prefetch_batches=2).get_data_loader())
    assert next(it).item() == 0.0 and next(it).item() == 1.0
    with pytest.raises(RuntimeError, match="broken sample"):
        next(it)


def test_batched_dataset_skips_collation():
    class Batched(torch.utils.data.Dataset):
        det_batched = True
        calls = 0

        def __len__(self):
            return 10

        def __getitem__(self, i):
            raise AssertionError("per-sample path must not be used")

        def __getitems__(self, idxs):
            Batched.calls += 1
            return torch.tensor(idxs) * 2
    out = list(_data.DataLoader(Batched(), batch_size=5).get_data_loader())
    assert Batched.calls == 2 and out[1].tolist() == [10, 12, 14, 16, 18]


def test_data_length_and_to_device():
    assert _data.data_length({"a": torch.zeros(7, 2)}) == 7
    assert _data.data_length([np.zeros((3, 1)), torch.zeros(3)]) == 3
    with pytest.raises(ValueError):
        _data.data_length({})
    with pytest.raises(TypeError):
        _data.data_length("nope")
    moved = _data.to_device({"x": np.ones((2, 2), dtype=np.float32), "y": [torch.zeros(1)], "z": "keep"}, torch.device("cpu"), set())
    assert isinstance(moved["x"], torch.Tensor) and moved["z"] == "keep"


# ---- storage ----------------------------------------------------------------------------------------------------------
def test_shared_fs_store_restore_delete(tmp_path: pathlib.Path):
    mgr = storage.build({"type": "shared_fs", "host_path": str(tmp_path), "save_trial_best": 3})
    with mgr.store_path() as (sid, path):
        os.makedirs(os.path.join(path, "sub"))
        with open(os.path.join(path, "sub", "w.bin"), "wb") as f:
            f.write(b"x" * 10)
        with open(os.path.join(path, "a.txt"), "w") as f:
            f.write("hi")
    meta = mgr.store(os.path.join(str(tmp_path), sid), "torch-2", "cloudpickle", sid)
    assert meta.resources == {"a.txt": 2, "sub/": 0, "sub/w.bin": 10}
    with mgr.restore_path(meta) as p:
        assert open(os.path.join(p, "a.txt")).read() == "hi"
    mgr.delete(meta)
    with pytest.raises(FileNotFoundError):
        with mgr.restore_path(meta):
            pass
    with pytest.raises(FileNotFoundError):
        mgr.delete(meta)


def test_store_path_rejects_empty_checkpoint(tmp_path):
    mgr = storage.build({"type": "shared_fs", "host_path": str(tmp_path)})
    with pytest.raises(RuntimeError, match="empty"):
        with mgr.store_path():
            pass


def test_shared_fs_storage_path_and_container_translation(tmp_path):
    mgr = storage.build({"type": "shared_fs", "host_path": "/host/ckpt", "storage_path": "exp1"}, container_path=str(tmp_path))
    assert mgr._base_path == os.path.join(str(tmp_path), "exp1")
    with pytest.raises(ValueError):
        storage.build({"type": "shared_fs", "host_path": "/host/ckpt", "storage_path": "../escape"})
    legacy = storage.build({"type": "shared_fs", "host_path": str(tmp_path), "checkpoint_path": "old"})
    assert legacy._base_path.endswith("old")
    with pytest.raises(TypeError):
        storage.build({"type": "shared_fs", "host_path": str(tmp_path), "bogus": 1})
    with pytest.raises(TypeError):
        storage.build({"type": "floppy"})


def test_validate_config_round_trip(tmp_path):
    storage.validate_config({"type": "shared_fs", "host_path": str(tmp_path)})
    assert os.listdir(tmp_path) == [], "the probe checkpoint is removed again"


def test_cloud_backends_fail_clearly_without_credentials(monkeypatch):
    for v in ('AWS_ACCESS_KEY_ID', 'AWS_SECRET_ACCESS_KEY'):
        monkeypatch.delenv(v, raising=False)
    for cfg in ({"type": "s3", "bucket": "b"}, {"type": "gcs", "bucket": "b"}, {"type": "hdfs", "hdfs_url": "http://x", "hdfs_path": "/p"}):
        try:
            mgr = storage.build(cfg)
        except (ImportError, RuntimeError, TypeError) as e:
            assert any(w in str(e).lower() for w in ("credentials", "google", "hdfs", "not available", "install", "unknown"))
            continue
        assert isinstance(mgr, storage.StorageManager)


# ---- reducers / utilities ---------------------------------------------------------------------------------------------
def test_reduce_metrics_matches_reference_semantics():
    m = np.array([1.0, 2.0, 6.0])
    assert _reducer._reduce_metrics(_reducer.Reducer.AVG, m, [1, 1, 2]) == pytest.approx((1 + 2 + 12) / 4)
    assert _reducer._reduce_metrics(_reducer.Reducer.SUM, m) == 9.0
    assert _reducer._reduce_metrics(_reducer.Reducer.MAX, m) == 6.0
    assert _reducer._reduce_metrics(_reducer.Reducer.MIN, m) == 1.0


def test_custom_metric_reducer_protocol():
    red = _reducer.make_metric_reducer(lambda flat: sum(flat))     # a function over every batch value of every rank
    red.reset()
    red.accumulate(1)
    red.accumulate(2)
    assert red.cross_slot_reduce([red.per_slot_reduce(), [10]]) == 13

    class Mean(_reducer.MetricReducer):
        def reset(self):
            self.s, self.n = 0.0, 0

        def accumulate(self, metric):
            self.s += float(metric)
            self.n += 1

        def per_slot_reduce(self):
            return self.s, self.n

        def cross_slot_reduce(self, per_slot_metrics):
            return sum(s for s, _ in per_slot_metrics) / sum(n for _, n in per_slot_metrics)
    m = Mean()
    m.reset()
    m.accumulate(2.0)
    m.accumulate(4.0)
    assert m.cross_slot_reduce([m.per_slot_reduce(), (12.0, 2)]) == 4.5


def test_make_metrics_and_json_encode():
    out = util.make_metrics(8, [{"loss": np.float32(1.0), "acc": torch.tensor(0.5)}, {"loss": 3.0, "acc": 1.0}])
    assert out["num_inputs"] == 8 and out["avg_metrics"] == {"loss": 2.0, "acc": 0.75}
    assert len(out["batch_metrics"]) == 2
    enc = util.json_encode({"nan": float("nan"), "np": np.int64(3), "arr": np.array([1.5])})
    assert '"nan": null' in enc and '"np": 3' in enc
    assert util.sizeof_fmt(96 * 1024 * 1024).startswith("96")
    assert list(util.chunks([1, 2, 3, 4, 5], 2)) == [[1, 2], [3, 4], [5]]
