"""TensorBoard metric writer + sync manager (reference: harness/determined/tensorboard/base.py:6-55, build.py:10-96,
metric_writers/pytorch.py:9-42) and the debug profiler (reference: layers/_harness_profiler.py:55-182)."""
import json
import pathlib
import time

from determined_b200.layers import HarnessProfiler
from determined_b200.tensorboard.build import BatchMetricWriter, TensorboardManager, get_sync_path


def test_sync_path_layout():
    assert get_sync_path("cid", "3", "7") == pathlib.Path("cid/tensorboard/experiment/3/trial/7")


def test_metric_writer_and_incremental_sync(tmp_path):
    base, store = tmp_path / "local", tmp_path / "store"
    mgr = TensorboardManager(base, get_sync_path("cid", "3", "7"), store)
    w = BatchMetricWriter(mgr.base_path)
    w.on_train_step_end(1, 4, [{"loss": 1.0, "note": "skipped"}, {"loss": 0.8}, {"loss": 0.7}, {"loss": 0.5}])
    events = mgr.list_tfevents()
    assert len(events) == 1
    mgr.sync()
    dst = store / "cid/tensorboard/experiment/3/trial/7" / events[0].name
    assert dst.exists() and dst.stat().st_size == events[0].stat().st_size
    assert mgr.to_sync() == []                                   # nothing grew: nothing to copy again
    w.on_validation_step_end(1, 4, {"validation_loss": 0.6, "accuracy": 0.9})
    assert mgr.to_sync() == events                               # the file grew, so it is mirrored again
    mgr._sizes.clear()
    mgr.sync()
    assert dst.stat().st_size == events[0].stat().st_size
    # the scalars are tagged the way the TensorBoard task expects them
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator

    acc = EventAccumulator(str(base))
    acc.Reload()
    tags = set(acc.Tags()["scalars"])
    assert {"Determined/loss", "Determined/validation_loss", "Determined/val_accuracy"} <= tags
    assert [e.step for e in acc.Scalars("Determined/loss")] == [0, 1, 2, 3]


def test_manager_without_storage_root_is_a_no_op(tmp_path):
    mgr = TensorboardManager(tmp_path / "l", get_sync_path("c", "1", "1"), None)
    (mgr.base_path / "events.out.tfevents.1").write_bytes(b"x")
    mgr.sync()
    assert len(mgr.list_tfevents()) == 1


def test_profiler_samples_and_serializes(tmp_path):
    p = HarnessProfiler(use_gpu=False, period_s=0.02)
    p.start()
    time.sleep(0.15)
    p.stop()
    assert len(p.samples) >= 3
    path = p.serialize(str(tmp_path / "step-1-RUN_STEP"))
    rows = json.loads(open(path).read())
    assert {"t", "cpu_percent", "rss", "disk_read_bps", "net_recv_bps", "gpus"} <= set(rows[0])
    assert rows[0]["rss"] > 0 and rows[-1]["t"] >= rows[0]["t"]
    p.start()                                                     # one profiler object serves every workload in turn
    p.stop()
    assert len(p.samples) <= 2


def test_event_files_sync_to_s3_and_are_fetched_by_the_tensorboard_task(tmp_path, monkeypatch):
    """Object-store path (reference tensorboard/s3.py + fetchers/s3.py): the trial's manager uploads new / grown event files,
    the TensorBoard task mirrors them into a local directory and picks up growth on the next poll."""
    from tests.fake_objstore import make_s3

    from determined_b200.exec.tensorboard import resolve_sources
    from determined_b200.tensorboard.build import object_store_for

    srv, url, objects, stats = make_s3()
    try:
        cfg = {"type": "s3", "bucket": "tb", "access_key": "AKIDEXAMPLE", "secret_key": "s3cr3t", "endpoint_url": url, "prefix": "runs"}
        client, prefix = object_store_for(cfg)
        assert prefix == "runs/" and object_store_for({"type": "shared_fs", "host_path": "/tmp"}) == (None, "")
        mgr = TensorboardManager(tmp_path / "local", get_sync_path("cid", "3", "7"), None, client, prefix)
        w = BatchMetricWriter(mgr.base_path)
        w.on_train_step_end(1, 2, [{"loss": 1.0}, {"loss": 0.5}])
        mgr.sync()
        name = mgr.list_tfevents()[0].name
        key = f"runs/cid/tensorboard/experiment/3/trial/7/{name}"
        assert key in objects and stats["bad_signature"] == 0
        first_size = len(objects[key])
        # the TensorBoard task's side
        sources = [{"experiment_id": 3, "trial_id": None, "storage": cfg, "prefix": "cid/tensorboard/experiment/3"}]
        runs, fetchers = resolve_sources(sources, tmp_path / "scratch")
        assert [n for n, _ in runs] == ["experiment-3"] and len(fetchers) == 1
        assert fetchers[0].fetch_new() == 1
        local = pathlib.Path(runs[0][1]) / "trial" / "7" / name
        assert local.stat().st_size == first_size
        assert fetchers[0].fetch_new() == 0                         # unchanged: nothing is downloaded again
        w.on_validation_step_end(1, 2, {"validation_loss": 0.4})
        mgr.sync()
        assert len(objects[key]) > first_size
        assert fetchers[0].fetch_new() == 1 and local.stat().st_size == len(objects[key])
        from tensorboard.backend.event_processing.event_accumulator import EventAccumulator

        acc = EventAccumulator(str(local.parent))
        acc.Reload()
        assert {"Determined/loss", "Determined/validation_loss"} <= set(acc.Tags()["scalars"])
        # shared_fs sources are served in place, unknown back-ends are skipped
        runs, fetchers = resolve_sources([{"experiment_id": 4, "trial_id": 9, "prefix": "cid/tensorboard/experiment/4/trial/9",
                                           "storage": {"type": "shared_fs", "host_path": str(tmp_path), "storage_path": "sp"}},
                                          {"experiment_id": 5, "prefix": "x", "storage": {"type": "hdfs", "hdfs_url": "http://h", "hdfs_path": "/"}}],
                                         tmp_path / "scratch2")
        assert runs == [("experiment-4-trial-9", str(tmp_path / "sp" / "cid/tensorboard/experiment/4/trial/9"))] and fetchers == []
    finally:
        srv.shutdown()
