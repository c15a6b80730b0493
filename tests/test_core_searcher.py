"""C++ core: NumPy-exact RNG, config validation, every search method checked by simulation (the approach of the
reference's master/pkg/searcher/*_test.go: drive the searcher with synthetic validations and compare the multiset of
per-trial operation sequences written as "100B V 300B V")."""
import hashlib
import uuid

import numpy as np
import pytest

from determined_b200.master import _core as core

HP = {"global_batch_size": 32, "lr": {"type": "log", "minval": -4, "maxval": -1, "base": 10},
      "layers": {"type": "int", "minval": 1, "maxval": 5}, "act": {"type": "categorical", "vals": ["relu", "tanh"]},
      "drop": {"type": "double", "minval": 0.0, "maxval": 0.5}}


def test_nprand_matches_numpy_randomstate():
    for seed in (0, 1, 42, 2 ** 31 + 5):
        ref = np.random.RandomState(seed)
        rng = core.NpRand(seed)
        st = ref.get_state()
        ref2 = np.random.RandomState(seed)
        assert [rng.bits32() for _ in range(700)] == [int(x) for x in ref2.randint(0, 2 ** 32, size=700, dtype=np.uint64)]
        rng = core.NpRand(seed)
        ref = np.random.RandomState(seed)
        assert [rng.unit_interval() for _ in range(50)] == [ref.random_sample() for _ in range(50)]
        assert [rng.uniform(-2.0, 3.0) for _ in range(20)] == [ref.uniform(-2.0, 3.0) for _ in range(20)]
        del st
    rng, ref = core.NpRand(7), np.random.RandomState(7)
    # bounded ints: numpy's legacy randint uses the same masked rejection on 32-bit draws
    assert [rng.int64n(10) for _ in range(100)] == [int(ref.randint(0, 10)) for _ in range(100)]
    assert [rng.int64(3, 1000) for _ in range(100)] == [int(ref.randint(3, 1000)) for _ in range(100)]
    rng, ref = core.NpRand(9), np.random.RandomState(9)
    assert rng.bytes(16) == ref.bytes(16)


def test_uuid4_is_valid_and_deterministic():
    a, b = core.NpRand(5), core.NpRand(5)
    u = a.uuid4()
    assert u == b.uuid4()
    parsed = uuid.UUID(u)
    assert parsed.version == 4 and parsed.variant == uuid.RFC_4122


def test_md5_tiebreak_hash():
    assert core.md5_low64("task-1") == int.from_bytes(hashlib.md5(b"task-1").digest()[:8], "little")


def _sim(searcher, hp=HP, **kw):
    return core.simulate(searcher, hp, **kw)


def test_single_and_random():
    r = _sim({"name": "single", "metric": "loss", "max_length": {"batches": 500}})
    assert r["results"] == {"500B V": 1}
    r = _sim({"name": "random", "metric": "loss", "max_trials": 7, "max_length": {"records": 64000}}, random_order=True)
    assert r["results"] == {"64000R V": 7}


def test_grid_cartesian_product_and_values():
    hp = {"global_batch_size": 32, "a": {"type": "int", "minval": 0, "maxval": 10, "count": 3},
          "b": {"type": "categorical", "vals": ["x", "y"]}, "c": {"type": "double", "minval": 0.0, "maxval": 1.0, "count": 2},
          "d": {"type": "log", "minval": -2, "maxval": 0, "base": 10, "count": 3}}
    s = core.Searcher({"name": "grid", "metric": "loss", "max_length": {"batches": 10}}, hp, seed=1)
    ops = s.initial_operations()
    creates = [o for o in ops if o["kind"] == "create"]
    assert len(creates) == 3 * 2 * 2 * 3 and len(ops) == 4 * len(creates)
    assert sorted({c["hparams"]["a"] for c in creates}) == [0, 5, 10]
    assert sorted({c["hparams"]["c"] for c in creates}) == [0.0, 1.0]
    assert sorted({round(c["hparams"]["d"], 6) for c in creates}) == [0.01, 0.1, 1.0]
    # int count is clamped to the size of the range
    hp2 = {"global_batch_size": 1, "a": {"type": "int", "minval": 1, "maxval": 2, "count": 10}}
    s2 = core.Searcher({"name": "grid", "metric": "loss", "max_length": {"batches": 1}}, hp2, seed=1)
    assert len([o for o in s2.initial_operations() if o["kind"] == "create"]) == 2


def test_sampling_is_reproducible_and_int_max_exclusive():
    cfg = {"name": "random", "metric": "loss", "max_trials": 200, "max_length": {"batches": 1}}
    a = core.Searcher(cfg, HP, seed=99).initial_operations()
    b = core.Searcher(cfg, HP, seed=99).initial_operations()
    c = core.Searcher(cfg, HP, seed=100).initial_operations()
    assert a == b and a != c
    layers = [o["hparams"]["layers"] for o in a if o["kind"] == "create"]
    assert min(layers) == 1 and max(layers) == 4
    lrs = [o["hparams"]["lr"] for o in a if o["kind"] == "create"]
    assert 1e-4 <= min(lrs) and max(lrs) < 1e-1
    seeds = [o["trial_seed"] for o in a if o["kind"] == "create"]
    assert all(0 <= s < 2 ** 31 for s in seeds)


def test_sync_halving_constant_validation():
    cfg = {"name": "sync_halving", "metric": "error", "num_rungs": 4, "max_length": {"batches": 800}, "budget": {"batches": 48000},
           "divisor": 4, "smaller_is_better": True}
    r = _sim(cfg, validation="constant")
    # rung lengths 12/50/200/800; promotions train only the difference
    keys = set(r["results"])
    assert any(k.startswith("12B V") for k in keys)
    assert "12B V 38B V 150B V 600B V" in r["results"]
    assert sum(r["results"].values()) == r["trials"]


def test_async_halving_shapes():
    cfg = {"name": "async_halving", "metric": "error", "num_rungs": 3, "max_length": {"batches": 900}, "max_trials": 12, "divisor": 3}
    r = _sim(cfg, validation="constant")
    assert r["trials"] == 12 and sum(r["results"].values()) == 12
    assert set(r["results"]) <= {"100B V", "100B V 200B V", "100B V 200B V 600B V"}
    assert r["results"]["100B V 200B V 600B V"] >= 1
    r2 = _sim(cfg, validation="random", random_order=True, simulation_seed=3)
    assert r2["trials"] == 12


def test_adaptive_family_and_pbt_complete():
    for cfg in (
        {"name": "adaptive", "metric": "error", "max_length": {"batches": 640}, "budget": {"batches": 64000}, "divisor": 4,
         "max_rungs": 3, "mode": "standard"},
        {"name": "adaptive_simple", "metric": "error", "max_length": {"batches": 640}, "max_trials": 20, "divisor": 4,
         "max_rungs": 3, "mode": "conservative"},
        {"name": "adaptive_asha", "metric": "error", "max_length": {"batches": 640}, "max_trials": 16, "divisor": 4,
         "max_rungs": 3, "mode": "aggressive"},
        {"name": "adaptive_asha", "metric": "error", "max_length": {"epochs": 32}, "max_trials": 16, "max_rungs": 5, "divisor": 4,
         "mode": "standard", "max_concurrent_trials": 8},
        {"name": "pbt", "metric": "error", "population_size": 6, "num_rounds": 4, "length_per_round": {"batches": 100},
         "replace_function": {"truncate_fraction": 0.34}, "explore_function": {"resample_probability": 0.3, "perturb_factor": 0.2}},
    ):
        for order in (False, True):
            r = _sim(cfg, validation="random", random_order=order, simulation_seed=11)
            assert r["trials"] >= 1 and sum(r["results"].values()) == r["trials"], cfg["name"]
    r = _sim({"name": "adaptive_asha", "metric": "error", "max_length": {"batches": 640}, "max_trials": 16, "divisor": 4,
              "max_rungs": 3, "mode": "aggressive"}, validation="constant")
    assert r["trials"] == 16
    # pbt: truncated trials are cloned from a checkpoint of a survivor
    r = _sim({"name": "pbt", "metric": "error", "population_size": 4, "num_rounds": 3, "length_per_round": {"batches": 50},
              "replace_function": {"truncate_fraction": 0.5}}, validation="random")
    assert any("C" in k for k in r["results"])


def test_facade_shutdown_and_early_exit():
    s = core.Searcher({"name": "random", "metric": "loss", "max_trials": 2, "max_length": {"batches": 10}}, HP, seed=3)
    ops = s.initial_operations()
    creates = [o for o in ops if o["kind"] == "create"]
    assert s.progress() == 0.0
    for tid, c in enumerate(creates, 1):
        assert s.trial_created(c, tid) == []
    s.workload_completed({"kind": "RUN_STEP"}, 10)
    assert s.progress() == pytest.approx(0.5)
    assert s.trial_exited_early(2, "ERRORED") == []          # random search tolerates early exits
    assert s.trial_closed(creates[0]["request_id"]) == []
    last = s.trial_closed(creates[1]["request_id"])
    assert last == [{"kind": "shutdown", "failure": False}]
    evs = s.events()
    assert [e["type"] for e in evs] == ["trial_created", "trial_created", "workload_completed", "trial_closed", "trial_closed"]
    with pytest.raises(core.CoreError):
        s.operation_completed(99, {"kind": "validate", "request_id": "x"}, {})


def test_config_resolution_and_validation():
    cfg = core.resolve_config({"entrypoint": "model_def:T", "hyperparameters": {"global_batch_size": 64},
                               "searcher": {"name": "single", "metric": "loss", "max_length": {"batches": 10}}}, default_seed=5)
    assert cfg["scheduling_unit"] == 100 and cfg["max_restarts"] == 5 and cfg["resources"]["slots_per_trial"] == 1
    assert cfg["optimizations"]["aggregation_frequency"] == 1 and cfg["optimizations"]["mixed_precision"] == "O0"
    assert cfg["checkpoint_storage"]["save_trial_best"] == 1 and cfg["reproducibility"]["experiment_seed"] == 5
    # a checkpoint_storage section that only sets retention policies keeps the default location; one of another type
    # replaces the location but keeps the policy defaults
    base = {"entrypoint": "model_def:T", "hyperparameters": {"global_batch_size": 64},
            "searcher": {"name": "single", "metric": "loss", "max_length": {"batches": 10}}}
    cs = core.resolve_config(dict(base, checkpoint_storage={"save_trial_latest": 5}))["checkpoint_storage"]
    assert cs["type"] == "shared_fs" and cs["host_path"] and cs["save_trial_latest"] == 5
    cs = core.resolve_config(dict(base, checkpoint_storage={"type": "s3", "bucket": "b"}))["checkpoint_storage"]
    assert cs["type"] == "s3" and "host_path" not in cs and cs["save_trial_best"] == 1
    errs = core.check_config({"entrypoint": "m:T", "hyperparameters": {}, "searcher": {"name": "random", "metric": "x"}})
    joined = " ".join(errs)
    assert "global_batch_size" in joined and "max_trials" in joined and "max_length" in joined
    assert any("unknown field" in e for e in core.check_config({"bogus": 1}))
    errs = core.check_config({"entrypoint": "m:T", "hyperparameters": {"global_batch_size": 8},
                              "searcher": {"name": "single", "metric": "l", "max_length": {"epochs": 2}}})
    assert any("records_per_epoch" in e for e in errs)
    errs = core.check_config({"entrypoint": "m:T", "hyperparameters": {"global_batch_size": 8, "x": {"type": "double", "minval": 0, "maxval": 1}},
                              "searcher": {"name": "grid", "metric": "l", "max_length": {"batches": 2}}})
    assert any("count" in e for e in errs)


def test_lttb():
    pts = [[float(i), float((i * 7) % 13)] for i in range(1000)]
    out = core.lttb(pts, 50)
    assert len(out) == 50 and out[0] == pts[0] and out[-1] == pts[-1]
    assert core.lttb(pts[:10], 50) == pts[:10]
