"""Exact expectations for the search methods: the multiset of per-trial operation sequences ("100B V 300B V") each
method must produce under a constant validation metric, and the bracket arithmetic of the adaptive family.

The expected values are facts about the algorithms (rung lengths, promotion counts, bracket sizes); they are the same
ones the reference pins in master/pkg/searcher/{asha,sha,adaptive,adaptive_simple,adaptive_asha}_test.go, re-stated here
against the C++ core through its C API.
"""
import collections

import pytest

from determined_b200.master import _core as core

HP = {"global_batch_size": 32}


def multiset(*pairs):
    out = collections.Counter()
    for seq, n in pairs:
        out[seq] += n
    return dict(out)


def sim(cfg, **kw):
    return core.simulate(cfg, HP, validation="constant", **kw)["results"]


@pytest.mark.parametrize("max_length,unit_seqs", [
    ({"records": 576000}, ("64000R V", "64000R V 128000R V", "64000R V 128000R V 384000R V")),
    ({"batches": 9000}, ("1000B V", "1000B V 2000B V", "1000B V 2000B V 6000B V")),
    ({"epochs": 12}, ("1E V", "1E V 3E V", "1E V 3E V 8E V")),
])
def test_async_halving_rungs_and_promotions(max_length, unit_seqs):
    cfg = {"name": "async_halving", "metric": "error", "num_rungs": 3, "max_length": max_length, "divisor": 3, "max_trials": 12}
    got = sim(cfg)
    a, b, c = unit_seqs
    assert got == multiset((a, 8), (b, 3), (c, 1))


def test_sync_halving_with_stragglers_records_and_batches():
    cfg = {"name": "sync_halving", "metric": "error", "num_rungs": 4, "max_length": {"records": 5120050}, "budget": {"records": 3072050},
           "divisor": 4, "train_stragglers": True}
    assert sim(cfg) == multiset(("80000R V", 9), ("80000R V 240003R V", 1), ("80000R V 240003R V 960009R V 3840038R V", 1))
    cfg = {"name": "sync_halving", "metric": "error", "num_rungs": 4, "max_length": {"batches": 80000}, "budget": {"batches": 48000},
           "divisor": 4, "train_stragglers": True}
    assert sim(cfg) == multiset(("1250B V", 9), ("1250B V 3750B V", 1), ("1250B V 3750B V 15000B V 60000B V", 1))


def test_adaptive_simple_corner_cases():
    base = {"name": "adaptive_simple", "metric": "error", "smaller_is_better": True, "max_length": {"batches": 100}, "max_trials": 1,
            "divisor": 4, "max_rungs": 3}
    assert sim(dict(base, mode="conservative")) == multiset(("100B V", 1), ("25B V 75B V", 1), ("6B V 19B V 75B V", 1))
    assert sim(dict(base, mode="aggressive")) == multiset(("6B V 19B V 75B V", 1))


def test_adaptive_standard_two_rungs():
    # budget 6400 over brackets [1 rung, 2 rungs] of max length 3200: one trial straight to 3200, two start at 800 and the
    # better one is promoted
    cfg = {"name": "adaptive", "metric": "error", "smaller_is_better": True, "max_length": {"batches": 3200}, "budget": {"batches": 6400},
           "mode": "standard", "max_rungs": 2, "divisor": 4}
    assert sim(cfg) == multiset(("800B V 2400B V", 1), ("800B V", 1), ("3200B V", 1))


@pytest.mark.parametrize("name,extra", [
    ("adaptive", {"budget": {"batches": 102400}, "train_stragglers": True, "mode": "aggressive"}),
    ("adaptive_simple", {"max_trials": 50, "mode": "conservative"}),
    ("adaptive_asha", {"max_trials": 128, "mode": "aggressive"}),
    ("async_halving", {"max_trials": 40, "num_rungs": 3}),
])
def test_same_seed_same_search(name, extra):
    cfg = dict({"name": name, "metric": "error", "smaller_is_better": True, "max_length": {"batches": 6400}, "divisor": 4}, **extra)
    if name not in ("async_halving",):
        cfg["max_rungs"] = 3
    hp = dict(HP, lr={"type": "log", "minval": -4, "maxval": -1, "base": 10}, layers={"type": "int", "minval": 1, "maxval": 9})
    a = core.simulate(cfg, hp, validation="constant", seed=17, simulation_seed=17)
    b = core.simulate(cfg, hp, validation="constant", seed=17, simulation_seed=17)
    assert a == b and a["trials"] >= 1


def test_adaptive_asha_bracket_arithmetic_through_behaviour():
    """Trials per bracket are weighted divisor^(rungs-1)/rungs and every bracket's bottom rung length is
    max_length / divisor^(rungs-1): with 20 trials, divisor 3 and brackets of 3, 2, 1 rungs the split is 12 / 5 / 3, so
    exactly 12 trials start at 100 batches, 5 at 300 and 3 run the full 900."""
    cfg = {"name": "adaptive_asha", "metric": "error", "smaller_is_better": True, "max_length": {"batches": 900}, "max_trials": 20,
           "divisor": 3, "mode": "conservative", "max_rungs": 3}
    got = sim(cfg)
    starts = collections.Counter(k.split()[0] for k, n in got.items() for _ in range(n))
    assert starts == {"100B": 12, "300B": 5, "900B": 3}
    # brackets of 3 and 2 rungs (standard mode, max_rungs 3), divisor 4: 50 trials split 37 / 13
    cfg = dict(cfg, max_trials=50, max_rungs=3, mode="standard", divisor=4, max_length={"batches": 1600})
    got = sim(cfg)
    starts = collections.Counter(k.split()[0] for k, n in got.items() for _ in range(n))
    assert starts == {"100B": 37, "400B": 13}
    # brackets of 4, 3 and 2 rungs, divisor 4: 100 trials split 70 / 22 / 8
    cfg = dict(cfg, max_trials=100, max_rungs=4, mode="standard", divisor=4, max_length={"batches": 6400})
    got = sim(cfg)
    starts = collections.Counter(k.split()[0] for k, n in got.items() for _ in range(n))
    assert starts == {"100B": 70, "400B": 22, "1600B": 8}
    # brackets of 4, 3 and 2 rungs, divisor 3: weights 27/4, 9/3, 3/2 -> 50 trials split 31 / 13 / 6 (remainder to the first)
    cfg = dict(cfg, max_trials=50, max_rungs=4, mode="standard", divisor=3, max_length={"batches": 2700})
    got = sim(cfg)
    starts = collections.Counter(k.split()[0] for k, n in got.items() for _ in range(n))
    assert starts == {"100B": 31, "300B": 13, "900B": 6}
