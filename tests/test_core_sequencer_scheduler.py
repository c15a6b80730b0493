"""C++ core: workload sequencer state machine and the three schedulers + fitting, driven through the C ABI.
Test ideas follow master/internal/trial_workload_sequencer_test.go and resourcemanagers/*_test.go of the reference."""
import pytest

from determined_b200.master import _core as core

RID = "11111111-1111-4111-8111-111111111111"


def make_seq(min_val=0, min_ckpt=0, policy="best", scheduling_unit=100, initial_validation=False, gbs=64,
             searcher_len=None, records_per_epoch=0):
    cfg = core.resolve_config({
        "entrypoint": "m:T", "hyperparameters": {"global_batch_size": gbs},
        "searcher": {"name": "single", "metric": "loss", "max_length": searcher_len or {"batches": 1000}},
        "min_validation_period": {"batches": min_val}, "min_checkpoint_period": {"batches": min_ckpt},
        "checkpoint_policy": policy, "scheduling_unit": scheduling_unit, "perform_initial_validation": initial_validation,
        "records_per_epoch": records_per_epoch})
    create = {"kind": "create", "request_id": RID, "trial_seed": 1, "hparams": {"global_batch_size": gbs}, "checkpoint": None}
    s = core.Sequencer(cfg, 7, create)
    s.set_trial_id(3)
    return s


def complete(seq, w, best=False, exited=None):
    msg = {"workload": w}
    if exited:
        msg["exited_reason"] = exited
    if w["kind"] == "RUN_STEP":
        msg["metrics"] = {"avg_metrics": {"loss": 1.0}}
    elif w["kind"] == "COMPUTE_VALIDATION_METRICS":
        msg["validation_metrics"] = {"num_inputs": 1, "validation_metrics": {"loss": 0.5}}
    elif w["kind"] == "CHECKPOINT_MODEL":
        msg["checkpoint_metrics"] = {"uuid": f"ckpt-{w['step_id']}", "resources": {}}
    return seq.workload_completed(msg, best)


def run_until_uptodate(seq, best=False, limit=100):
    kinds = []
    done_ops = []
    while not seq.up_to_date() and len(kinds) < limit:
        w = seq.workload()
        kinds.append((w["kind"], w["num_batches"], w["step_id"]))
        r = complete(seq, w, best)
        if r["op"]:
            done_ops.append(r["op"]["kind"])
    return kinds, done_ops


def test_train_is_sliced_into_scheduling_units_and_validate_checkpoints_first():
    s = make_seq(scheduling_unit=100)
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 250}})
    s.operation_requested({"kind": "validate", "request_id": RID})
    kinds, ops = run_until_uptodate(s)
    assert kinds == [("RUN_STEP", 100, 1), ("RUN_STEP", 100, 2), ("RUN_STEP", 50, 3), ("CHECKPOINT_MODEL", 0, 3),
                     ("COMPUTE_VALIDATION_METRICS", 0, 3)]
    assert ops == ["train", "validate"]
    assert s.state()["total_batches_processed"] == 250 and s.state()["latest_checkpoint"]["uuid"] == "ckpt-3"
    assert s.preclose_checkpoint_workload() is None  # nothing trained since the last checkpoint


def test_min_validation_and_checkpoint_periods():
    s = make_seq(min_val=100, min_ckpt=200, scheduling_unit=100, policy="none")
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 400}})
    kinds, _ = run_until_uptodate(s)
    assert [k for k, _, _ in kinds] == ["RUN_STEP", "COMPUTE_VALIDATION_METRICS", "RUN_STEP", "COMPUTE_VALIDATION_METRICS",
                                        "CHECKPOINT_MODEL", "RUN_STEP", "COMPUTE_VALIDATION_METRICS", "RUN_STEP"]


def test_checkpoint_policy_best_all_none():
    for policy, best, expect_ckpt in (("all", False, True), ("best", True, True), ("best", False, False), ("none", True, False)):
        s = make_seq(min_val=100, policy=policy)
        s.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 200}})
        complete(s, s.workload())                       # train 100
        w = s.workload()
        assert w["kind"] == "COMPUTE_VALIDATION_METRICS"
        complete(s, w, best=best)
        assert (s.workload()["kind"] == "CHECKPOINT_MODEL") == expect_ckpt, (policy, best)


def test_initial_validation_and_records_units():
    s = make_seq(initial_validation=True, gbs=64, searcher_len={"records": 6400})
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"records": 6400}})
    assert s.workload()["kind"] == "COMPUTE_VALIDATION_METRICS"
    complete(s, s.workload())
    w = s.workload()
    assert w["kind"] == "RUN_STEP" and w["num_batches"] == 100   # 6400 records / 64
    r = complete(s, w)
    assert r["op"]["kind"] == "train" and s.up_to_date()


def test_rollback_to_last_checkpoint_snapshot():
    s = make_seq(min_ckpt=100, policy="none")
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 300}})
    complete(s, s.workload())                 # step 1
    w = s.workload()
    assert w["kind"] == "CHECKPOINT_MODEL"
    complete(s, w)                            # snapshot here: 100 batches, step 1
    complete(s, s.workload())                 # step 2 (would be lost by a crash)
    assert s.state()["total_batches_processed"] == 200
    assert s.roll_back() == 1
    st = s.state()
    assert st["total_batches_processed"] == 100 and st["latest_checkpoint"]["uuid"] == "ckpt-1"
    w = s.workload()
    assert w["kind"] == "RUN_STEP" and w["step_id"] == 2 and w["total_batches_processed"] == 100


def test_graceful_stop_checkpoints_then_is_up_to_date_and_errored_exit_is_immediate():
    s = make_seq()
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 500}})
    complete(s, s.workload(), exited="USER_CANCELED")
    assert not s.up_to_date() and s.workload()["kind"] == "CHECKPOINT_MODEL"
    complete(s, s.workload())
    assert s.up_to_date()
    s2 = make_seq()
    s2.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 500}})
    complete(s2, s2.workload(), exited="ERRORED")
    assert s2.up_to_date()
    assert s2.terminate_workload()["kind"] == "TERMINATE"


def test_illegal_completion_is_rejected():
    s = make_seq()
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 500}})
    with pytest.raises(core.CoreError):
        s.workload_completed({"workload": {"kind": "COMPUTE_VALIDATION_METRICS", "experiment_id": 7, "trial_id": 3, "step_id": 0,
                                           "num_batches": 0, "total_batches_processed": 0}})


# ---- schedulers -----------------------------------------------------------------------------------------------------
def agents(*slots, label=""):
    return [{"id": f"agent{i}", "slots": s, "label": label} for i, s in enumerate(slots)]


def test_fair_share_splits_by_weight_and_respects_max_slots():
    req = {"scheduler": "fair_share", "agents": agents(8),
           "groups": [{"id": "g1", "weight": 1}, {"id": "g2", "weight": 1}],
           "tasks": [{"id": f"a{i}", "group": "g1", "slots_needed": 1} for i in range(8)] +
                    [{"id": f"b{i}", "group": "g2", "slots_needed": 1} for i in range(8)]}
    out = core.schedule(req)
    assert sum(t.startswith("a") for t in out["to_allocate"]) == 4 and sum(t.startswith("b") for t in out["to_allocate"]) == 4
    req["groups"][0]["max_slots"] = 2
    out = core.schedule(req)
    assert sum(t.startswith("a") for t in out["to_allocate"]) == 2 and sum(t.startswith("b") for t in out["to_allocate"]) == 6
    req["groups"] = [{"id": "g1", "weight": 3}, {"id": "g2", "weight": 1}]
    out = core.schedule(req)
    assert sum(t.startswith("a") for t in out["to_allocate"]) == 6


def test_fair_share_preempts_group_over_its_share():
    req = {"scheduler": "fair_share", "agents": agents(4),
           "groups": [{"id": "old"}, {"id": "new"}],
           "tasks": [{"id": f"o{i}", "group": "old", "slots_needed": 1, "allocations": [{"agent": "agent0", "slots": 1}]} for i in range(4)] +
                    [{"id": f"n{i}", "group": "new", "slots_needed": 1} for i in range(2)]}
    out = core.schedule(req)
    assert len(out["to_release"]) == 2 and all(t.startswith("o") for t in out["to_release"])


def test_priority_fifo_and_preemption():
    base = {"scheduler": "priority", "agents": agents(4),
            "groups": [{"id": "lo", "priority": 50}, {"id": "hi", "priority": 10}],
            "tasks": [{"id": "lo1", "group": "lo", "slots_needed": 4, "allocations": [{"agent": "agent0", "slots": 4}]},
                      {"id": "hi1", "group": "hi", "slots_needed": 2}]}
    assert core.schedule(dict(base, preemption=False)) == {"to_allocate": [], "to_release": [], "fits": {}}
    out = core.schedule(dict(base, preemption=True))
    assert out["to_release"] == ["lo1"] and out["to_allocate"] == []
    # strict ordering: a blocked higher-priority task blocks lower priorities
    req = {"scheduler": "priority", "agents": agents(4),
           "groups": [{"id": "p1", "priority": 1}, {"id": "p2", "priority": 2}],
           "tasks": [{"id": "big", "group": "p1", "slots_needed": 8}, {"id": "small", "group": "p2", "slots_needed": 1}]}
    assert core.schedule(req)["to_allocate"] == []
    req["tasks"][0]["slots_needed"] = 2
    assert core.schedule(req)["to_allocate"] == ["big", "small"]


def test_round_robin_alternates_groups():
    req = {"scheduler": "round_robin", "agents": agents(3),
           "groups": [{"id": "g1"}, {"id": "g2"}],
           "tasks": [{"id": "a1", "group": "g1", "slots_needed": 1}, {"id": "a2", "group": "g1", "slots_needed": 1},
                     {"id": "a3", "group": "g1", "slots_needed": 1}, {"id": "b1", "group": "g2", "slots_needed": 1}]}
    assert core.schedule(req)["to_allocate"] == ["a1", "b1", "a2"]


def test_fitting_single_agent_first_then_dedicated_multi_agent():
    # best fit packs onto the fuller agent
    req = {"scheduler": "fair_share", "fitting_policy": "best", "agents": agents(8, 8), "groups": [{"id": "g"}],
           "tasks": [{"id": "t0", "group": "g", "slots_needed": 2, "allocations": [{"agent": "agent1", "slots": 2}]},
                     {"id": "t1", "group": "g", "slots_needed": 2}]}
    assert core.schedule(req)["fits"]["t1"] == [{"agent": "agent1", "slots": 2}]
    req["fitting_policy"] = "worst"
    assert core.schedule(req)["fits"]["t1"] == [{"agent": "agent0", "slots": 2}]
    # 16 slots over two idle 8-slot agents
    req = {"scheduler": "fair_share", "agents": agents(8, 8), "groups": [{"id": "g"}],
           "tasks": [{"id": "big", "group": "g", "slots_needed": 16}]}
    fits = core.schedule(req)["fits"]["big"]
    assert sorted(f["agent"] for f in fits) == ["agent0", "agent1"] and all(f["slots"] == 8 for f in fits)
    # 12 slots cannot be split over whole 8-slot agents
    req["tasks"][0]["slots_needed"] = 12
    assert core.schedule(req)["to_allocate"] == []
    # zero-slot tasks are scheduled independently
    req = {"scheduler": "fair_share", "agents": agents(1), "groups": [{"id": "g"}],
           "tasks": [{"id": "cmd", "group": "g", "slots_needed": 0}, {"id": "t", "group": "g", "slots_needed": 1}]}
    assert sorted(core.schedule(req)["to_allocate"]) == ["cmd", "t"]
    # label mismatch
    req = {"scheduler": "fair_share", "agents": agents(4, label="a100"), "groups": [{"id": "g"}],
           "tasks": [{"id": "t", "group": "g", "slots_needed": 1, "label": "b200"}]}
    assert core.schedule(req)["to_allocate"] == []


def test_reference_scenario_validation_and_checkpoint_periods_with_rollback_and_cached_checkpoint():
    """The long scripted scenario the reference pins (master/internal/trial_workload_sequencer_test.go:17-268):
    Train(500) + Validate + Checkpoint requested up front, min_validation_period 200, min_checkpoint_period 400,
    checkpoint_policy none, scheduling_unit 100. Expected workload order, the pre-close checkpoint at each point, the
    roll-back to the last completed checkpoint, and the searcher Checkpoint op being satisfied by the checkpoint that was
    taken before the validation (a cached completion)."""
    s = make_seq(min_val=200, min_ckpt=400, policy="none", scheduling_unit=100, searcher_len={"batches": 500})
    assert s.up_to_date()
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 500}})
    assert not s.up_to_date()
    s.operation_requested({"kind": "validate", "request_id": RID})
    s.operation_requested({"kind": "checkpoint", "request_id": RID})

    def expect(kind, step, total):
        w = s.workload()
        assert (w["kind"], w["step_id"], w["total_batches_processed"]) == (kind, step, total), w
        return w

    w = expect("RUN_STEP", 1, 0)
    assert s.preclose_checkpoint_workload() is None                     # nothing trained yet
    assert complete(s, w)["op"] is None
    w = expect("RUN_STEP", 2, 100)
    pc = s.preclose_checkpoint_workload()
    assert (pc["kind"], pc["step_id"], pc["total_batches_processed"]) == ("CHECKPOINT_MODEL", 1, 100)
    assert complete(s, w)["op"] is None
    w = expect("COMPUTE_VALIDATION_METRICS", 2, 200)                     # min_validation_period
    assert s.preclose_checkpoint_workload()["step_id"] == 2
    assert complete(s, w)["op"] is None
    assert complete(s, expect("RUN_STEP", 3, 200))["op"] is None
    assert complete(s, expect("RUN_STEP", 4, 300))["op"] is None
    assert complete(s, expect("COMPUTE_VALIDATION_METRICS", 4, 400))["op"] is None
    assert complete(s, expect("CHECKPOINT_MODEL", 4, 400))["op"] is None  # min_checkpoint_period
    assert s.preclose_checkpoint_workload() is None
    r = complete(s, expect("RUN_STEP", 5, 400))
    assert r["op"]["kind"] == "train"                                    # the searcher's Train(500) is done
    expect("CHECKPOINT_MODEL", 5, 500)                                   # a searcher Validate checkpoints first
    assert s.roll_back() == 4                                            # back to the snapshot of the step-4 checkpoint
    r = complete(s, expect("RUN_STEP", 5, 400))                          # step 5 is replayed
    assert r["op"]["kind"] == "train"
    assert complete(s, expect("CHECKPOINT_MODEL", 5, 500))["op"] is None
    r = complete(s, expect("COMPUTE_VALIDATION_METRICS", 5, 500))
    assert r["op"]["kind"] == "validate"
    assert s.preclose_checkpoint_workload() is None
    r = s.complete_cached_checkpoints()
    assert r["op"]["kind"] == "checkpoint"                               # satisfied by the step-5 checkpoint already taken
    assert s.up_to_date()
    with pytest.raises(core.CoreError):
        s.workload()


def test_failed_checkpoint_marks_early_exit_and_short_record_lengths_round_up_to_one_batch():
    s = make_seq(min_ckpt=100, policy="none")
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"batches": 500}})
    complete(s, s.workload())
    w = s.workload()
    assert w["kind"] == "CHECKPOINT_MODEL"
    r = s.workload_completed({"workload": w, "exited_reason": "ERRORED"}, False)
    assert r["op"] is None and s.state()["exiting_early"] is True
    # Train(24 records) at global_batch_size 64 is one batch
    s = make_seq(gbs=64, searcher_len={"records": 24})
    s.operation_requested({"kind": "train", "request_id": RID, "length": {"records": 24}})
    w = s.workload()
    assert (w["kind"], w["num_batches"]) == ("RUN_STEP", 1)
    assert complete(s, w)["op"]["kind"] == "train"
