"""Scheduling scenarios pinned by the reference's scheduler unit tests (master/internal/resourcemanagers/priority_test.go,
fair_share_test.go, round_robin_test.go, fitting_test.go), replayed against the C++ schedulers through the C ABI. Each case
states the cluster, the groups, the tasks in arrival order, and which tasks the pass must start / ask to release."""
import copy

from determined_b200.master import _core as core

LOW, HIGH = 50, 40          # a smaller number is a higher priority


def cluster(*slots, label=""):
    return [{"id": f"agent{i + 1}", "slots": s, "label": label} for i, s in enumerate(slots)]


def two_groups():
    return [{"id": "group1", "priority": LOW}, {"id": "group2", "priority": HIGH}]


def task(i, slots, group, **kw):
    return dict({"id": f"task{i}", "slots_needed": slots, "group": group}, **kw)


def priority(agents, groups, tasks, preemption=False):
    return core.schedule({"scheduler": "priority", "preemption": preemption, "agents": agents, "groups": groups, "tasks": tasks})


def allocate(tasks, out):
    """Apply a pass: every started task now holds what the pass fitted for it."""
    tasks = copy.deepcopy(tasks)
    for t in tasks:
        if t["id"] in out["to_allocate"]:
            t["allocations"] = out["fits"].get(t["id"], [])
            if t["slots_needed"] == 0:
                t["allocations"] = t["allocations"] or [{"agent": "agent1", "slots": 0}]
    return tasks


# ---- priority, preemption off ------------------------------------------------------------------------------------------
def test_priority_order_backfill_within_a_level_and_zero_slot_tasks():
    tasks = [task(1, 4, "group1"), task(2, 1, "group1"), task(3, 1, "group2"), task(4, 0, "group2"), task(5, 4, "group2"),
             task(6, 0, "group1")]
    out = priority(cluster(4, 4), two_groups(), tasks)
    # the higher priority first (3, 5 and the slot-less 4); of the lower one, task 1 no longer fits on one agent, 2 does
    assert sorted(out["to_allocate"]) == ["task2", "task3", "task4", "task5", "task6"]
    # three more one-slot tasks of the lower priority: two slots are left
    tasks = allocate(tasks, out) + [task(7, 1, "group1"), task(8, 1, "group1"), task(9, 1, "group1")]
    assert priority(cluster(4, 4), two_groups(), tasks)["to_allocate"] == ["task7", "task8"]


def test_priority_a_waiting_higher_priority_task_blocks_the_lower_ones():
    tasks = [task(1, 4, "group1"), task(2, 1, "group1"), task(3, 12, "group2")]
    out = priority(cluster(4, 4), two_groups(), tasks)
    assert out["to_allocate"] == [] and out["to_release"] == []


def test_priority_labels_separate_the_queues():
    # the higher-priority task wants another label: it neither runs here nor holds these agents back
    tasks = [task(1, 4, "group1", label="label1"), task(2, 1, "group1", label="label1"), task(3, 4, "group2", label="label2")]
    out = priority(cluster(4, 4, label="label1"), two_groups(), tasks)
    assert sorted(out["to_allocate"]) == ["task1", "task2"]


def test_priority_all_slots_taken_then_lower_priority_waits_then_a_finish_frees_room():
    tasks = [task(1, 4, "group1"), task(2, 1, "group1"), task(3, 1, "group2"), task(4, 0, "group2"), task(5, 4, "group2"),
             task(6, 1, "group1"), task(7, 1, "group1")]
    out = priority(cluster(4, 4), two_groups(), tasks)
    assert sorted(out["to_allocate"]) == ["task2", "task3", "task4", "task5", "task6", "task7"]       # 1 + 4 + 1 + 1 + 1 = 8 slots
    tasks = allocate(tasks, out)
    again = priority(cluster(4, 4), two_groups(), tasks)
    assert again["to_allocate"] == [] and again["to_release"] == []          # full, and nothing is preempted
    # a one-slot task finishes: the 4-slot task still does not fit, a new one-slot task does
    tasks = [t for t in tasks if t["id"] != "task7"] + [task(8, 1, "group1")]
    assert priority(cluster(4, 4), two_groups(), tasks)["to_allocate"] == ["task8"]
    # everything finishes: the big one finally starts
    tasks = [task(1, 4, "group1")]
    assert priority(cluster(4, 4), two_groups(), tasks)["to_allocate"] == ["task1"]


# ---- priority, preemption on -------------------------------------------------------------------------------------------
def test_priority_preemption_releases_the_lower_priority_task():
    tasks = [task(1, 4, "group1", allocations=[{"agent": "agent1", "slots": 4}]), task(2, 0, "group1"), task(3, 4, "group2")]
    out = priority(cluster(4), two_groups(), tasks, preemption=True)
    assert out["to_allocate"] == ["task2"] and out["to_release"] == ["task1"]


def test_priority_preemption_does_not_help_a_task_that_can_never_fit():
    tasks = [task(1, 4, "group1"), task(2, 8, "group2")]
    out = priority(cluster(4), two_groups(), tasks, preemption=True)
    assert out["to_allocate"] == [] and out["to_release"] == []


def test_priority_preempts_one_task_for_one_slot():
    tasks = [task(1, 4, "group1"), task(2, 1, "group1"), task(3, 1, "group2"), task(4, 1, "group2"), task(5, 4, "group2"),
             task(6, 1, "group1")]
    out = priority(cluster(4, 4), two_groups(), tasks, preemption=True)
    assert sorted(out["to_allocate"]) == ["task2", "task3", "task4", "task5", "task6"] and out["to_release"] == []
    tasks = allocate(tasks, out) + [task(7, 1, "group2")]
    out = priority(cluster(4, 4), two_groups(), tasks, preemption=True)
    # one slot is needed: exactly one lower-priority task goes, the most recent one
    assert out["to_allocate"] == [] and out["to_release"] == ["task6"]
    tasks = [t for t in tasks if t["id"] != "task6"]
    out = priority(cluster(4, 4), two_groups(), tasks, preemption=True)
    assert out["to_allocate"] == ["task7"] and out["to_release"] == []


def test_priority_preempts_every_lower_priority_task_when_all_slots_are_wanted():
    tasks = [task(1, 4, "group1"), task(2, 1, "group1"), task(3, 1, "group1"), task(4, 1, "group1"), task(5, 1, "group1"), task(6, 0, "group1")]
    out = priority(cluster(4, 4), two_groups(), tasks, preemption=True)
    assert sorted(out["to_allocate"]) == ["task1", "task2", "task3", "task4", "task5", "task6"]
    tasks = allocate(tasks, out) + [task(7, 4, "group2"), task(8, 4, "group2")]
    out = priority(cluster(4, 4), two_groups(), tasks, preemption=True)
    # both agents are wanted whole: every slot-holding lower-priority task is released, the slot-less one stays
    assert out["to_allocate"] == [] and sorted(out["to_release"]) == ["task1", "task2", "task3", "task4", "task5"]
    tasks = [t for t in tasks if t["id"] not in out["to_release"]]
    out = priority(cluster(4, 4), two_groups(), tasks, preemption=True)
    assert sorted(out["to_allocate"]) == ["task7", "task8"] and out["to_release"] == []


def test_priority_preempts_nothing_for_an_equal_priority():
    tasks = [task(1, 4, "group1"), task(2, 4, "group1")]
    out = priority(cluster(4, 4), two_groups(), tasks, preemption=True)
    tasks = allocate(tasks, out) + [task(3, 4, "group1")]
    out = priority(cluster(4, 4), two_groups(), tasks, preemption=True)
    assert out["to_allocate"] == [] and out["to_release"] == []


def test_priority_preemption_across_many_priorities_takes_the_lowest_first():
    groups = [{"id": f"p{p}", "priority": p} for p in (10, 20, 30, 40)]
    tasks = [task(1, 2, "p20"), task(2, 2, "p30"), task(3, 2, "p40"), task(4, 2, "p40")]
    out = priority(cluster(4, 4), groups, tasks, preemption=True)
    assert sorted(out["to_allocate"]) == ["task1", "task2", "task3", "task4"]
    tasks = allocate(tasks, out) + [task(5, 4, "p10")]
    out = priority(cluster(4, 4), groups, tasks, preemption=True)
    assert out["to_allocate"] == [] and sorted(out["to_release"]) == ["task3", "task4"]          # the two priority-40 tasks, not 20 / 30


def test_priority_slotless_tasks_are_not_preempted_for_slots():
    tasks = [task(1, 0, "group1"), task(2, 4, "group1")]
    out = priority(cluster(4), two_groups(), tasks, preemption=True)
    assert sorted(out["to_allocate"]) == ["task1", "task2"]
    tasks = allocate(tasks, out) + [task(3, 4, "group2")]
    out = priority(cluster(4), two_groups(), tasks, preemption=True)
    assert out["to_release"] == ["task2"]            # the slot holder, not the slot-less task


# ---- fair share ----------------------------------------------------------------------------------------------------------
def fair(agents, groups, tasks):
    return core.schedule({"scheduler": "fair_share", "agents": agents, "groups": groups, "tasks": tasks})


def test_fair_share_max_slots_caps_a_group():
    groups = [{"id": "group1", "max_slots": 1}, {"id": "group2"}]
    tasks = [task(1, 1, "group1"), task(2, 1, "group1"), task(3, 1, "group2"), task(4, 1, "group2")]
    out = fair(cluster(4), groups, tasks)
    assert sum(t in out["to_allocate"] for t in ("task1", "task2")) == 1 and {"task3", "task4"} <= set(out["to_allocate"])


def test_fair_share_weights():
    groups = [{"id": "group1", "weight": 10}, {"id": "group2", "weight": 5}]
    tasks = [task(i, 1, "group1") for i in range(1, 9)] + [task(i, 1, "group2") for i in range(9, 17)]
    out = fair(cluster(8, 4), groups, tasks)
    got1 = sum(t["id"] in out["to_allocate"] for t in tasks if t["group"] == "group1")
    got2 = sum(t["id"] in out["to_allocate"] for t in tasks if t["group"] == "group2")
    assert (got1, got2) == (8, 4)


def test_fair_share_multi_slot_tasks_and_no_deadlock_on_odd_shares():
    groups = [{"id": "group1"}, {"id": "group2"}]
    # each group's half is 2 slots; both tasks need 2 and both run
    out = fair(cluster(4), groups, [task(1, 2, "group1"), task(2, 2, "group2")])
    assert sorted(out["to_allocate"]) == ["task1", "task2"]
    # two 3-slot tasks on 4 slots: the halves are too small for either, yet one of them must run rather than none
    out = fair(cluster(4), groups, [task(1, 3, "group1"), task(2, 3, "group2")])
    assert len(out["to_allocate"]) == 1


def test_fair_share_a_task_larger_than_the_cluster_does_not_block_others():
    groups = [{"id": "group1"}, {"id": "group2"}]
    out = fair(cluster(4), groups, [task(1, 5, "group1"), task(2, 1, "group2")])
    assert out["to_allocate"] == ["task2"]


def test_fair_share_running_tasks_keep_their_slots_when_shares_are_met():
    groups = [{"id": "group1"}, {"id": "group2"}]
    tasks = [task(1, 2, "group1", allocations=[{"agent": "agent1", "slots": 2}]),
             task(2, 2, "group2", allocations=[{"agent": "agent1", "slots": 2}]), task(3, 1, "group1"), task(4, 1, "group2")]
    out = fair(cluster(4), groups, tasks)
    assert out["to_allocate"] == [] and out["to_release"] == []


def test_fair_share_labels():
    groups = [{"id": "group1"}, {"id": "group2"}]
    tasks = [task(1, 1, "group1", label="a"), task(2, 1, "group2", label="b"), task(3, 1, "group2")]
    agents = [{"id": "agent1", "slots": 2, "label": "a"}, {"id": "agent2", "slots": 2, "label": ""}]
    out = fair(agents, groups, tasks)
    assert sorted(out["to_allocate"]) == ["task1", "task3"]
    assert out["fits"]["task1"][0]["agent"] == "agent1" and out["fits"]["task3"][0]["agent"] == "agent2"


def test_fair_share_non_preemptible_tasks_are_never_released():
    groups = [{"id": "group1"}, {"id": "group2"}]
    held = [{"agent": "agent1", "slots": 2}]
    tasks = [task(1, 2, "group1", allocations=held, non_preemptible=True), task(2, 2, "group1", allocations=held, non_preemptible=True),
             task(3, 2, "group2")]
    out = fair(cluster(4), groups, tasks)
    assert out["to_release"] == [] and out["to_allocate"] == []
    # the same picture with preemptible tasks: group1 is over its half and gives one back
    for t in tasks[:2]:
        t["non_preemptible"] = False
    out = fair(cluster(4), groups, tasks)
    assert len(out["to_release"]) == 1 and out["to_release"][0] in ("task1", "task2")


# ---- round robin ----------------------------------------------------------------------------------------------------------
def test_round_robin_respects_labels():
    groups = [{"id": "group1"}, {"id": "group2"}]
    tasks = [task(1, 1, "group1", label="a"), task(2, 1, "group2", label="b"), task(3, 1, "group1", label="a")]
    agents = [{"id": "agent1", "slots": 1, "label": "a"}, {"id": "agent2", "slots": 1, "label": "b"}]
    out = core.schedule({"scheduler": "round_robin", "agents": agents, "groups": groups, "tasks": tasks})
    assert sorted(out["to_allocate"]) == ["task1", "task2"]
    assert out["fits"]["task1"][0]["agent"] == "agent1" and out["fits"]["task2"][0]["agent"] == "agent2"
