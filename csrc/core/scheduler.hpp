// scheduler.hpp — resource-pool scheduling as pure functions over value types: fair-share (max-min progressive
// filling over weighted groups with max_slots caps and a multi-slot deadlock breaker), strict priority with optional
// preemption, round-robin; and task→agent fitting (hard constraints, best/worst-fit score, md5 hash tie-break,
// single-agent fit first, else whole-agent "dedicated" multi-agent fit).
// Parity: reference master/internal/resourcemanagers/{fair_share.go:54-328, priority.go:80-203, round_robin.go:21-73,
// fitting.go:70-197, fitting_methods.go:9-73}.
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <vector>

#include "json.hpp"

namespace det {

struct AgentState {
  std::string id;
  std::string label;
  int num_slots = 0;
  std::map<int, std::string> devices;  // device index -> container id ("" = free)
  int max_zero_slot_containers = 100;
  std::map<std::string, bool> zero_slot_containers;
  bool enabled = true;

  int num_empty_slots() const { int n = 0; for (auto& d : devices) n += d.second.empty(); return n; }
  int num_used_slots() const { return static_cast<int>(devices.size()) - num_empty_slots(); }
  int num_zero_slot_containers() const { return static_cast<int>(zero_slot_containers.size()); }
  std::vector<int> allocate_free_devices(int slots, const std::string& container_id);
  void release_container(const std::string& container_id);
  static AgentState make(const std::string& id, int slots, const std::string& label = "", int max_zero = 100);
};

struct Allocation {
  std::string agent_id;
  int slots = 0;
  std::vector<int> devices;
  std::string container_id;
};

struct AllocateRequest {
  std::string id;        // task id
  std::string name;
  std::string group;     // group id (experiment)
  int slots_needed = 0;
  std::string label;
  bool non_preemptible = false;
  bool single_agent = false;
  long long registered = 0;            // monotone arrival stamp
  std::vector<Allocation> allocations; // empty = pending
  bool allocated() const { return !allocations.empty(); }
};

struct Group {
  std::string id;
  bool has_max_slots = false;
  int max_slots = 0;
  double weight = 1.0;
  bool has_priority = false;
  int priority = 42;
  long long registered = 0;
};

struct Fit {
  std::string agent_id;
  int slots = 0;
  double score = 0;
  uint64_t hash_distance = 0;
};

enum class FittingPolicy { Best, Worst };

std::vector<Fit> find_fits(const AllocateRequest& req, const std::map<std::string, AgentState>& agents, FittingPolicy policy);
void add_task_to_agents(std::map<std::string, AgentState>& agents, const std::vector<Fit>& fits, const std::string& container_prefix);

struct ScheduleDecision {
  std::vector<std::string> to_allocate;  // task ids, in order
  std::vector<std::string> to_release;   // task ids
};

constexpr int kMaxUserSchedulingPriority = 99;
constexpr int kDefaultSchedulingPriority = 42;

ScheduleDecision fairshare_schedule(const std::vector<AllocateRequest>& tasks, const std::map<std::string, Group>& groups,
                                    const std::map<std::string, AgentState>& agents, FittingPolicy policy);
ScheduleDecision priority_schedule(const std::vector<AllocateRequest>& tasks, const std::map<std::string, Group>& groups,
                                   const std::map<std::string, AgentState>& agents, FittingPolicy policy, bool preemption);
ScheduleDecision round_robin_schedule(const std::vector<AllocateRequest>& tasks, const std::map<std::string, Group>& groups,
                                      const std::map<std::string, AgentState>& agents, FittingPolicy policy);

// JSON front door used by the bindings and the tests: {"scheduler": "fair_share"|"priority"|"round_robin",
// "fitting_policy": "best"|"worst", "preemption": bool, "agents": [...], "groups": [...], "tasks": [...]}.
Json schedule_json(const Json& input);

uint64_t md5_low64(const std::string& s);

}  // namespace det
