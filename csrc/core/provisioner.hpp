// provisioner.hpp — elastic agent pool: decide how many agent instances to launch or terminate from the scheduler's
// view of demand, and drive a provider that can start/stop instances.
//
// Role of the reference's provisioner (master/internal/provisioner/scale_decider.go:14-259 + provisioner.go; AWS / GCP
// providers aws.go, gcp.go). The decision rules are the ones the reference documents — launch enough instances for the
// slots that pending tasks still need (bounded by max_instances), terminate instances that stopped, never connected
// within max_agent_starting_period, stayed idle longer than max_idle_agent_period, or exceed the desired pool size
// (idle first, then disconnected, then newest) — restated over plain JSON snapshots so they can be unit-tested without
// any cloud. Cloud SDKs do not exist offline, so the only provider shipped is LocalProcessProvider (one det-agent
// process per "instance", artificial or real slots), enough to exercise scale-up / scale-down end to end on one box.
#pragma once
#include <algorithm>
#include <string>
#include <vector>

#include "json.hpp"

namespace det {

struct ProvisionerConfig {
  int max_instances = 5;
  int slots_per_instance = 8;
  double max_idle_agent_period_s = 300;      // reference default: 5 min
  double max_agent_starting_period_s = 300;  // reference default: 5 min
  static ProvisionerConfig from_json(const Json& j) {
    ProvisionerConfig c;
    if (j.contains("max_instances")) c.max_instances = static_cast<int>(j["max_instances"].as_int(c.max_instances));
    if (j.contains("slots_per_instance")) c.slots_per_instance = static_cast<int>(j["slots_per_instance"].as_int(c.slots_per_instance));
    if (j.contains("max_idle_agent_period")) c.max_idle_agent_period_s = j["max_idle_agent_period"].as_double(c.max_idle_agent_period_s);
    if (j.contains("max_agent_starting_period")) c.max_agent_starting_period_s = j["max_agent_starting_period"].as_double(c.max_agent_starting_period_s);
    return c;
  }
};

// Snapshot in, decision out.
//   instances: [{"id", "state": "running|starting|stopped", "launch_time": seconds, "agent": "<agent id or empty>"}]
//   agents:    [{"id", "idle_since": seconds or null (busy), "connected": bool}]
//   demand:    {"pending_slots": N}   — slots that pending tasks need beyond what idle slots can satisfy
//   now:       seconds (any monotonic origin shared with launch_time / idle_since)
// Decision: {"launch": n, "terminate": [instance ids], "reasons": {id: why}}
inline Json scale_decide(const ProvisionerConfig& cfg, const Json& snapshot) {
  const double now = snapshot["now"].as_double(0);
  struct Inst { std::string id, state, agent; double launch; bool connected = false, idle = false; double idle_for = 0; };
  std::vector<Inst> inst;
  for (auto& j : snapshot["instances"].arr()) {
    Inst i;
    i.id = j["id"].as_string(); i.state = j["state"].str_or("running"); i.agent = j["agent"].str_or("");
    i.launch = j["launch_time"].as_double(now);
    inst.push_back(i);
  }
  for (auto& a : snapshot["agents"].arr()) {
    for (auto& i : inst) {
      if (i.agent.empty() || i.agent != a["id"].as_string()) continue;
      i.connected = a["connected"].as_bool(true);
      if (!a["idle_since"].is_null()) { i.idle = true; i.idle_for = now - a["idle_since"].as_double(now); }
    }
  }
  Json reasons = Json::object();
  std::vector<std::string> term;
  auto mark = [&](const Inst& i, const std::string& why) {
    if (std::find(term.begin(), term.end(), i.id) == term.end()) { term.push_back(i.id); reasons[i.id] = why; }
  };
  for (auto& i : inst) {
    if (i.state == "stopped") mark(i, "stopped");
    else if (!i.connected && now - i.launch > cfg.max_agent_starting_period_s) mark(i, "agent never connected / disconnected too long");
    else if (i.connected && i.idle && i.idle_for > cfg.max_idle_agent_period_s) mark(i, "idle too long");
  }
  const int pending = static_cast<int>(snapshot["demand"]["pending_slots"].as_int(0));
  int alive = 0;
  for (auto& i : inst) if (std::find(term.begin(), term.end(), i.id) == term.end()) ++alive;
  int want_new = pending > 0 ? (pending + cfg.slots_per_instance - 1) / cfg.slots_per_instance : 0;
  // instances still starting will bring slots of their own
  int starting = 0;
  for (auto& i : inst)
    if (!i.connected && std::find(term.begin(), term.end(), i.id) == term.end()) ++starting;
  want_new = std::max(0, want_new - starting);
  int launch = std::max(0, std::min(want_new, cfg.max_instances - alive));
  // over the cap (e.g. max_instances lowered): shed idle, then disconnected, then newest
  int excess = alive - cfg.max_instances;
  if (excess > 0) {
    std::vector<const Inst*> order;
    for (auto& i : inst) if (std::find(term.begin(), term.end(), i.id) == term.end()) order.push_back(&i);
    std::stable_sort(order.begin(), order.end(), [](const Inst* a, const Inst* b) {
      auto rank = [](const Inst* x) { return x->connected && x->idle ? 0 : (!x->connected ? 1 : 2); };
      if (rank(a) != rank(b)) return rank(a) < rank(b);
      return a->launch > b->launch;
    });
    for (int k = 0; k < excess && k < static_cast<int>(order.size()); ++k) mark(*order[k], "pool above max_instances");
    launch = 0;
  }
  Json out = Json::object();
  out["launch"] = static_cast<long long>(launch);
  Json t = Json::array();
  for (auto& id : term) t.push_back(Json(id));
  out["terminate"] = t;
  out["reasons"] = reasons;
  return out;
}

}  // namespace det
