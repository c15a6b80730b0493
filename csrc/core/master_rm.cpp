// master_rm.cpp — resource pool actor (agents, slots, groups, task list, 500 ms scheduler tick), trial logger actor.
#include <algorithm>

#include "master.hpp"

namespace det {

namespace {

class RootActor : public Actor {
 public:
  void receive(Context&, const Message& msg) override { msg.respond(Json()); }
};

// ---- resource pool ----------------------------------------------------------------------------------------------
class ResourcePoolActor : public Actor {
 public:
  explicit ResourcePoolActor(Master* m) : m_(m) {}
  void pre_start(Context& ctx) override { ctx.system().notify_after(ctx.self(), m_->cfg.scheduler_tick_ms, "tick"); }

  void receive(Context& ctx, const Message& msg) override {
    const std::string& t = msg.type;
    if (t == "tick") {
      schedule(ctx);
      ctx.system().notify_after(ctx.self(), m_->cfg.scheduler_tick_ms, "tick");
    } else if (t == "kick") {
      kick_pending_ = false;
      schedule(ctx);
    } else if (t == "add_agent") {
      if (agents_.count(msg.body["id"].as_string())) {   // heartbeat of a known agent: keep its slot assignments
        agent_info_[msg.body["id"].as_string()]["last_seen"] = msg.body["registered_time"];
        msg.respond(Json(true));
        return;
      }
      AgentState a = AgentState::make(msg.body["id"].as_string(), 0, msg.body["label"].str_or(""),
                                      static_cast<int>(msg.body["max_zero_slot_containers"].as_int(100)));
      Json devs = msg.body["devices"];
      for (auto& d : devs.arr()) a.devices[static_cast<int>(d["id"].as_int())] = "";
      a.num_slots = static_cast<int>(a.devices.size());
      agents_[a.id] = a;
      agent_info_[a.id] = msg.body;
      m_->log("info", "agent connected: " + a.id + " with " + std::to_string(a.num_slots) + " slots");
      msg.respond(Json(true));
      kick(ctx);
    } else if (t == "remove_agent") {
      const std::string id = msg.body["id"].as_string();
      agents_.erase(id);
      agent_info_.erase(id);
      for (auto& req : tasks_) {  // containers on a vanished agent are treated as terminated
        bool hit = false;
        for (auto& al : req.allocations) hit |= al.agent_id == id;
        if (hit && handlers_.count(req.id)) ctx.tell(handlers_[req.id], "agent_lost", Json(id));
      }
      msg.respond(Json(true));
    } else if (t == "enable_slot" || t == "disable_slot") {
      auto it = agents_.find(msg.body["agent"].as_string());
      if (it != agents_.end()) {
        const int slot = static_cast<int>(msg.body["slot"].as_int(-1));
        if (slot < 0) it->second.enabled = (t == "enable_slot");
        else disabled_slots_[it->first][slot] = (t == "disable_slot");
        if (slot >= 0 && t == "disable_slot" && it->second.devices.count(slot) && it->second.devices[slot].empty()) it->second.devices[slot] = "__disabled__";
        if (slot >= 0 && t == "enable_slot" && it->second.devices.count(slot) && it->second.devices[slot] == "__disabled__") it->second.devices[slot].clear();
      }
      msg.respond(Json(it != agents_.end()));
    } else if (t == "set_group") {
      Group& g = groups_[msg.body["group"].as_string()];
      if (g.id.empty()) { g.id = msg.body["group"].as_string(); g.registered = ++stamp_; }
      if (msg.body.contains("max_slots")) { g.has_max_slots = !msg.body["max_slots"].is_null(); g.max_slots = static_cast<int>(msg.body["max_slots"].as_int(0)); }
      if (msg.body.contains("weight")) g.weight = msg.body["weight"].as_double(1.0);
      if (msg.body.contains("priority") && !msg.body["priority"].is_null()) { g.has_priority = true; g.priority = static_cast<int>(msg.body["priority"].as_int(42)); }
      msg.respond(Json(true));
    } else if (t == "allocate") {
      AllocateRequest r;
      r.id = msg.body["task_id"].as_string();
      r.name = msg.body["name"].str_or(r.id);
      r.group = msg.body["group"].str_or(r.id);
      if (!groups_.count(r.group)) { Group g; g.id = r.group; g.registered = ++stamp_; groups_[r.group] = g; }
      r.slots_needed = static_cast<int>(msg.body["slots_needed"].as_int(0));
      r.label = msg.body["label"].str_or("");
      r.non_preemptible = msg.body["non_preemptible"].as_bool(false);
      r.single_agent = msg.body["single_agent"].as_bool(false);
      r.registered = ++stamp_;
      tasks_.erase(std::remove_if(tasks_.begin(), tasks_.end(), [&](const AllocateRequest& x) { return x.id == r.id; }), tasks_.end());
      tasks_.push_back(r);
      handlers_[r.id] = msg.sender;
      msg.respond(Json(true));
      kick(ctx);
    } else if (t == "release") {
      release(msg.body["task_id"].as_string());
      msg.respond(Json(true));
      kick(ctx);
    } else if (t == "summary") {
      msg.respond(summary());
    } else {
      msg.respond(Json());
    }
  }

 private:
  // The periodic tick (reference: 500 ms, resource_pool.go) stays as the backstop; a change that can make a waiting task
  // schedulable — a new request, freed slots, a new agent — also asks for a pass a few milliseconds later, coalesced, so a
  // short trial does not sit out half a tick on average before it gets its slot (a 16-trial search is dozens of hand-overs).
  void kick(Context& ctx) {
    if (kick_pending_) return;
    kick_pending_ = true;
    ctx.system().notify_after(ctx.self(), 5, "kick");
  }
  bool kick_pending_ = false;

  void release(const std::string& task_id) {
    for (auto& req : tasks_) {
      if (req.id != task_id) continue;
      for (auto& al : req.allocations) {
        auto it = agents_.find(al.agent_id);
        if (it != agents_.end()) it->second.release_container(al.container_id);
      }
    }
    tasks_.erase(std::remove_if(tasks_.begin(), tasks_.end(), [&](const AllocateRequest& x) { return x.id == task_id; }), tasks_.end());
    handlers_.erase(task_id);
  }

  void schedule(Context& ctx) {
    if (tasks_.empty()) return;
    const FittingPolicy policy = m_->cfg.fitting_policy == "worst" ? FittingPolicy::Worst : FittingPolicy::Best;
    ScheduleDecision d;
    try {
      if (m_->cfg.scheduler == "priority") d = priority_schedule(tasks_, groups_, agents_, policy, m_->cfg.preemption);
      else if (m_->cfg.scheduler == "round_robin") d = round_robin_schedule(tasks_, groups_, agents_, policy);
      else d = fairshare_schedule(tasks_, groups_, agents_, policy);
    } catch (const std::exception& e) {
      m_->log("error", std::string("scheduler failed: ") + e.what());
      return;
    }
    for (auto& id : d.to_allocate) {
      for (auto& req : tasks_) {
        if (req.id != id || req.allocated()) continue;
        auto fits = find_fits(req, agents_, policy);
        if (fits.empty()) break;
        Json allocs = Json::array();
        int rank = 0;
        for (auto& f : fits) {
          Allocation al;
          al.agent_id = f.agent_id;
          al.slots = f.slots;
          al.container_id = req.id + "-r" + std::to_string(run_counter_[req.id]) + "-c" + std::to_string(rank);
          al.devices = agents_.at(f.agent_id).allocate_free_devices(f.slots, al.container_id);
          req.allocations.push_back(al);
          Json a = Json::object();
          a["agent"] = al.agent_id; a["container_id"] = al.container_id; a["rank"] = rank++;
          a["addresses"] = agent_info_[al.agent_id]["addresses"];
          Json devs = Json::array();
          const Json& info = agent_info_[al.agent_id]["devices"];
          for (int dev : al.devices)
            for (auto& dj : info.arr()) if (dj["id"].as_int() == dev) devs.push_back(dj);
          a["devices"] = devs;
          allocs.push_back(a);
        }
        ++run_counter_[req.id];
        Json body = Json::object();
        body["allocations"] = allocs;
        if (handlers_.count(id)) ctx.tell(handlers_[id], "resources_allocated", body);
        break;
      }
    }
    for (auto& id : d.to_release)
      if (handlers_.count(id)) ctx.tell(handlers_[id], "release_resources");
  }

  Json summary() {
    Json out = Json::object();
    Json agents = Json::array();
    for (auto& kv : agents_) {
      Json a = Json::object();
      a["id"] = kv.first; a["label"] = kv.second.label; a["enabled"] = kv.second.enabled;
      a["num_slots"] = kv.second.num_slots; a["num_used_slots"] = kv.second.num_used_slots();
      a["registered_time"] = agent_info_[kv.first]["registered_time"];
      a["addresses"] = agent_info_[kv.first]["addresses"];
      Json slots = Json::array();
      for (auto& d : kv.second.devices) {
        Json s = Json::object();
        s["id"] = d.first;
        s["container_id"] = (d.second.empty() || d.second == "__disabled__") ? Json() : Json(d.second);
        s["enabled"] = d.second != "__disabled__";
        for (auto& dj : agent_info_[kv.first]["devices"].arr()) if (dj["id"].as_int() == d.first) s["device"] = dj;
        slots.push_back(s);
      }
      a["slots"] = slots;
      agents.push_back(a);
    }
    out["agents"] = agents;
    Json tasks = Json::array();
    for (auto& r : tasks_) {
      Json t = Json::object();
      t["id"] = r.id; t["name"] = r.name; t["group"] = r.group; t["slots_needed"] = r.slots_needed;
      t["state"] = r.allocated() ? "RUNNING" : "PENDING";
      Json cs = Json::array();
      for (auto& al : r.allocations) { Json c = Json::object(); c["agent"] = al.agent_id; c["container_id"] = al.container_id; c["slots"] = al.slots; cs.push_back(c); }
      t["containers"] = cs;
      tasks.push_back(t);
    }
    out["tasks"] = tasks;
    out["scheduler"] = m_->cfg.scheduler;
    out["fitting_policy"] = m_->cfg.fitting_policy;
    return out;
  }

  Master* m_;
  std::map<std::string, AgentState> agents_;
  std::map<std::string, Json> agent_info_;
  std::map<std::string, std::map<int, bool>> disabled_slots_;
  std::map<std::string, Group> groups_;
  std::vector<AllocateRequest> tasks_;
  std::map<std::string, ActorRef> handlers_;
  std::map<std::string, int> run_counter_;
  long long stamp_ = 0;
};

// ---- trial logger: batches rows (<= 1000 or 20 ms) before they hit the log store ----------------------------------------
class TrialLoggerActor : public Actor {
 public:
  explicit TrialLoggerActor(Master* m) : m_(m) {}
  void receive(Context& ctx, const Message& msg) override {
    if (msg.type == "log") {
      const long long trial = msg.body["trial_id"].as_int();
      pending_[trial].push_back(msg.body);
      ++count_;
      if (count_ >= 1000) flush();
      else if (!timer_armed_) { timer_armed_ = true; ctx.system().notify_after(ctx.self(), 20, "flush"); }
    } else if (msg.type == "flush") {
      timer_armed_ = false;
      flush();
      msg.respond(Json(true));
    }
  }
  void post_stop(Context&) override { flush(); }

 private:
  void flush() {
    for (auto& kv : pending_) m_->store.append_logs(kv.first, kv.second);
    pending_.clear();
    count_ = 0;
  }
  Master* m_;
  std::map<long long, std::vector<Json>> pending_;
  size_t count_ = 0;
  bool timer_armed_ = false;
};

}  // namespace

std::unique_ptr<Actor> make_resource_pool(Master* m) { return std::unique_ptr<Actor>(new ResourcePoolActor(m)); }
std::unique_ptr<Actor> make_trial_logger(Master* m) { return std::unique_ptr<Actor>(new TrialLoggerActor(m)); }
std::unique_ptr<Actor> make_root() { return std::unique_ptr<Actor>(new RootActor()); }

}  // namespace det
