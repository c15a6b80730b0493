// master_provisioner.cpp — elastic agent pool driven by the scale decider (provisioner.hpp).
//
// Every tick the actor snapshots the resource manager (which agents exist, which are idle and since when, how many
// slots pending tasks still need), asks scale_decide() what to do, and has the provider act on it. The provider
// shipped here starts one det-agent PROCESS per "instance" on this box (artificial CPU slots, or a slice of the GPUs
// through --visible-gpus) — enough to exercise scale-up on demand and scale-down on idleness end to end.
// `provider: {type: aws | gcp, ...}` switches to cloud VMs through the vendors' command-line clients
// (cloud_providers.hpp; reference master/internal/provisioner/aws.go, aws_spot.go, gcp.go): instances are found by tag /
// label, their own id is their agent id, and the same decider drives them.
#include <signal.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>

#include "cloud_providers.hpp"
#include "master.hpp"

namespace det {

static Json pobj(std::initializer_list<std::pair<const char*, Json>> kv) {
  Json j = Json::object();
  for (auto& p : kv) j[p.first] = p.second;
  return j;
}

namespace {

double mono_now() {
  using namespace std::chrono;
  return duration_cast<duration<double>>(steady_clock::now().time_since_epoch()).count();
}

class ProvisionerActor : public Actor {
 public:
  ProvisionerActor(Master* m, const Json& cfg) : m_(m), raw_(cfg), cfg_(ProvisionerConfig::from_json(cfg)) {
    tick_ms_ = static_cast<int>(cfg["tick_ms"].as_int(1000));
    agent_bin_ = cfg["agent_binary"].str_or(m->cfg.repo_root + "/determined_b200/_lib/det-agent");
    const std::string ptype = cfg["provider"]["type"].str_or("");
    if (ptype == "aws" || ptype == "gcp")
      cloud_.reset(new CloudCliProvider(cloud_config_from_json(cfg["provider"], m->cluster_id, m->cfg.host == "0.0.0.0" ? "127.0.0.1" : m->cfg.host,
                                                               m->http.port(), cfg_.slots_per_instance)));
    else if (!ptype.empty() && ptype != "local")
      throw ConfigError("provisioner.provider.type must be local, aws or gcp");
  }
  void pre_start(Context& ctx) override { ctx.system().notify_after(ctx.self(), tick_ms_, "tick"); }
  void post_stop(Context&) override {
    for (auto& kv : inst_) terminate(kv.first, kv.second.pid);
  }
  void receive(Context& ctx, const Message& msg) override {
    if (msg.type == "tick") {
      if (cloud_) cloud_step(); else step();
      ctx.system().notify_after(ctx.self(), tick_ms_, "tick");
    } else if (msg.type == "summary") {
      Json out = Json::array();
      for (auto& kv : inst_) out.push_back(pobj({{"id", kv.first}, {"pid", static_cast<long long>(kv.second.pid)}, {"launch_time", kv.second.launch}}));
      for (auto& kv : cloud_state_) out.push_back(pobj({{"id", kv.first}, {"state", kv.second}, {"launch_time", first_seen_[kv.first]}}));
      msg.respond(pobj({{"instances", out}, {"launched_total", static_cast<long long>(launched_total_)},
                        {"terminated_total", static_cast<long long>(terminated_total_)}, {"config", raw_}}));
    }
  }

 private:
  struct Inst { pid_t pid; double launch; };

  void step() {
    const double now = mono_now();
    // reap instances whose process died on its own
    for (auto it = inst_.begin(); it != inst_.end();) {
      int st = 0;
      if (waitpid(it->second.pid, &st, WNOHANG) == it->second.pid) {
        m_->system.ask(m_->rm, "remove_agent", pobj({{"id", it->first}}));
        it = inst_.erase(it);
      } else {
        ++it;
      }
    }
    Json summary = m_->system.ask(m_->rm, "summary");
    Json agents = Json::array();
    long long free_slots = 0;
    for (auto& a : summary["agents"].arr()) {
      const std::string id = a["id"].as_string();
      const long long used = a["num_used_slots"].as_int(0), total = a["num_slots"].as_int(0);
      if (a["enabled"].as_bool(true)) free_slots += total - used;
      if (!inst_.count(id)) continue;
      int containers = 0;
      for (auto& s : a["slots"].arr()) if (!s["container_id"].is_null()) ++containers;
      const bool idle = used == 0 && containers == 0;
      if (!idle) idle_since_.erase(id);
      else if (!idle_since_.count(id)) idle_since_[id] = now;
      agents.push_back(pobj({{"id", id}, {"connected", true}, {"idle_since", idle ? Json(idle_since_[id]) : Json()}}));
    }
    long long pending = 0;
    for (auto& t : summary["tasks"].arr())
      if (t["state"].as_string() == "PENDING") pending += t["slots_needed"].as_int(0);
    pending = std::max<long long>(0, pending - free_slots);
    Json instances = Json::array();
    for (auto& kv : inst_) instances.push_back(pobj({{"id", kv.first}, {"state", "running"}, {"launch_time", kv.second.launch}, {"agent", kv.first}}));
    // an instance whose agent has not registered yet must not look "connected"
    Json snapshot = pobj({{"now", now}, {"instances", instances}, {"agents", agents}, {"demand", pobj({{"pending_slots", pending}})}});
    Json d = scale_decide(cfg_, snapshot);
    for (auto& id : d["terminate"].arr()) {
      auto it = inst_.find(id.as_string());
      if (it == inst_.end()) continue;
      m_->log("info", "provisioner: terminating " + it->first + " (" + d["reasons"][it->first].str_or("") + ")");
      terminate(it->first, it->second.pid);
      m_->system.ask(m_->rm, "remove_agent", pobj({{"id", it->first}}));
      idle_since_.erase(it->first);
      inst_.erase(it);
      ++terminated_total_;
    }
    for (long long i = 0; i < d["launch"].as_int(0); ++i) launch(now);
  }

  // Cloud instances: the provider's listing is the truth about what exists; agents pair with instances by id.
  void cloud_step() {
    const double now = mono_now();
    std::vector<CloudInstance> listed;
    std::string err;
    if (!cloud_->list(&listed, &err)) { m_->log("warn", "provisioner: " + err); return; }
    cloud_state_.clear();
    for (auto& i : listed) {
      cloud_state_[i.id] = i.state;
      if (!first_seen_.count(i.id)) first_seen_[i.id] = now;
    }
    for (auto it = first_seen_.begin(); it != first_seen_.end();)
      if (!cloud_state_.count(it->first)) { m_->system.ask(m_->rm, "remove_agent", pobj({{"id", it->first}})); idle_since_.erase(it->first); it = first_seen_.erase(it); } else ++it;
    // launches the listing does not show yet (eventual consistency) still count as starting instances for a while
    if (static_cast<int>(listed.size()) >= expected_count_ || now - last_launch_ > 60.0) expected_count_ = 0;
    Json summary = m_->system.ask(m_->rm, "summary");
    Json agents = Json::array();
    long long free_slots = 0;
    for (auto& a : summary["agents"].arr()) {
      const std::string id = a["id"].as_string();
      const long long used = a["num_used_slots"].as_int(0), total = a["num_slots"].as_int(0);
      if (a["enabled"].as_bool(true)) free_slots += total - used;
      if (!cloud_state_.count(id)) continue;
      int containers = 0;
      for (auto& sl : a["slots"].arr()) if (!sl["container_id"].is_null()) ++containers;
      const bool idle = used == 0 && containers == 0;
      if (!idle) idle_since_.erase(id);
      else if (!idle_since_.count(id)) idle_since_[id] = now;
      agents.push_back(pobj({{"id", id}, {"connected", true}, {"idle_since", idle ? Json(idle_since_[id]) : Json()}}));
    }
    long long pending = 0;
    for (auto& t : summary["tasks"].arr())
      if (t["state"].as_string() == "PENDING") pending += t["slots_needed"].as_int(0);
    pending = std::max<long long>(0, pending - free_slots);
    Json instances = Json::array();
    for (auto& i : listed) instances.push_back(pobj({{"id", i.id}, {"state", i.state}, {"launch_time", first_seen_[i.id]}, {"agent", i.id}}));
    for (int k = static_cast<int>(listed.size()); k < expected_count_; ++k)
      instances.push_back(pobj({{"id", "pending-launch-" + std::to_string(k)}, {"state", "starting"}, {"launch_time", last_launch_}, {"agent", ""}}));
    Json d = scale_decide(cfg_, pobj({{"now", now}, {"instances", instances}, {"agents", agents}, {"demand", pobj({{"pending_slots", pending}})}}));
    std::vector<std::string> kill;
    for (auto& id : d["terminate"].arr())
      if (cloud_state_.count(id.as_string())) {
        kill.push_back(id.as_string());
        m_->log("info", "provisioner: terminating " + id.as_string() + " (" + d["reasons"][id.as_string()].str_or("") + ")");
      }
    if (!kill.empty()) {
      if (!cloud_->terminate(kill, &err)) m_->log("warn", "provisioner: " + err);
      else terminated_total_ += static_cast<long long>(kill.size());
      for (auto& id : kill) { m_->system.ask(m_->rm, "remove_agent", pobj({{"id", id}})); idle_since_.erase(id); }
    }
    const int n = static_cast<int>(d["launch"].as_int(0));
    if (n > 0) {
      if (!cloud_->launch(n, &err)) { m_->log("warn", "provisioner: " + err); return; }
      launched_total_ += n;
      expected_count_ = static_cast<int>(listed.size()) + n;
      last_launch_ = now;
      m_->log("info", "provisioner: requested " + std::to_string(n) + " " + cloud_->config().type + " instance(s)");
    }
  }

  void launch(double now) {
    const std::string id = "prov-" + std::to_string(++seq_);
    std::vector<std::string> argv = {agent_bin_, "--master-host", m_->cfg.host == "0.0.0.0" ? "127.0.0.1" : m_->cfg.host,
                                     "--master-port", std::to_string(m_->http.port()), "--agent-id", id};
    if (raw_["artificial"].as_bool(true)) { argv.push_back("--artificial-slots"); argv.push_back(std::to_string(cfg_.slots_per_instance)); }
    pid_t pid = fork();
    if (pid < 0) { m_->log("error", "provisioner: fork failed"); return; }
    if (pid == 0) {
      setpgid(0, 0);
      std::vector<char*> av;
      for (auto& a : argv) av.push_back(const_cast<char*>(a.c_str()));
      av.push_back(nullptr);
      execvp(av[0], av.data());
      _exit(127);
    }
    inst_[id] = Inst{pid, now};
    ++launched_total_;
    m_->log("info", "provisioner: launched " + id + " (pid " + std::to_string(pid) + ", " + std::to_string(cfg_.slots_per_instance) + " slots)");
  }

  static void terminate(const std::string&, pid_t pid) {
    if (pid <= 0) return;
    kill(-pid, SIGTERM);
    for (int i = 0; i < 50; ++i) {
      int st = 0;
      if (waitpid(pid, &st, WNOHANG) == pid) return;
      usleep(20 * 1000);
    }
    kill(-pid, SIGKILL);
    int st = 0;
    waitpid(pid, &st, 0);
  }

  Master* m_;
  Json raw_;
  ProvisionerConfig cfg_;
  int tick_ms_ = 1000;
  std::string agent_bin_;
  std::map<std::string, Inst> inst_;
  std::unique_ptr<CloudCliProvider> cloud_;
  std::map<std::string, std::string> cloud_state_;
  std::map<std::string, double> first_seen_;
  int expected_count_ = 0;
  double last_launch_ = 0;
  std::map<std::string, double> idle_since_;
  long long seq_ = 0, launched_total_ = 0, terminated_total_ = 0;
};

}  // namespace

std::unique_ptr<Actor> make_provisioner(Master* m, const Json& cfg) { return std::unique_ptr<Actor>(new ProvisionerActor(m, cfg)); }

}  // namespace det
