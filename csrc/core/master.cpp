// master.cpp — Master bootstrap, configuration, container routing, command tasks, checkpoint GC, restore.
#include "elastic_logs.hpp"
#include "task_spec.hpp"
#include "master.hpp"

#include <sys/stat.h>

#include <algorithm>
#include <chrono>
#include <ctime>
#include <fstream>
#include <set>

namespace det {

static Json obj(std::initializer_list<std::pair<const char*, Json>> kv) {
  Json j = Json::object();
  for (auto& p : kv) j[p.first] = p.second;
  return j;
}

std::string now_rfc3339() {
  using namespace std::chrono;
  auto now = system_clock::now();
  std::time_t t = system_clock::to_time_t(now);
  char buf[40];
  std::tm tm{};
  gmtime_r(&t, &tm);
  strftime(buf, sizeof buf, "%Y-%m-%dT%H:%M:%S", &tm);
  char out[64];
  snprintf(out, sizeof out, "%s.%06ldZ", buf, static_cast<long>(duration_cast<microseconds>(now.time_since_epoch()).count() % 1000000));
  return out;
}
std::string random_hex(int bytes) {
  static thread_local std::mt19937_64 rng(std::random_device{}());
  static const char* hex = "0123456789abcdef";
  std::string s;
  for (int i = 0; i < bytes; ++i) { unsigned v = static_cast<unsigned>(rng() & 0xff); s.push_back(hex[v >> 4]); s.push_back(hex[v & 15]); }
  return s;
}
std::string md5_hex(const std::string& s) {
  // two 64-bit halves out of the md5 helper (low half of md5(s), low half of md5(s + "#")) — used only for
  // salted password digests inside this store, not for interchange
  char buf[40];
  snprintf(buf, sizeof buf, "%016llx%016llx", static_cast<unsigned long long>(md5_low64(s)), static_cast<unsigned long long>(md5_low64(s + "#")));
  return buf;
}

MasterConfig MasterConfig::from_json(const Json& j) {
  static const std::set<std::string> known = {"host", "port", "data_dir", "scheduler", "fitting_policy", "preemption", "checkpoint_storage",
                                              "python", "repo_root", "slot_type", "visible_gpus", "artificial_slots", "auth_required",
                                              "cluster_name", "scheduler_tick_ms", "terminate_timeout_ms", "all_ready_timeout_ms", "resource_pools", "db", "telemetry", "provisioner", "resource_manager",
                                              "security", "task_container_defaults", "logging", "warm_workers",
                                              // reference master.yaml keys: honoured (log, tensorboard_timeout, enable_cors) or accepted because they name
                                              // things this master does not have (a database, a static root, a harness wheel)
                                              "log", "tensorboard_timeout", "enable_cors", "config_file", "root", "harness_path"};
  MasterConfig c;
  for (auto& kv : j.obj())
    if (!known.count(kv.first)) throw ConfigError("unknown field in master config: '" + kv.first + "'");
  c.host = j["host"].str_or(c.host);
  c.port = static_cast<int>(j["port"].as_int(c.port));
  c.data_dir = j["data_dir"].str_or(c.data_dir);
  if (j["scheduler"].is_object()) {  // reference shape: scheduler: {type, fit, preemption}
    c.scheduler = j["scheduler"]["type"].str_or(c.scheduler);
    c.fitting_policy = j["scheduler"]["fit"].str_or(c.fitting_policy);
    c.preemption = j["scheduler"]["preemption"].as_bool(false);
  } else {
    c.scheduler = j["scheduler"].str_or(c.scheduler);
  }
  c.fitting_policy = j["fitting_policy"].str_or(c.fitting_policy);
  if (j.contains("preemption")) c.preemption = j["preemption"].as_bool(false);
  c.checkpoint_storage = j["checkpoint_storage"];
  c.python = j["python"].str_or(c.python);
  c.repo_root = j["repo_root"].str_or(c.repo_root);
  c.warm_workers = j["warm_workers"].str_or(c.warm_workers);
  if (c.warm_workers != "off" && c.warm_workers != "gpu" && c.warm_workers != "lazy" && c.warm_workers != "eager")
    throw ConfigError("warm_workers must be off, gpu, lazy or eager");
  c.slot_type = j["slot_type"].str_or(c.slot_type);
  c.visible_gpus = j["visible_gpus"].str_or("");
  c.artificial_slots = static_cast<int>(j["artificial_slots"].as_int(0));
  c.auth_required = j["auth_required"].as_bool(false) || j["security"]["auth_required"].as_bool(false);
  if (j["security"].is_object()) {
    static const std::set<std::string> sec_known = {"auth_required", "tls", "default_task"};
    for (auto& kv : j["security"].obj())
      if (!sec_known.count(kv.first)) throw ConfigError("unknown field in master config: 'security." + kv.first + "'");
    const Json& t = j["security"]["tls"];
    if (t.is_object()) {
      c.tls_cert = t["cert"].str_or("");
      c.tls_key = t["key"].str_or("");
      c.tls_cert_name = t["master_cert_name"].str_or("");
      if (c.tls_cert.empty() != c.tls_key.empty()) throw ConfigError("security.tls needs both cert and key");
    }
  }
  c.default_task = j["security"]["default_task"];
  if (c.default_task.is_object() && !c.default_task["uid"].is_number()) throw ConfigError("security.default_task needs uid (and gid, user, group)");
  c.task_container_defaults = j["task_container_defaults"];
  if (c.task_container_defaults.is_object()) {
    static const std::set<std::string> tcd_known = {"dtrain_network_interface", "nccl_port_range", "gloo_port_range", "shm_size_bytes", "network_mode",
                                                    "cpu_pod_spec", "gpu_pod_spec", "image", "registry_auth", "force_pull_image"};
    for (auto& kv : c.task_container_defaults.obj())
      if (!tcd_known.count(kv.first)) throw ConfigError("unknown field in master config: 'task_container_defaults." + kv.first + "'");
    if (c.task_container_defaults.contains("shm_size_bytes") && c.task_container_defaults["shm_size_bytes"].as_int(0) <= 0)
      throw ConfigError("task_container_defaults.shm_size_bytes must be > 0");
    for (const char* key : {"nccl_port_range", "gloo_port_range"}) {
      const std::string r = c.task_container_defaults[key].str_or("");
      if (r.empty()) continue;
      unsigned lo = 0, hi = 0;
      char tail = 0;
      if (sscanf(r.c_str(), "%u:%u%c", &lo, &hi, &tail) != 2 || lo > 65535 || hi > 65535)
        throw ConfigError(std::string("task_container_defaults.") + key + ": expected a port range of the form MIN:MAX, got '" + r + "'");
      if (lo > hi) throw ConfigError(std::string("task_container_defaults.") + key + ": the minimum exceeds the maximum");
    }
  }
  c.tensorboard_timeout_s = static_cast<int>(j["tensorboard_timeout"].as_int(c.tensorboard_timeout_s));
  if (c.tensorboard_timeout_s < 0) throw ConfigError("tensorboard_timeout must be >= 0");
  c.enable_cors = j["enable_cors"].as_bool(false);
  if (j["log"].is_object()) {
    c.log_level = j["log"]["level"].str_or(c.log_level);
    if (c.log_level == "warning") c.log_level = "warn";
    static const std::set<std::string> levels = {"trace", "debug", "info", "warn", "error", "fatal"};
    if (!levels.count(c.log_level)) throw ConfigError("log.level must be one of trace, debug, info, warn, error, fatal");
  }
  c.cluster_name = j["cluster_name"].str_or(c.cluster_name);
  c.provisioner = j["provisioner"];
  c.telemetry = j["telemetry"];
  c.logging = j["logging"];
  if (c.logging.is_object()) {
    const std::string t = c.logging["type"].str_or("default");
    if (t != "default" && t != "elastic") throw ConfigError("logging.type must be default or elastic");
    if (t == "elastic" && c.logging["host"].str_or("").empty()) throw ConfigError("logging.type elastic needs logging.host");
  }
  c.resource_manager = j["resource_manager"];
  if (c.resource_manager.is_object()) {
    const std::string t = c.resource_manager["type"].str_or("agent");
    if (t != "agent" && t != "kubernetes") throw ConfigError("resource_manager.type must be agent or kubernetes");
  }
  c.scheduler_tick_ms = static_cast<int>(j["scheduler_tick_ms"].as_int(c.scheduler_tick_ms));
  c.terminate_timeout_ms = static_cast<int>(j["terminate_timeout_ms"].as_int(c.terminate_timeout_ms));
  c.all_ready_timeout_ms = static_cast<int>(j["all_ready_timeout_ms"].as_int(c.all_ready_timeout_ms));
  if (c.scheduler != "fair_share" && c.scheduler != "priority" && c.scheduler != "round_robin") throw ConfigError("scheduler must be fair_share, priority or round_robin");
  if (c.fitting_policy != "best" && c.fitting_policy != "worst") throw ConfigError("fitting_policy must be best or worst");
  return c;
}
Json MasterConfig::to_json() const {
  return obj({{"host", host}, {"port", port}, {"data_dir", data_dir}, {"scheduler", scheduler}, {"fitting_policy", fitting_policy},
              {"preemption", preemption}, {"checkpoint_storage", checkpoint_storage}, {"python", python}, {"repo_root", repo_root},
              {"slot_type", slot_type}, {"artificial_slots", artificial_slots}, {"auth_required", auth_required},
              {"cluster_name", cluster_name}, {"scheduler_tick_ms", scheduler_tick_ms}, {"resource_manager", resource_manager},
              {"task_container_defaults", task_container_defaults}, {"tensorboard_timeout", tensorboard_timeout_s}, {"enable_cors", enable_cors},
              {"log", obj({{"level", log_level}})}});
}

// ---- command tasks (generic commands, notebooks, shells, tensorboards) ---------------------------------------------
namespace {
class CommandActor : public Actor {
 public:
  CommandActor(Master* m, long long id, const Json& spec) : m_(m), id_(id), spec_(spec) { task_id_ = "command-" + std::to_string(id); }
  void pre_start(Context& ctx) override {
    set_state("PENDING");
    m_->system.tell(m_->rm, "allocate", obj({{"task_id", task_id_}, {"name", spec_["description"].str_or(task_id_)}, {"group", task_id_},
                                             {"slots_needed", spec_["slots"].as_int(0)}, {"non_preemptible", true}, {"single_agent", true},
                                             {"label", spec_["config"]["resources"]["agent_label"].str_or("")}}), ctx.self());
  }
  void receive(Context& ctx, const Message& msg) override {
    if (msg.type == "resources_allocated") {
      const Json& a = msg.body["allocations"].at(0);
      ContainerSpec spec;
      spec.container_id = a["container_id"].as_string();
      spec.task_id = task_id_;
      for (auto& v : spec_["argv"].arr()) spec.argv.push_back(v.as_string());
      spec.cwd = spec_["cwd"].str_or("");
      for (auto& kv : spec_["env"].obj()) spec.env[kv.first] = kv.second.as_string();
      if (!spec.env.count("PYTHONPATH") && !m_->cfg.repo_root.empty()) {   // the package is importable in every task, as in the reference's images
        const char* pp = getenv("PYTHONPATH");
        spec.env["PYTHONPATH"] = m_->cfg.repo_root + (pp && *pp ? std::string(":") + pp : "");
      }
      m_->apply_owner_identity(m_->store.get("commands", id_)["owner"].str_or(""), spec.env);
      for (auto& d : a["devices"].arr()) spec.devices.push_back(Device{static_cast<int>(d["id"].as_int()), d["brand"].as_string(), d["uuid"].as_string(), d["type"].as_string()});
      if (spec_["config"].is_object() || spec_["has_context"].as_bool(false)) {   // commands / notebooks / shells take bind_mounts, environment and resources like trials do
        const std::string slot_type = (!spec.devices.empty() && spec.devices[0].type == "gpu" && spec.devices[0].brand != "Artificial") ? "gpu" : "cpu";
        std::string warning;
        const std::string bad = apply_task_config(spec_["config"], slot_type, 0, spec_["has_context"].as_bool(false), spec.env, spec.argv, &warning);
        if (!bad.empty()) m_->log("error", task_id_ + ": " + bad);
        if (!warning.empty()) m_->log("warn", task_id_ + ": " + warning);
      }
      // Service tasks (notebooks, TensorBoards, anything submitted with "proxy": true) get a port of their own: it replaces
      // "{port}" in argv, is exported as DET_TASK_PORT, and /proxy/<id>/... on the master forwards to it once the
      // readiness check has seen the service answer (reference: command.go readiness checks + internal/proxy).
      port_ = static_cast<int>(spec_["proxy_port"].as_int(0));
      if (port_ > 0) {
        const std::string ps = std::to_string(port_);
        for (auto& arg : spec.argv) {
          size_t at;
          while ((at = arg.find("{port}")) != std::string::npos) arg.replace(at, 6, ps);
          while ((at = arg.find("{prefix}")) != std::string::npos) arg.replace(at, 8, "/proxy/" + std::to_string(id_) + "/");
        }
        spec.env["DET_TASK_PORT"] = ps;
        spec.env["DET_TENSORBOARD_PORT"] = ps;
        spec.env["DET_PROXY_PREFIX"] = "/proxy/" + std::to_string(id_) + "/";
        const Json& addrs = a["addresses"];
        host_ = a["agent"].as_string() == "agent-local" ? "127.0.0.1" : (addrs.is_array() && addrs.size() ? addrs.at(0).str_or("127.0.0.1") : "127.0.0.1");
        m_->store.update("commands", id_, obj({{"proxy_host", host_}, {"proxy_port", port_}, {"is_ready", false},
                                               {"service_address", "/proxy/" + std::to_string(id_) + "/"}}));
        ctx.system().notify_after(ctx.self(), 300, "check_ready");
      }
      container_ = spec.container_id;
      m_->register_container(container_, ctx.self(), -id_);
      set_state("RUNNING");
      m_->start_container(a["agent"].as_string(), spec);
      if (spec_["type"].str_or("") == "tensorboard" && m_->cfg.tensorboard_timeout_s > 0) {
        // reference tensorboard_manager.go: a TensorBoard nobody has asked anything of for tensorboard_timeout seconds is stopped
        m_->store.update("commands", id_, obj({{"last_request_s", static_cast<long long>(std::time(nullptr))}}));
        ctx.system().notify_after(ctx.self(), 1000 * std::max(1, std::min(m_->cfg.tensorboard_timeout_s, 30)), "idle_check");
      }
    } else if (msg.type == "check_ready") {
      if (container_.empty() || ready_) return;
      std::string body;
      std::string ready_path = spec_["ready_path"].str_or("/");
      if (spec_["keep_prefix"].as_bool(false)) ready_path = "/proxy/" + std::to_string(id_) + ready_path;
      const int st = http_request(host_, port_, "GET", ready_path, "", &body, 2);
      if (st > 0 && st < 500) {
        ready_ = true;
        m_->store.update("commands", id_, obj({{"is_ready", true}, {"ready_time", now_rfc3339()}}));
        m_->log("info", task_id_ + " is ready at /proxy/" + std::to_string(id_) + "/");
      } else {
        ctx.system().notify_after(ctx.self(), 500, "check_ready");
      }
    } else if (msg.type == "idle_check") {
      if (container_.empty()) return;
      const long long idle = static_cast<long long>(std::time(nullptr)) - m_->store.get("commands", id_)["last_request_s"].as_int(0);
      if (idle >= m_->cfg.tensorboard_timeout_s) {
        m_->log("info", task_id_ + ": idle for " + std::to_string(idle) + " s; stopping it (tensorboard_timeout)");
        m_->signal_container(container_, SIGKILL);
      } else {
        ctx.system().notify_after(ctx.self(), 1000 * std::max(1, std::min(m_->cfg.tensorboard_timeout_s, 30)), "idle_check");
      }
    } else if (msg.type == "container_event") {
      if (msg.body["state"].as_string() == "TERMINATED") {
        m_->unregister_container(container_);
        container_.clear();
        m_->store.update("commands", id_, obj({{"exit_code", msg.body["exit_code"]}, {"end_time", now_rfc3339()}, {"is_ready", false}}));
        set_state("TERMINATED");
        ctx.tell(m_->rm, "release", obj({{"task_id", task_id_}}));
        ctx.stop_self();
      }
    } else if (msg.type == "kill" || msg.type == "release_resources") {
      if (!container_.empty()) m_->signal_container(container_, SIGKILL);
      else { set_state("TERMINATED"); ctx.tell(m_->rm, "release", obj({{"task_id", task_id_}})); ctx.stop_self(); }
      msg.respond(Json(true));
    } else if (msg.type == "log") {
      m_->store.append("command_logs", std::to_string(id_), msg.body);
    }
  }

 private:
  void set_state(const std::string& s) { m_->store.update("commands", id_, obj({{"state", s}})); }
  Master* m_;
  long long id_;
  Json spec_;
  std::string task_id_, container_, host_;
  int port_ = 0;
  bool ready_ = false;
};
}  // namespace

std::unique_ptr<Actor> make_command(Master* m, long long id, const Json& spec) { return std::unique_ptr<Actor>(new CommandActor(m, id, spec)); }

// ---- Master ------------------------------------------------------------------------------------------------------------
Master::Master(const MasterConfig& c) : cfg(c), store(c.data_dir), system(16) {
  containers.reset(new ContainerManager([this](const ContainerEvent& ev) { route_container_event(ev); }, [this](const LogLine& l) { route_log(l); }));
}
Master::~Master() { stop(); }

void Master::report_telemetry(const std::string& event, const Json& properties) {
  if (!cfg.telemetry.is_object() || !cfg.telemetry["enabled"].as_bool(false)) return;
  const std::string sink = cfg.telemetry["sink"].str_or(cfg.data_dir + "/telemetry.jsonl");
  Json e = obj({{"event", event}, {"timestamp", now_rfc3339()}, {"cluster_id", cluster_id}, {"version", "0.13.10.dev0+b200.1"},
                {"properties", properties}});
  std::lock_guard<std::mutex> g(log_mu_);
  std::ofstream f(sink, std::ios::app);
  if (f) f << e.dump() << "\n";
}
void Master::log(const std::string& level, const std::string& msg) {
  static const std::map<std::string, int> rank = {{"trace", 0}, {"debug", 1}, {"info", 2}, {"warn", 3}, {"error", 4}, {"fatal", 5}};
  auto at = rank.find(level), min = rank.find(cfg.log_level);
  if (at != rank.end() && min != rank.end() && at->second < min->second) return;      // below log.level
  Json e = obj({{"time", now_rfc3339()}, {"level", level}, {"message", msg}});
  {
    std::lock_guard<std::mutex> g(log_mu_);
    e["id"] = log_seq_++;
    log_ring_.push_back(e);
    if (log_ring_.size() > 25000) log_ring_.pop_front();
  }
  fprintf(stderr, "[%s] %s: %s\n", e["time"].as_string().c_str(), level.c_str(), msg.c_str());
}
std::vector<Json> Master::recent_logs(size_t limit, long long after_id) {
  std::lock_guard<std::mutex> g(log_mu_);
  std::vector<Json> out;
  if (after_id >= 0) {      // a follower: everything newer than what it has seen, oldest first
    for (auto& e : log_ring_)
      if (e["id"].as_int(0) > after_id && out.size() < limit) out.push_back(e);
    return out;
  }
  size_t start = log_ring_.size() > limit ? log_ring_.size() - limit : 0;
  for (size_t i = start; i < log_ring_.size(); ++i) out.push_back(log_ring_[i]);
  return out;
}

ActorRef Master::experiment_actor(long long id) { return system.get("experiments/" + std::to_string(id)); }

void Master::register_container(const std::string& cid, const ActorRef& owner, long long trial_id) {
  std::lock_guard<std::mutex> g(container_mu_);
  container_owner_[cid] = {owner, trial_id};
}
void Master::unregister_container(const std::string& cid) {
  std::lock_guard<std::mutex> g(container_mu_);
  container_owner_.erase(cid);
}
void Master::start_container(const std::string& agent_id, const ContainerSpec& spec) {
  if (agent_id == "agent-local") { containers->start(spec); return; }
  if (kube && KubernetesPods::owns(agent_id)) {
    // multi-container trials talk NCCL / rendezvous between nodes: give those pods the host network (reference task.go:286-289)
    const auto it = spec.env.find("DET_NUM_CONTAINERS");
    kube->create(agent_id, spec, it != spec.env.end() && atoi(it->second.c_str()) > 1);
    return;
  }
  Json j = obj({{"type", "start"}, {"container_id", spec.container_id}, {"task_id", spec.task_id}, {"description", spec.description},
                {"cwd", spec.cwd}});
  Json argv = Json::array();
  for (auto& a : spec.argv) argv.push_back(Json(a));
  j["argv"] = argv;
  Json env = Json::object();
  for (auto& kv : spec.env) env[kv.first] = kv.second;
  j["env"] = env;
  Json devs = Json::array();
  for (auto& d : spec.devices) devs.push_back(d.to_json());
  j["devices"] = devs;
  std::lock_guard<std::mutex> g(agent_cmd_mu_);
  container_agent_[spec.container_id] = agent_id;
  agent_cmds_[agent_id].push_back(j);
  agent_cmd_cv_.notify_all();
}
void Master::signal_container(const std::string& container_id, int sig) {
  if (kube && kube->has_container(container_id)) { kube->remove(container_id, sig == SIGKILL); return; }
  {
    std::lock_guard<std::mutex> g(agent_cmd_mu_);
    auto it = container_agent_.find(container_id);
    if (it != container_agent_.end()) {
      agent_cmds_[it->second].push_back(obj({{"type", "signal"}, {"container_id", container_id}, {"signal", static_cast<long long>(sig)}}));
      agent_cmd_cv_.notify_all();
      return;
    }
  }
  if (containers) containers->signal(container_id, sig);
}
Json Master::take_agent_commands(const std::string& agent_id, int wait_ms) {
  std::unique_lock<std::mutex> lk(agent_cmd_mu_);
  agent_cmd_cv_.wait_for(lk, std::chrono::milliseconds(wait_ms), [&] { return !agent_cmds_[agent_id].empty(); });
  Json out = Json::array();
  for (auto& c : agent_cmds_[agent_id]) out.push_back(c);
  agent_cmds_[agent_id].clear();
  return out;
}
void Master::route_container_event(const ContainerEvent& ev) {
  if (ev.state == "TERMINATED") {
    std::lock_guard<std::mutex> g(agent_cmd_mu_);
    container_agent_.erase(ev.container_id);
  }
  ActorRef owner;
  {
    std::lock_guard<std::mutex> g(container_mu_);
    auto it = container_owner_.find(ev.container_id);
    if (it != container_owner_.end()) owner = it->second.first;
  }
  if (owner) system.tell(owner, "container_event", obj({{"container_id", ev.container_id}, {"state", ev.state}, {"exit_code", ev.exit_code}, {"failure", ev.failure}}));
}
void Master::route_log(const LogLine& l) {
  ActorRef owner;
  long long trial_id = 0;
  {
    std::lock_guard<std::mutex> g(container_mu_);
    auto it = container_owner_.find(l.container_id);
    if (it != container_owner_.end()) { owner = it->second.first; trial_id = it->second.second; }
  }
  std::string agent_id = "agent-local";
  {
    std::lock_guard<std::mutex> g(agent_cmd_mu_);
    auto it = container_agent_.find(l.container_id);
    if (it != container_agent_.end()) agent_id = it->second;
  }
  // level: what the line itself says ("WARNING: ...", "[ERROR] ...", python logging's "LEVEL:name:message"), INFO otherwise
  std::string level = "INFO";
  {
    size_t at = 0;
    while (at < l.message.size() && (l.message[at] == '[' || l.message[at] == ' ')) ++at;
    for (const char* name : {"TRACE", "DEBUG", "INFO", "WARNING", "ERROR", "CRITICAL"}) {
      const size_t n = strlen(name);
      if (l.message.compare(at, n, name) == 0 && at + n < l.message.size() && (l.message[at + n] == ':' || l.message[at + n] == ']')) { level = name; break; }
    }
  }
  Json row = obj({{"trial_id", trial_id}, {"container_id", l.container_id}, {"timestamp", now_rfc3339()}, {"stdtype", l.stdtype},
                  {"rank", l.rank >= 0 ? Json(l.rank) : Json()}, {"message", l.message}, {"source", "agent"}, {"agent_id", agent_id},
                  {"level", level}});
  if (trial_id > 0) system.tell(logger, "log", row);
  else if (owner) system.tell(owner, "log", row);
}

// A submitted context ([{"path": "...", "content": "<text>", "mode": 0644}]; directories are implied by paths) laid out under
// `dir`. Paths that would leave the directory are dropped.
void Master::write_context_files(const std::string& dir, const Json& files) {
  mkdirs(dir);
  if (!files.is_array()) return;
  for (auto& f : files.arr()) {
    const std::string rel = f["path"].str_or("");
    if (rel.empty() || rel.find("..") != std::string::npos || rel[0] == '/') continue;
    const std::string full = dir + "/" + rel;
    auto slash = full.rfind('/');
    if (slash != std::string::npos) mkdirs(full.substr(0, slash));
    if (f["type"].str_or("file") == "dir") { mkdirs(full); continue; }
    std::ofstream out(full, std::ios::binary);
    out << f["content"].str_or("");
    out.close();
    if (f["mode"].is_number()) chmod(full.c_str(), static_cast<mode_t>(f["mode"].as_int()));
  }
}

// Tasks of a user who is linked to an agent user run under that identity on the agents (agent.hpp: drop_to_task_user).
void Master::apply_owner_identity(const std::string& owner, std::map<std::string, std::string>& env) {
  Json aug = owner.empty() ? Json() : store.get("users", owner)["agent_user_group"];
  if (!aug.is_object() || !aug["uid"].is_number()) aug = cfg.default_task;      // security.default_task: everyone who is not linked
  if (!aug.is_object() || !aug["uid"].is_number()) return;
  env["DET_RUN_AS_UID"] = std::to_string(aug["uid"].as_int());
  env["DET_RUN_AS_GID"] = std::to_string(aug["gid"].as_int(aug["uid"].as_int()));
  if (aug["user"].is_string()) env["DET_RUN_AS_USER"] = aug["user"].as_string();
  if (aug["group"].is_string()) env["DET_RUN_AS_GROUP"] = aug["group"].as_string();
}

long long Master::create_experiment(const Json& config, const Json& model_definition, long long parent_id, bool activate, const std::string& owner) {
  const long long id = store.next_id("experiments");
  const std::string dir = model_def_dir(id);
  write_context_files(dir, model_definition);
  store.put("experiments", id, obj({{"id", id}, {"state", activate ? "ACTIVE" : "PAUSED"}, {"config", config}, {"parent_id", parent_id > 0 ? Json(parent_id) : Json()},
                                    {"start_time", now_rfc3339()}, {"end_time", Json()}, {"archived", false}, {"progress", 0.0}, {"owner", owner},
                                    {"description", config["description"]}, {"labels", config["labels"]}, {"best_validation", Json()}}));
  system.actor_of("experiments/" + std::to_string(id), make_experiment(this, id, false), experiments_root);
  log("info", "created experiment " + std::to_string(id));
  report_telemetry("experiment_created", obj({{"id", id}, {"searcher", config["searcher"]["name"]}, {"slots_per_trial", config["resources"]["slots_per_trial"]},
                                              {"num_hparams", static_cast<long long>(config["hyperparameters"].size())}}));
  return id;
}

// Pick checkpoints outside save_experiment_best / save_trial_best / save_trial_latest (never ones used for warm
// starts), mark them DELETED and run the GC task that removes their files. (reference checkpoint_gc.go:15-88,
// postgres.go:1043-1156)
void Master::spawn_checkpoint_gc(long long experiment_id) {
  Json exp = store.get("experiments", experiment_id);
  if (exp.is_null()) return;
  const Json& cs = exp["config"]["checkpoint_storage"];
  const int keep_exp = static_cast<int>(cs["save_experiment_best"].as_int(0)), keep_best = static_cast<int>(cs["save_trial_best"].as_int(1)),
            keep_latest = static_cast<int>(cs["save_trial_latest"].as_int(1));
  const std::string metric = exp["config"]["searcher"]["metric"].str_or("");
  const bool smaller = exp["config"]["searcher"]["smaller_is_better"].is_null() ? true : exp["config"]["searcher"]["smaller_is_better"].as_bool(true);
  auto ckpts = store.list("checkpoints", [&](const Json& r) { return r["experiment_id"].as_int() == experiment_id && r["state"].as_string() == "COMPLETED"; });
  std::set<std::string> warm;
  for (auto& t : store.list("trials")) if (t["warm_start_checkpoint"].is_object()) warm.insert(t["warm_start_checkpoint"]["uuid"].str_or(""));
  auto metric_of = [&](const Json& c, double* out) {
    for (auto& v : store.series("validations", std::to_string(c["trial_id"].as_int())))
      if (v["total_batches"].as_int() == c["total_batches"].as_int()) {
        const Json& m = v["metrics"]["validation_metrics"][metric];
        if (m.is_number()) { *out = smaller ? m.as_double() : -m.as_double(); return true; }
      }
    return false;
  };
  std::set<std::string> keep;
  std::map<long long, std::vector<Json>> by_trial;
  for (auto& c : ckpts) by_trial[c["trial_id"].as_int()].push_back(c);
  std::vector<std::pair<double, std::string>> exp_ranked;
  for (auto& kv : by_trial) {
    auto& v = kv.second;
    std::sort(v.begin(), v.end(), [](const Json& a, const Json& b) { return a["total_batches"].as_int() > b["total_batches"].as_int(); });
    for (int i = 0; i < keep_latest && i < static_cast<int>(v.size()); ++i) keep.insert(v[static_cast<size_t>(i)]["uuid"].as_string());
    std::vector<std::pair<double, std::string>> ranked;
    for (auto& c : v) { double m; if (metric_of(c, &m)) { ranked.push_back({m, c["uuid"].as_string()}); exp_ranked.push_back({m, c["uuid"].as_string()}); } }
    std::sort(ranked.begin(), ranked.end());
    for (int i = 0; i < keep_best && i < static_cast<int>(ranked.size()); ++i) keep.insert(ranked[static_cast<size_t>(i)].second);
  }
  std::sort(exp_ranked.begin(), exp_ranked.end());
  for (int i = 0; i < keep_exp && i < static_cast<int>(exp_ranked.size()); ++i) keep.insert(exp_ranked[static_cast<size_t>(i)].second);
  Json to_delete = Json::array();
  for (auto& c : ckpts) {
    const std::string uuid = c["uuid"].as_string();
    if (keep.count(uuid) || warm.count(uuid)) continue;
    store.update("checkpoints", uuid, obj({{"state", "DELETED"}}));
    to_delete.push_back(obj({{"uuid", uuid}, {"resources", c["resources"]}}));
  }
  if (to_delete.size() == 0) return;
  const long long cmd_id = store.next_id("commands");
  Json spec = obj({{"type", "gc"}, {"description", "checkpoint GC (experiment " + std::to_string(experiment_id) + ")"}, {"slots", 0},
                   {"argv", Json(Json::array_t{Json(cfg.python), Json("-m"), Json("determined_b200.exec.gc_checkpoints")})},
                   {"env", obj({{"DET_EXPERIMENT_ID", std::to_string(experiment_id)}, {"DET_EXPERIMENT_CONFIG", exp["config"].dump()},
                                {"DET_DELETE", to_delete.dump()}, {"PYTHONPATH", cfg.repo_root}})}});
  store.put("commands", cmd_id, obj({{"id", cmd_id}, {"type", "gc"}, {"state", "PENDING"}, {"spec", spec}, {"start_time", now_rfc3339()}}));
  system.actor_of("commands/" + std::to_string(cmd_id), make_command(this, cmd_id, spec), commands_root);
  log("info", "checkpoint GC for experiment " + std::to_string(experiment_id) + ": deleting " + std::to_string(to_delete.size()) + " checkpoints");
}

bool Master::check_auth(const HttpRequest& req, std::string* user) {
  auto it = req.headers.find("authorization");
  std::string token;
  if (it != req.headers.end() && it->second.rfind("Bearer ", 0) == 0) token = it->second.substr(7);
  if (!token.empty()) {
    Json s = store.get("sessions", token);
    if (!s.is_null()) {     // a session of a deactivated (or removed) user is no session
      const Json u = store.get("users", s["username"].as_string());
      if (!u.is_null() && u["active"].as_bool(true)) { if (user) *user = s["username"].as_string(); return true; }
    }
  }
  if (user) *user = "determined";
  return !cfg.auth_required;
}

void Master::restore_experiments() {
  // non-terminal experiments come back through searcher-event replay (reference core.go:453-466)
  for (auto& e : store.list("experiments", [](const Json& r) { return !is_terminal_state(r["state"].as_string()); })) {
    const long long id = e["id"].as_int();
    system.actor_of("experiments/" + std::to_string(id), make_experiment(this, id, true), experiments_root);
  }
  for (auto& c : store.list("commands", [](const Json& r) { return r["state"].as_string() != "TERMINATED"; }))
    store.update("commands", c["id"].as_int(), obj({{"state", "TERMINATED"}, {"exit_code", -1}}));  // processes did not survive the master
}

int Master::start() {
  started_at = static_cast<long long>(std::time(nullptr));
  if (const char* tf = getenv("DET_ACTOR_TRACE")) system.enable_trace(tf);
  Json cl = store.get("cluster", "id");
  if (cl.is_null()) { cluster_id = random_hex(16); store.put("cluster", "id", obj({{"cluster_id", cluster_id}})); }
  else cluster_id = cl["cluster_id"].as_string();
  for (const char* u : {"determined", "admin"})
    if (!store.exists("users", u)) store.put("users", u, obj({{"username", u}, {"admin", std::string(u) == "admin"}, {"active", true}, {"password_hash", ""}, {"salt", ""}}));
  experiments_root = system.actor_of("experiments", make_root());
  commands_root = system.actor_of("commands", make_root());
  rm = system.actor_of("rm", make_resource_pool(this));
  logger = system.actor_of("trial_logger", make_trial_logger(this));
  const bool k8s = cfg.resource_manager.is_object() && cfg.resource_manager["type"].str_or("agent") == "kubernetes";
  if (k8s) {
    // Pods replace agents: every schedulable node with allocatable slot resources joins the pool as "k8s-<node>".
    KubeConfig kc = KubeConfig::from_json(cfg.resource_manager);
    if (!kc.pod_spec_overrides.is_object()) {
      // task_container_defaults.gpu_pod_spec / cpu_pod_spec (a Pod manifest; its .spec is what is merged into task pods)
      const bool gpu_pool = cfg.slot_type != "cpu";
      const Json& pod = cfg.task_container_defaults[gpu_pool ? "gpu_pod_spec" : "cpu_pod_spec"];
      if (pod.is_object()) kc.pod_spec_overrides = pod["spec"].is_object() ? pod["spec"] : pod;
    }
    kube.reset(new KubernetesPods(
        kc, cluster_id, [this](const ContainerEvent& ev) { route_container_event(ev); }, [this](const LogLine& l) { route_log(l); },
        [this, kc](const std::string& node, int slots, bool present) {
          const std::string id = "k8s-" + node;
          if (!present) { system.tell(rm, "remove_agent", obj({{"id", id}})); return; }
          Json devs = Json::array();
          for (int i = 0; i < slots; ++i) devs.push_back(Device{i, kc.slot_resource, node + "-" + std::to_string(i), "gpu"}.to_json());
          system.tell(rm, "add_agent", obj({{"id", id}, {"label", ""}, {"devices", devs}, {"registered_time", now_rfc3339()},
                                            {"addresses", Json(Json::array_t{Json(node)})}, {"driver_version", "kubernetes"}}));
        }));
  } else {
    devices = detect_devices(cfg.slot_type, cfg.visible_gpus, cfg.artificial_slots);
    containers->set_pycache_dir(cfg.data_dir + "/pycache");
    containers->set_slot_count(static_cast<int>(devices.size()));
    containers->enable_warm(cfg.warm_workers, cfg.python, cfg.repo_root);
    if (cfg.warm_workers == "eager")
      for (auto& d : devices) containers->prewarm(d.type == "gpu" && d.brand != "Artificial" ? std::to_string(d.id) : std::string("cpu"));
    Json devs = Json::array();
    for (auto& d : devices) devs.push_back(d.to_json());
    system.ask(rm, "add_agent", obj({{"id", "agent-local"}, {"label", ""}, {"devices", devs}, {"registered_time", now_rfc3339()},
                                     {"addresses", Json(Json::array_t{Json("127.0.0.1")})}, {"driver_version", nvidia_driver_version()}}));
  }
  if (cfg.logging.is_object() && cfg.logging["type"].str_or("default") == "elastic") {
    auto es = std::make_shared<ElasticLogs>(cfg.logging["host"].as_string(), static_cast<int>(cfg.logging["port"].as_int(9200)),
                                            cfg.logging["index"].str_or("determined-triallogs"));
    store.log_append_hook = [this, es](long long trial, const std::vector<Json>& lines) {
      std::string err;
      if (!es->append(trial, lines, &err)) log("error", "trial logs: " + err);
    };
    store.log_read_hook = [this, es](long long trial, size_t offset, size_t limit) {
      std::string err;
      auto rows = es->read(trial, offset, limit, &err);
      if (!err.empty()) log("error", "trial logs: " + err);
      return rows;
    };
    log("info", "trial logs go to elasticsearch at " + cfg.logging["host"].as_string() + ":" + std::to_string(cfg.logging["port"].as_int(9200)));
  }
  register_routes();
  if (!cfg.tls_cert.empty()) {
    std::string err;
    if (!http.enable_tls(cfg.tls_cert, cfg.tls_key, &err)) throw std::runtime_error("security.tls: " + err);
  }
  http.set_cors(cfg.enable_cors);
  const int port = http.listen(cfg.host, cfg.port);
  if (port >= 0 && cfg.provisioner.is_object()) provisioner = system.actor_of("provisioner", make_provisioner(this, cfg.provisioner));
  if (port < 0) throw std::runtime_error("cannot bind " + cfg.host + ":" + std::to_string(cfg.port));
  log("info", std::string("master listening on ") + (http.tls_enabled() ? "https://" : "http://") + cfg.host + ":" + std::to_string(port) + " with " + std::to_string(devices.size()) + " local slots (" + cfg.scheduler + ")");
  if (kube) kube->start();
  report_telemetry("master_started", obj({{"scheduler", cfg.scheduler}, {"resource_manager", kube ? "kubernetes" : "agent"},
                                          {"local_slots", static_cast<long long>(devices.size())}, {"provisioner", cfg.provisioner.is_object()}}));
  restore_experiments();
  return port;
}
void Master::run_forever() { http.serve_forever(); }
void Master::stop() {
  http.stop();
  if (kube) kube->stop();
  if (containers) containers->kill_all();
  system.shutdown();
}

}  // namespace det
