// det-agent — node agent: detects devices, registers with the master, reports slots; prints the detection result
// with --detect-only. On the single-box deployment the master embeds the same machinery as "agent-local"; det-agent
// serves extra nodes: it long-polls the master for container commands (start / signal), runs them under its own
// ContainerManager with the slot's CUDA_VISIBLE_DEVICES mask, and posts container state changes and log lines back.
//   det-agent [--master-host H] [--master-port N] [--agent-id ID] [--label L] [--slot-type auto|gpu|cpu|none]
//             [--visible-gpus LIST] [--artificial-slots N] [--detect-only] [--container-master-host H]
//             [--security-tls-enabled] [--security-tls-skip-verify] [--security-tls-master-cert FILE]
//             [--security-tls-master-cert-name NAME] [--config-file agent.yaml-as-json]
// Every flag can also come from DET_AGENT_<FLAG> (upper case, dashes as underscores), flags win.
// (reference: agent/cmd/determined-agent/run.go, agent/internal/options.go:12-103, agent.go:60-150)
#include <atomic>
#include <unistd.h>

#include <csignal>
#include <fstream>
#include <iostream>
#include <set>
#include <sstream>
#include <mutex>
#include <thread>

#include "agent.hpp"
#include "http.hpp"

static std::atomic<bool> g_stop{false};     // lock-free: safe to set from the signal handler
static void on_signal(int) { g_stop = true; }

int main(int argc, char** argv) {
  using namespace det;
  std::string host = "127.0.0.1", id, label, slot_type = "auto", visible, container_master_host, warm_mode = "gpu", warm_python, warm_root, pycache_dir = std::string(getenv("TMPDIR") ? getenv("TMPDIR") : "/tmp") + "/det-agent-pycache";
  int port = 8080, artificial = 0;
  bool detect_only = false;
  TlsClientOptions tls;
  std::map<std::string, std::string> task_env;
  // --config-file F (or DET_AGENT_CONFIG_FILE): the agent.yaml of the reference as JSON — master_host, master_port, agent_id,
  // label / resource_pool, slot_type, visible_gpus, artificial_slots, container_master_host, security.tls.{enabled, skip_verify,
  // master_cert, master_cert_name}, http_proxy / https_proxy / ftp_proxy / no_proxy. Environment variables and flags override it.
  std::string config_file = getenv("DET_AGENT_CONFIG_FILE") ? getenv("DET_AGENT_CONFIG_FILE") : "";
  for (int i = 1; i + 1 < argc; ++i) if (std::string(argv[i]) == "--config-file") config_file = argv[i + 1];
  std::string file_id;
  if (!config_file.empty()) {
    std::ifstream f(config_file);
    if (!f) { std::cerr << "det-agent: cannot read " << config_file << "\n"; return 2; }
    std::stringstream ss;
    ss << f.rdbuf();
    Json j;
    try { j = Json::parse(ss.str()); } catch (const std::exception& e) { std::cerr << "det-agent: " << config_file << ": " << e.what() << "\n"; return 2; }
    if (!j.is_object()) { std::cerr << "det-agent: " << config_file << ": a configuration is a mapping\n"; return 2; }
    static const std::set<std::string> known = {"master_host", "master_port", "agent_id", "label", "resource_pool", "slot_type", "visible_gpus",
                                                "artificial_slots", "container_master_host", "container_master_port", "security", "http_proxy",
                                                "https_proxy", "ftp_proxy", "no_proxy", "warm_workers", "config_file", "fluent", "api_enabled",
                                                "bind_ip", "bind_port"};
    for (auto& kv : j.obj())
      if (!known.count(kv.first)) { std::cerr << "det-agent: unknown field in agent config: '" << kv.first << "'\n"; return 2; }
    const Json& c = j;
    host = c["master_host"].str_or(host);
    port = static_cast<int>(c["master_port"].as_int(port));
    file_id = c["agent_id"].str_or("");
    label = c["label"].str_or(c["resource_pool"].str_or(label));
    slot_type = c["slot_type"].str_or(slot_type);
    visible = c["visible_gpus"].str_or(visible);
    artificial = static_cast<int>(c["artificial_slots"].as_int(artificial));
    container_master_host = c["container_master_host"].str_or(container_master_host);
    warm_mode = c["warm_workers"].str_or(warm_mode);
    const Json& t = c["security"]["tls"];
    if (t.is_object()) {
      tls.enabled = t["enabled"].as_bool(tls.enabled);
      tls.skip_verify = t["skip_verify"].as_bool(tls.skip_verify);
      tls.ca_file = t["master_cert"].str_or(tls.ca_file);
      tls.cert_name = t["master_cert_name"].str_or(tls.cert_name);
    }
    for (const char* k : {"http_proxy", "https_proxy", "ftp_proxy", "no_proxy"})
      if (!c[k].str_or("").empty()) {
        std::string upper = k;
        for (auto& ch : upper) ch = static_cast<char>(toupper(ch));
        task_env[k] = c[k].as_string();
        task_env[upper] = c[k].as_string();
      }
  }
  auto envs = [](const char* k, std::string& dst) { if (const char* v = getenv(k)) dst = v; };
  auto envb = [](const char* k, bool& dst) { if (const char* v = getenv(k)) dst = std::string(v) == "1" || std::string(v) == "true"; };
  envs("DET_AGENT_MASTER_HOST", host); envs("DET_AGENT_LABEL", label); envs("DET_AGENT_SLOT_TYPE", slot_type);
  envs("DET_AGENT_VISIBLE_GPUS", visible); envs("DET_AGENT_CONTAINER_MASTER_HOST", container_master_host);
  envb("DET_AGENT_SECURITY_TLS_ENABLED", tls.enabled); envb("DET_AGENT_SECURITY_TLS_SKIP_VERIFY", tls.skip_verify);
  envs("DET_AGENT_SECURITY_TLS_MASTER_CERT", tls.ca_file); envs("DET_AGENT_SECURITY_TLS_MASTER_CERT_NAME", tls.cert_name);
  if (const char* v = getenv("DET_AGENT_MASTER_PORT")) port = atoi(v);
  char hostname[256] = "agent";
  gethostname(hostname, sizeof hostname);
  id = file_id.empty() ? std::string("agent-") + hostname : file_id;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
    if (a == "--config-file") next();
    else if (a == "--master-host") host = next();
    else if (a == "--master-port") port = atoi(next().c_str());
    else if (a == "--agent-id") id = next();
    else if (a == "--label") label = next();
    else if (a == "--slot-type") slot_type = next();
    else if (a == "--visible-gpus") visible = next();
    else if (a == "--artificial-slots") artificial = atoi(next().c_str());
    else if (a == "--detect-only") detect_only = true;
    else if (a == "--container-master-host") container_master_host = next();
    else if (a == "--security-tls-enabled") tls.enabled = true;
    else if (a == "--security-tls-skip-verify") tls.skip_verify = true;
    else if (a == "--security-tls-master-cert") tls.ca_file = next();
    else if (a == "--security-tls-master-cert-name") tls.cert_name = next();
    else if (a == "--repo-root") warm_root = next();
    else if (a == "--python") warm_python = next();
    else if (a == "--warm-workers") warm_mode = next();   // off | gpu | lazy | eager (agent.hpp)
    else if (a == "--pycache-dir") pycache_dir = next();   // byte-code cache of the task processes ("" = leave Python's defaults)
    else { std::cerr << "unknown flag " << a << "\n"; return 2; }
  }
  auto devices = detect_devices(slot_type, visible, artificial);
  Json devs = Json::array();
  for (auto& d : devices) devs.push_back(d.to_json());
  Json body = Json::object();
  body["label"] = label;
  body["devices"] = devs;
  body["driver_version"] = nvidia_driver_version();
  Json addrs = Json::array();
  addrs.push_back(std::string(hostname));
  body["addresses"] = addrs;
  if (detect_only) { std::cout << body.dump(1) << "\n"; return 0; }
  signal(SIGINT, on_signal);
  signal(SIGTERM, on_signal);
  std::string out;
  while (!g_stop) {
    const int st = http_request(host, port, "POST", "/api/v1/agents/" + id + "/register", body.dump(), &out, 10, "", tls);
    if (st == 200) break;
    std::cerr << "det-agent: master not reachable at " << host << ":" << port << " (status " << st << (st < 0 && !out.empty() ? ": " + out : std::string()) << "); retrying\n";
    sleep(2);
  }
  std::cerr << "det-agent " << id << ": registered " << devices.size() << " slots with " << host << ":" << port << "\n";

  // outbox of container events and log lines, flushed to the master in batches
  std::mutex out_mu;
  Json pending_events = Json::array(), pending_logs = Json::array();
  ContainerManager cm(
      [&](const ContainerEvent& ev) {
        std::lock_guard<std::mutex> g(out_mu);
        Json e = Json::object();
        e["container_id"] = ev.container_id; e["state"] = ev.state; e["exit_code"] = static_cast<long long>(ev.exit_code); e["failure"] = ev.failure;
        pending_events.push_back(e);
      },
      [&](const LogLine& l) {
        std::lock_guard<std::mutex> g(out_mu);
        Json e = Json::object();
        e["container_id"] = l.container_id; e["stdtype"] = l.stdtype; e["message"] = l.message;
        e["rank"] = l.rank >= 0 ? Json(static_cast<long long>(l.rank)) : Json();
        pending_logs.push_back(e);
      });
  cm.set_pycache_dir(pycache_dir);
  cm.set_task_env(task_env);
  cm.set_slot_count(static_cast<int>(devices.size()));
  cm.enable_warm(warm_mode, warm_python, warm_root);
  if (warm_mode == "eager" && !warm_python.empty())
    for (auto& d : devices) cm.prewarm(d.type == "gpu" && d.brand != "Artificial" ? std::to_string(d.id) : std::string("cpu"));
  std::thread flusher([&] {
    while (!g_stop) {
      usleep(100 * 1000);
      Json b = Json::object();
      {
        std::lock_guard<std::mutex> g(out_mu);
        if (pending_events.arr().empty() && pending_logs.arr().empty()) continue;
        b["events"] = pending_events; b["logs"] = pending_logs;
        pending_events = Json::array(); pending_logs = Json::array();
      }
      std::string o;
      for (int attempt = 0; attempt < 20 && !g_stop; ++attempt) {
        if (http_request(host, port, "POST", "/api/v1/agents/" + id + "/events", b.dump(), &o, 10, "", tls) == 200) break;
        sleep(1);
      }
    }
  });
  time_t last_register = time(nullptr);
  while (!g_stop) {
    std::string resp;
    const int st = http_request(host, port, "GET", "/api/v1/agents/" + id + "/commands?wait_ms=5000", "", &resp, 15, "", tls);
    if (st != 200) { sleep(2); }
    else {
      Json cmds = Json::parse(resp)["commands"];
      for (auto& c : cmds.arr()) {
        if (c["type"].as_string() == "start") {
          ContainerSpec spec;
          spec.container_id = c["container_id"].as_string(); spec.task_id = c["task_id"].as_string();
          spec.description = c["description"].as_string(); spec.cwd = c["cwd"].as_string();
          for (auto& a : c["argv"].arr()) spec.argv.push_back(a.as_string());
          for (auto& kv : c["env"].obj()) spec.env[kv.first] = kv.second.as_string();
          // tasks on this node reach the master the way this agent does: its address (or the configured override) and the
          // certificate this agent pins (reference agent/internal/options.go: container_master_host, security.tls.master_cert)
          if (!container_master_host.empty()) { spec.env["DET_MASTER_ADDR"] = container_master_host; spec.env["DET_MASTER_HOST"] = container_master_host; }
          if (tls.enabled) {
            spec.env["DET_USE_TLS"] = "true";
            if (tls.skip_verify) spec.env["DET_MASTER_CERT_FILE"] = "noverify";
            else if (!tls.ca_file.empty()) spec.env["DET_MASTER_CERT_FILE"] = tls.ca_file;
            if (!tls.cert_name.empty()) spec.env["DET_MASTER_CERT_NAME"] = tls.cert_name;
          }
          for (auto& d : c["devices"].arr())
            spec.devices.push_back(Device{static_cast<int>(d["id"].as_int()), d["brand"].as_string(), d["uuid"].as_string(), d["type"].as_string()});
          cm.start(spec);
        } else if (c["type"].as_string() == "signal") {
          cm.signal(c["container_id"].as_string(), static_cast<int>(c["signal"].as_int(SIGTERM)));
        }
      }
    }
    if (time(nullptr) - last_register >= 5) {  // heartbeat: re-register if the master restarted
      last_register = time(nullptr);
      http_request(host, port, "POST", "/api/v1/agents/" + id + "/register", body.dump(), &out, 10, "", tls);
    }
  }
  cm.kill_all();
  flusher.join();
  http_request(host, port, "POST", "/api/v1/agents/" + id + "/unregister", "{}", &out, 5, "", tls);
  return 0;
}
