// det-master — single-box master daemon.
//   det-master [--config master.json] [--port N] [--host H] [--data-dir D] [--scheduler S] [--artificial-slots N]
//              [--repo-root PATH] [--python EXE] [--slot-type auto|gpu|cpu|none] [--print-port-file F]
// Config precedence: defaults <- config file (JSON; unknown keys rejected) <- DET_MASTER_* environment <- flags
// (reference: master/cmd/determined-master/root.go:58-87, master/internal/config.go:27-148).
#include <csignal>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <sstream>

#include "master.hpp"

static det::Master* g_master = nullptr;
static void on_signal(int) {
  if (g_master) g_master->stop();
  _exit(0);
}

int main(int argc, char** argv) {
  using namespace det;
  Json j = Json::object();
  std::string port_file;
  try {
    for (int i = 1; i < argc; ++i) {
      if (std::string(argv[i]) == "--config" && i + 1 < argc) {
        std::ifstream f(argv[i + 1]);
        if (!f) { std::cerr << "cannot read " << argv[i + 1] << "\n"; return 2; }
        std::stringstream ss;
        ss << f.rdbuf();
        j = Json::parse(ss.str());
      }
    }
    auto env = [&](const char* name, const char* key, bool is_int) {
      if (const char* v = getenv(name)) j[key] = is_int ? Json(static_cast<long long>(atoll(v))) : Json(std::string(v));
    };
    env("DET_MASTER_PORT", "port", true); env("DET_MASTER_HOST", "host", false); env("DET_MASTER_DATA_DIR", "data_dir", false);
    env("DET_MASTER_SCHEDULER", "scheduler", false); env("DET_MASTER_ARTIFICIAL_SLOTS", "artificial_slots", true);
    for (int i = 1; i < argc; ++i) {
      const std::string a = argv[i];
      auto next = [&]() -> std::string { if (i + 1 >= argc) throw ConfigError("flag " + a + " needs a value"); return argv[++i]; };
      if (a == "--config") next();
      else if (a == "--port") j["port"] = static_cast<long long>(atoll(next().c_str()));
      else if (a == "--host") j["host"] = next();
      else if (a == "--data-dir") j["data_dir"] = next();
      else if (a == "--scheduler") j["scheduler"] = next();
      else if (a == "--fitting-policy") j["fitting_policy"] = next();
      else if (a == "--preemption") j["preemption"] = true;
      else if (a == "--artificial-slots") j["artificial_slots"] = static_cast<long long>(atoll(next().c_str()));
      else if (a == "--repo-root") j["repo_root"] = next();
      else if (a == "--python") j["python"] = next();
      else if (a == "--warm-workers") j["warm_workers"] = next();   // off | gpu | lazy | eager
      else if (a == "--slot-type") j["slot_type"] = next();
      else if (a == "--telemetry") j["telemetry"] = det::Json::parse(next());     // {"enabled": true, "sink": "/path/events.jsonl"}
      else if (a == "--provisioner") j["provisioner"] = det::Json::parse(next());   // JSON object, see docs/cluster-config.md
      else if (a == "--resource-manager") j["resource_manager"] = det::Json::parse(next());   // {"type": "kubernetes", ...}
      else if (a == "--visible-gpus") j["visible_gpus"] = next();
      else if (a == "--auth-required") j["auth_required"] = true;
      else if (a == "--logging") j["logging"] = det::Json::parse(next());   // {"type": "elastic", "host": "...", "port": 9200}
      else if (a == "--security-tls-cert" || a == "--security-tls-key" || a == "--security-tls-master-cert-name") {
        det::Json sec = j["security"].is_object() ? j["security"] : det::Json::object();
        det::Json t = sec["tls"].is_object() ? sec["tls"] : det::Json::object();
        t[a == "--security-tls-cert" ? "cert" : (a == "--security-tls-key" ? "key" : "master_cert_name")] = next();
        sec["tls"] = t;
        j["security"] = sec;
      }
      else if (a == "--print-port-file") port_file = next();
      else if (a == "--help" || a == "-h") { std::cout << "usage: det-master [--config F] [--port N] [--host H] [--data-dir D] [--scheduler fair_share|priority|round_robin] [--artificial-slots N] [--repo-root P] [--python EXE]\n"; return 0; }
      else throw ConfigError("unknown flag " + a);
    }
    MasterConfig cfg = MasterConfig::from_json(j);
    Master master(cfg);
    g_master = &master;
    signal(SIGINT, on_signal);
    signal(SIGTERM, on_signal);
    signal(SIGPIPE, SIG_IGN);
    const int port = master.start();
    if (!port_file.empty()) { std::ofstream pf(port_file); pf << port << "\n"; }
    master.run_forever();
  } catch (const std::exception& e) {
    std::cerr << "det-master: " << e.what() << "\n";
    return 1;
  }
  return 0;
}
