// det-agent — node agent: detects devices, registers with the master, reports slots; prints the detection result
// with --detect-only. On the single-box deployment the master embeds the same machinery as "agent-local"; det-agent
// serves extra nodes and device inspection.
//   det-agent [--master-host H] [--master-port N] [--agent-id ID] [--label L] [--slot-type auto|gpu|cpu|none]
//             [--visible-gpus LIST] [--artificial-slots N] [--detect-only]
// (reference: agent/cmd/determined-agent/run.go, agent/internal/options.go:12-103, agent.go:60-150)
#include <unistd.h>

#include <csignal>
#include <iostream>

#include "agent.hpp"
#include "http.hpp"

static volatile bool g_stop = false;
static void on_signal(int) { g_stop = true; }

int main(int argc, char** argv) {
  using namespace det;
  std::string host = "127.0.0.1", id, label, slot_type = "auto", visible;
  int port = 8080, artificial = 0;
  bool detect_only = false;
  char hostname[256] = "agent";
  gethostname(hostname, sizeof hostname);
  id = std::string("agent-") + hostname;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    auto next = [&]() -> std::string { return i + 1 < argc ? argv[++i] : ""; };
    if (a == "--master-host") host = next();
    else if (a == "--master-port") port = atoi(next().c_str());
    else if (a == "--agent-id") id = next();
    else if (a == "--label") label = next();
    else if (a == "--slot-type") slot_type = next();
    else if (a == "--visible-gpus") visible = next();
    else if (a == "--artificial-slots") artificial = atoi(next().c_str());
    else if (a == "--detect-only") detect_only = true;
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
    const int st = http_request(host, port, "POST", "/api/v1/agents/" + id + "/register", body.dump(), &out, 10);
    if (st == 200) break;
    std::cerr << "det-agent: master not reachable at " << host << ":" << port << " (status " << st << "); retrying\n";
    sleep(2);
  }
  std::cerr << "det-agent " << id << ": registered " << devices.size() << " slots with " << host << ":" << port << "\n";
  while (!g_stop) {  // heartbeat: re-register if the master restarted
    sleep(5);
    if (http_request(host, port, "GET", "/api/v1/master", "", &out, 5) != 200) continue;
    http_request(host, port, "POST", "/api/v1/agents/" + id + "/register", body.dump(), &out, 10);
  }
  http_request(host, port, "POST", "/api/v1/agents/" + id + "/unregister", "{}", &out, 5);
  return 0;
}
