// kubernetes.hpp — Kubernetes as the container back-end: every task container becomes a pod.
//
// Parity map (reference → here): master/internal/kubernetes/spec.go (pod spec generation: env, GPU limits, host
// networking, volumes) → make_pod_spec; pods.go + pod.go (pod lifecycle per task container, state fan-out) →
// KubernetesPods::reconcile; informer.go / event listener (watch) → a 1 s list-and-diff loop over the label selector;
// request_queue.go (serialised create/delete workers with retry) → the request queue + worker thread below;
// pod log streamer → incremental reads of /pods/<name>/log; resourcemanagers/kubernetes_resource_manager.go → nodes are
// registered with the ordinary resource pool as agents ("k8s-<node>", one slot per allocatable GPU), so the fair-share /
// priority / round-robin schedulers and the fitting rules are shared with the agent back-end.
//
// Transport is plain HTTP to the API server address in the config — `kubectl proxy` on the master's pod or host
// (127.0.0.1:8001) terminates TLS and authentication; an optional bearer token is forwarded as-is.
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <functional>
#include <map>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <vector>

#include "agent.hpp"
#include "http.hpp"
#include "json.hpp"

namespace det {

struct KubeConfig {
  std::string api_host = "127.0.0.1";
  int api_port = 8001;
  std::string token;                       // optional bearer token
  std::string ns = "default";
  std::string image = "determined-b200:latest";
  std::string master_host;                 // address pods use to reach the master (service name / IP); "" = keep
  std::string slot_resource = "nvidia.com/gpu";
  std::string slot_type = "gpu";           // "cpu": slots are counted in slot_resource but tasks run without CUDA
  int max_slots_per_pod = 8;
  std::string shared_fs_host_path;         // mounted at the same path inside the pod when set
  std::string service_account;
  Json pod_spec_overrides;                 // merged into spec (nodeSelector, tolerations, ...)
  int poll_ms = 1000;

  static KubeConfig from_json(const Json& j) {
    KubeConfig c;
    const std::string api = j["api_server"].str_or("127.0.0.1:8001");
    const auto colon = api.rfind(':');
    c.api_host = colon == std::string::npos ? api : api.substr(0, colon);
    c.api_port = colon == std::string::npos ? 8001 : atoi(api.substr(colon + 1).c_str());
    c.token = j["token"].str_or("");
    c.ns = j["namespace"].str_or(c.ns);
    c.image = j["image"].str_or(c.image);
    c.master_host = j["master_service_name"].str_or(j["master_host"].str_or(""));
    c.slot_resource = j["slot_resource"].str_or(c.slot_resource);
    c.slot_type = j["slot_type"].str_or(c.slot_type);
    c.max_slots_per_pod = static_cast<int>(j["max_slots_per_pod"].as_int(c.max_slots_per_pod));
    c.shared_fs_host_path = j["shared_fs_host_path"].str_or("");
    c.service_account = j["service_account"].str_or("");
    c.pod_spec_overrides = j["pod_spec"];
    c.poll_ms = static_cast<int>(j["poll_ms"].as_int(c.poll_ms));
    return c;
  }
};

inline std::string kube_pod_name(const std::string& container_id) {
  std::string n = "det-";
  for (char ch : container_id) {
    const char c = static_cast<char>(tolower(ch));
    n += (isalnum(static_cast<unsigned char>(c)) || c == '-') ? c : '-';   // RFC 1123 label
  }
  if (n.size() > 63) n = n.substr(0, 63);
  while (!n.empty() && n.back() == '-') n.pop_back();
  return n;
}

// The pod manifest of one task container (reference spec.go: configureEnvVars / configureResourcesRequirements /
// configureVolumes / createPodSpec).
inline Json make_pod_spec(const ContainerSpec& spec, const KubeConfig& kc, const std::string& cluster_id, const std::string& node,
                          bool host_network) {
  Json pod = Json::object();
  pod["apiVersion"] = "v1";
  pod["kind"] = "Pod";
  Json meta = Json::object();
  meta["name"] = kube_pod_name(spec.container_id);
  meta["namespace"] = kc.ns;
  Json labels = Json::object();
  labels["determined"] = cluster_id;
  labels["determined-task"] = kube_pod_name(spec.task_id).substr(4);
  meta["labels"] = labels;
  Json ann = Json::object();
  ann["determined.ai/container-id"] = spec.container_id;
  ann["determined.ai/description"] = spec.description;
  meta["annotations"] = ann;
  pod["metadata"] = meta;

  Json c = Json::object();
  c["name"] = "determined-container";
  c["image"] = kc.image;
  c["imagePullPolicy"] = "IfNotPresent";
  Json cmd = Json::array();
  for (auto& a : spec.argv) cmd.push_back(Json(a));
  c["command"] = cmd;
  if (!spec.cwd.empty()) c["workingDir"] = spec.cwd;
  Json env = Json::array();
  auto add_env = [&](const std::string& k, const std::string& v) {
    Json e = Json::object();
    e["name"] = k; e["value"] = v;
    env.push_back(e);
  };
  for (auto& kv : spec.env) {
    if (!kc.master_host.empty() && (kv.first == "DET_MASTER_ADDR" || kv.first == "DET_MASTER_HOST")) add_env(kv.first, kc.master_host);
    else add_env(kv.first, kv.second);
  }
  int gpus = 0;
  for (auto& d : spec.devices) gpus += d.type == "gpu" ? 1 : 0;
  {   // what the agent injects for process containers (reference agent/internal/containers.go:203-213); the device
      // plugin renumbers a pod's GPUs from 0
    std::string ids = "[";
    for (int i = 0; i < gpus; ++i) ids += (i ? "," : "") + std::to_string(i);
    add_env("DET_SLOT_IDS", ids + "]");
    const bool use_gpu = gpus > 0 && kc.slot_type != "cpu";
    add_env("DET_USE_GPU", use_gpu ? "true" : "false");
    if (!use_gpu) add_env("CUDA_VISIBLE_DEVICES", "");
  }
  add_env("DET_CONTAINER_ID", spec.container_id);
  add_env("DET_K8S_POD", "1");
  c["env"] = env;
  Json res = Json::object();
  if (gpus > 0) {
    Json lim = Json::object();
    lim[kc.slot_resource] = static_cast<long long>(gpus);
    res["limits"] = lim;
  }
  c["resources"] = res;
  Json mounts = Json::array(), volumes = Json::array();
  auto add_host_path = [&](const std::string& name, const std::string& path) {
    Json m = Json::object();
    m["name"] = name; m["mountPath"] = path;
    mounts.push_back(m);
    Json v = Json::object(), hp = Json::object();
    hp["path"] = path; hp["type"] = "DirectoryOrCreate";
    v["name"] = name; v["hostPath"] = hp;
    volumes.push_back(v);
  };
  if (!kc.shared_fs_host_path.empty()) add_host_path("det-shared-fs", kc.shared_fs_host_path);
  {   // /dev/shm sized for NCCL / DataLoader workers (reference: shm_size in the task container defaults)
    Json m = Json::object();
    m["name"] = "det-shm"; m["mountPath"] = "/dev/shm";
    mounts.push_back(m);
    Json v = Json::object(), ed = Json::object();
    ed["medium"] = "Memory";
    v["name"] = "det-shm"; v["emptyDir"] = ed;
    volumes.push_back(v);
  }
  c["volumeMounts"] = mounts;

  Json ps = Json::object();
  ps["restartPolicy"] = "Never";
  ps["containers"] = Json(Json::array_t{c});
  ps["volumes"] = volumes;
  if (!node.empty()) ps["nodeName"] = node;
  if (host_network) { ps["hostNetwork"] = true; ps["dnsPolicy"] = "ClusterFirstWithHostNet"; }
  if (!kc.service_account.empty()) ps["serviceAccountName"] = kc.service_account;
  if (kc.pod_spec_overrides.is_object())
    for (auto& kv : kc.pod_spec_overrides.obj()) ps[kv.first] = kv.second;
  pod["spec"] = ps;
  return pod;
}

// Exit status of a finished pod: the first terminated container state wins; a pod that failed without one reports 137.
inline int kube_pod_exit_code(const Json& pod, std::string* reason) {
  const Json& st = pod["status"];
  if (st["containerStatuses"].is_array())
    for (auto& cs : st["containerStatuses"].arr()) {
      const Json& t = cs["state"]["terminated"];
      if (t.is_object()) {
        if (reason) *reason = t["reason"].str_or("");
        return static_cast<int>(t["exitCode"].as_int(0));
      }
    }
  if (reason) *reason = st["reason"].str_or(st["phase"].str_or(""));
  return st["phase"].str_or("") == "Succeeded" ? 0 : 137;
}

class KubernetesPods {
 public:
  using EventFn = std::function<void(const ContainerEvent&)>;
  using LogFn = std::function<void(const LogLine&)>;
  using NodeFn = std::function<void(const std::string& node, int slots, bool present)>;

  KubernetesPods(KubeConfig kc, std::string cluster_id, EventFn on_event, LogFn on_log, NodeFn on_node)
      : kc_(std::move(kc)), cluster_id_(std::move(cluster_id)), on_event_(std::move(on_event)), on_log_(std::move(on_log)),
        on_node_(std::move(on_node)) {}
  ~KubernetesPods() { stop(); }

  void start() {
    running_ = true;
    worker_ = std::thread([this] { request_loop(); });
    informer_ = std::thread([this] { informer_loop(); });
  }
  void stop() {
    if (!running_.exchange(false)) return;
    cv_.notify_all();
    if (worker_.joinable()) worker_.join();
    if (informer_.joinable()) informer_.join();
  }

  static bool owns(const std::string& agent_id) { return agent_id.rfind("k8s-", 0) == 0; }
  bool has_container(const std::string& container_id) {
    std::lock_guard<std::mutex> g(mu_);
    return pods_.count(container_id) > 0;
  }

  void create(const std::string& agent_id, const ContainerSpec& spec, bool host_network) {
    const std::string node = agent_id.substr(4);
    Request r;
    r.kind = Request::Create;
    r.container_id = spec.container_id;
    r.body = make_pod_spec(spec, kc_, cluster_id_, node, host_network).dump();
    {
      std::lock_guard<std::mutex> g(mu_);
      PodState& p = pods_[spec.container_id];
      p.name = kube_pod_name(spec.container_id);
      p.state = "ASSIGNED";
      queue_.push_back(std::move(r));
    }
    on_event_(ContainerEvent{spec.container_id, "ASSIGNED", 0, ""});
    cv_.notify_all();
  }
  void remove(const std::string& container_id, bool force) {
    std::lock_guard<std::mutex> g(mu_);
    if (!pods_.count(container_id)) return;
    Request r;
    r.kind = Request::Delete;
    r.container_id = container_id;
    r.grace = force ? 0 : 30;
    queue_.push_back(std::move(r));
    cv_.notify_all();
  }

  Json summary() {
    std::lock_guard<std::mutex> g(mu_);
    Json out = Json::array();
    for (auto& kv : pods_) {
      Json p = Json::object();
      p["container_id"] = kv.first; p["pod"] = kv.second.name; p["state"] = kv.second.state;
      out.push_back(p);
    }
    return out;
  }

 private:
  struct Request {
    enum Kind { Create, Delete } kind = Create;
    std::string container_id, body;
    int grace = 30;
    int attempts = 0;
  };
  struct PodState {
    std::string name, state;   // ASSIGNED | PULLING | STARTING | RUNNING | TERMINATED
    size_t log_bytes = 0;
    bool delete_sent = false;
  };

  int api(const std::string& method, const std::string& path, const std::string& body, std::string* out) {
    return http_request(kc_.api_host, kc_.api_port, method, path, body, out, 30, kc_.token);
  }
  std::string pods_path() const { return "/api/v1/namespaces/" + kc_.ns + "/pods"; }

  // Serialised create/delete with retry (reference request_queue.go: failed creations are retried, deletions of pods that
  // are already gone count as done).
  void request_loop() {
    while (running_) {
      Request r;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait_for(lk, std::chrono::milliseconds(200), [&] { return !queue_.empty() || !running_; });
        if (!running_) return;
        if (queue_.empty()) continue;
        r = std::move(queue_.front());
        queue_.pop_front();
      }
      std::string resp;
      if (r.kind == Request::Create) {
        const int st = api("POST", pods_path(), r.body, &resp);
        if (st == 409) continue;   // already exists (a retry raced with a slow success)
        if (st < 200 || st >= 300) {
          if (++r.attempts < 5 && (st < 0 || st >= 500 || st == 429)) {
            std::this_thread::sleep_for(std::chrono::milliseconds(200 * r.attempts));
            std::lock_guard<std::mutex> g(mu_);
            queue_.push_back(std::move(r));
          } else {
            finish(r.container_id, 125, "pod creation failed: HTTP " + std::to_string(st) + " " + resp.substr(0, 200));
          }
        }
      } else {
        std::string name;
        {
          std::lock_guard<std::mutex> g(mu_);
          auto it = pods_.find(r.container_id);
          if (it == pods_.end()) continue;
          name = it->second.name;
          it->second.delete_sent = true;
        }
        const int st = api("DELETE", pods_path() + "/" + name + "?gracePeriodSeconds=" + std::to_string(r.grace), "", &resp);
        if (st == 404) finish(r.container_id, 137, "killed");
        else if ((st < 200 || st >= 300) && ++r.attempts < 5) {
          std::this_thread::sleep_for(std::chrono::milliseconds(200 * r.attempts));
          std::lock_guard<std::mutex> g(mu_);
          queue_.push_back(std::move(r));
        }
      }
    }
  }

  void finish(const std::string& container_id, int code, const std::string& failure) {
    {
      std::lock_guard<std::mutex> g(mu_);
      auto it = pods_.find(container_id);
      if (it == pods_.end() || it->second.state == "TERMINATED") return;
      it->second.state = "TERMINATED";
      pods_.erase(it);
    }
    on_event_(ContainerEvent{container_id, "TERMINATED", code, failure});
  }

  void informer_loop() {
    int node_countdown = 0;
    while (running_) {
      if (node_countdown-- <= 0) { refresh_nodes(); node_countdown = 10; }
      reconcile();
      for (int i = 0; i < kc_.poll_ms / 50 && running_; ++i) std::this_thread::sleep_for(std::chrono::milliseconds(50));
    }
  }

  // Nodes with allocatable slot resources are the "agents" of this back-end.
  void refresh_nodes() {
    std::string resp;
    if (api("GET", "/api/v1/nodes", "", &resp) != 200) return;
    Json list;
    try { list = Json::parse(resp); } catch (...) { return; }
    std::set<std::string> seen;
    if (list["items"].is_array())
      for (auto& n : list["items"].arr()) {
        const std::string name = n["metadata"]["name"].str_or("");
        if (name.empty() || n["spec"]["unschedulable"].as_bool(false)) continue;
        const Json& q = n["status"]["allocatable"][kc_.slot_resource];
        const int slots = q.is_string() ? atoi(q.as_string().c_str()) : static_cast<int>(q.as_int(0));
        if (slots <= 0) continue;
        seen.insert(name);
        if (!nodes_.count(name) || nodes_[name] != slots) { nodes_[name] = slots; on_node_(name, slots, true); }
      }
    for (auto it = nodes_.begin(); it != nodes_.end();)
      if (!seen.count(it->first)) { on_node_(it->first, 0, false); it = nodes_.erase(it); } else ++it;
  }

  void reconcile() {
    std::string resp;
    if (api("GET", pods_path() + "?labelSelector=determined%3D" + cluster_id_, "", &resp) != 200) return;
    Json list;
    try { list = Json::parse(resp); } catch (...) { return; }
    std::map<std::string, Json> live;
    if (list["items"].is_array())
      for (auto& p : list["items"].arr()) live[p["metadata"]["annotations"]["determined.ai/container-id"].str_or("")] = p;
    std::vector<std::pair<std::string, std::string>> transitions;
    std::vector<std::pair<std::string, Json>> finished;
    std::vector<std::pair<std::string, std::string>> to_stream;
    {
      std::lock_guard<std::mutex> g(mu_);
      for (auto& kv : pods_) {
        auto it = live.find(kv.first);
        if (it == live.end()) {
          if (kv.second.delete_sent) finished.push_back({kv.first, Json()});   // deleted and gone
          continue;
        }
        const Json& st = it->second["status"];
        const std::string phase = st["phase"].str_or("Pending");
        std::string next = kv.second.state;
        if (phase == "Pending") {
          bool pulling = false;
          if (st["containerStatuses"].is_array())
            for (auto& cs : st["containerStatuses"].arr()) {
              const std::string why = cs["state"]["waiting"]["reason"].str_or("");
              pulling |= why == "ContainerCreating" || why == "Pulling" || why == "ErrImagePull" || why == "ImagePullBackOff";
            }
          next = pulling ? "PULLING" : (st["conditions"].is_array() ? "STARTING" : "ASSIGNED");
        } else if (phase == "Running") {
          next = "RUNNING";
        } else if (phase == "Succeeded" || phase == "Failed") {
          finished.push_back({kv.first, it->second});
          to_stream.push_back({kv.first, kv.second.name});
          continue;
        }
        if (next != kv.second.state) { kv.second.state = next; transitions.push_back({kv.first, next}); }
        if (next == "RUNNING") to_stream.push_back({kv.first, kv.second.name});
      }
    }
    for (auto& t : transitions) on_event_(ContainerEvent{t.first, t.second, 0, ""});
    for (auto& s : to_stream) stream_logs(s.first, s.second);
    for (auto& f : finished) {
      std::string reason;
      const int code = f.second.is_object() ? kube_pod_exit_code(f.second, &reason) : 137;
      std::string name;
      {
        std::lock_guard<std::mutex> g(mu_);
        auto it = pods_.find(f.first);
        if (it != pods_.end()) name = it->second.name;
      }
      finish(f.first, code, code == 0 ? "" : (reason.empty() ? "pod failed" : reason));
      if (!name.empty() && f.second.is_object()) {   // finished pods are garbage-collected once their exit is recorded
        std::string ignore;
        api("DELETE", pods_path() + "/" + name + "?gracePeriodSeconds=0", "", &ignore);
      }
    }
  }

  // The API returns the whole log; forward what is new since the last read, line by line, parsing the harness's
  // "[rank=N] " prefix the way the agent's log pump does.
  void stream_logs(const std::string& container_id, const std::string& name) {
    std::string body;
    if (api("GET", pods_path() + "/" + name + "/log", "", &body) != 200) return;
    size_t from;
    {
      std::lock_guard<std::mutex> g(mu_);
      auto it = pods_.find(container_id);
      if (it == pods_.end()) return;
      from = it->second.log_bytes;
      if (body.size() <= from) return;
      const size_t last_nl = body.rfind('\n');
      if (last_nl == std::string::npos || last_nl < from) return;   // only complete lines
      it->second.log_bytes = last_nl + 1;
      body = body.substr(from, last_nl + 1 - from);
    }
    size_t pos = 0;
    while (pos < body.size()) {
      const size_t nl = body.find('\n', pos);
      std::string line = body.substr(pos, nl - pos);
      pos = nl + 1;
      LogLine l;
      l.container_id = container_id;
      l.stdtype = "stdout";
      if (line.rfind("[rank=", 0) == 0) {
        const size_t close = line.find(']');
        if (close != std::string::npos) {
          l.rank = atoi(line.substr(6, close - 6).c_str());
          line = line.substr(close + 1 < line.size() && line[close + 1] == ' ' ? close + 2 : close + 1);
        }
      }
      l.message = line;
      on_log_(l);
    }
  }

  KubeConfig kc_;
  std::string cluster_id_;
  EventFn on_event_;
  LogFn on_log_;
  NodeFn on_node_;
  std::atomic<bool> running_{false};
  std::mutex mu_;
  std::condition_variable cv_;
  std::deque<Request> queue_;
  std::map<std::string, PodState> pods_;
  std::map<std::string, int> nodes_;
  std::thread worker_, informer_;
};

}  // namespace det
