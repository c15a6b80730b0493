// master.hpp — the single-box master: configuration, shared services, actor declarations.
// Parity map (reference → here): master/internal/core.go Master.Run → Master::run; config.go → MasterConfig;
// resourcemanagers/resource_pool.go → ResourcePoolActor; experiment.go → ExperimentActor; trial.go → TrialActor;
// trial_logger.go → TrialLoggerActor; command/*.go → CommandActor; checkpoint_gc.go → Master::spawn_checkpoint_gc;
// agent/internal/agent.go → the embedded local agent (ContainerManager) plus det-agent for extra nodes.
#pragma once
#include <atomic>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <random>
#include <string>
#include <vector>

#include "actor.hpp"
#include "agent.hpp"
#include "http.hpp"
#include "json.hpp"
#include "kubernetes.hpp"
#include "model.hpp"
#include "provisioner.hpp"
#include "rw_coordinator.hpp"
#include "scheduler.hpp"
#include "searcher.hpp"
#include "sequencer.hpp"
#include "store.hpp"

namespace det {

struct MasterConfig {
  std::string host = "127.0.0.1";
  int port = 8080;
  std::string data_dir = "/tmp/determined-master";
  std::string scheduler = "fair_share";  // fair_share | priority | round_robin
  std::string fitting_policy = "best";
  bool preemption = false;
  Json checkpoint_storage;               // master-level default
  std::string python = "python";
  std::string repo_root;                 // added to PYTHONPATH of launched harness processes
  std::string warm_workers = "gpu";      // off | gpu | lazy | eager: pre-started harness processes per slot (agent.hpp)
  std::string slot_type = "auto";
  std::string visible_gpus;
  int artificial_slots = 0;
  bool auth_required = false;
  std::string tls_cert, tls_key;         // security.tls.{cert,key}: serve HTTPS (reference master/internal/config.go:252-256)
  std::string tls_cert_name;             // security.tls.master_cert_name: the name tasks and agents expect in the certificate
  std::string cluster_name = "b200-box";
  Json provisioner;                      // {max_instances, slots_per_instance, max_idle_agent_period, ...} or null
  Json resource_manager;                 // null / {type: agent} or {type: kubernetes, namespace, api_server, image, ...}
  Json logging;                          // {type: default | elastic, host, port, index}: where trial logs are kept
  Json telemetry;                        // {enabled: bool, sink: path}; events go to a local JSONL file (no network here)
  int scheduler_tick_ms = 500;
  int terminate_timeout_ms = 60000;
  int all_ready_timeout_ms = 600000;     // reference allReadyTimeout: 10 minutes from container start to a working harness
  Json task_container_defaults;          // {shm_size_bytes, dtrain_network_interface, nccl_port_range, gloo_port_range, ...}
  Json default_task;                     // security.default_task: {uid, gid, user, group} of tasks whose owner has no linked agent user
  int tensorboard_timeout_s = 300;       // idle TensorBoards are stopped after this long without a request (reference tensorboard_timeout)
  bool enable_cors = false;              // Access-Control-Allow-* on every response, OPTIONS pre-flights answered
  std::string log_level = "info";        // log.level: trace | debug | info | warn | error | fatal

  static MasterConfig from_json(const Json& j);  // strict: unknown keys are rejected
  Json to_json() const;
};

std::string now_rfc3339();
std::string random_hex(int bytes);
std::string md5_hex(const std::string& s);

class Master {
 public:
  explicit Master(const MasterConfig& cfg);
  ~Master();
  int start();         // binds the API port, starts actors, restores experiments; returns the bound port
  void run_forever();  // blocks in the HTTP accept loop
  void stop();

  MasterConfig cfg;
  Store store;
  ActorSystem system;
  HttpServer http;
  std::unique_ptr<ContainerManager> containers;
  std::unique_ptr<KubernetesPods> kube;  // set when resource_manager.type == kubernetes
  std::vector<Device> devices;
  ActorRef rm, logger, experiments_root, commands_root, provisioner;
  RWCoordinator rw;
  std::mutex agent_cmd_mu_;
  std::condition_variable agent_cmd_cv_;
  std::map<std::string, std::vector<Json>> agent_cmds_;      // agent id -> pending commands
  std::map<std::string, std::string> container_agent_;       // container id -> agent id (remote ones only)   // data-layer reader/writer leases (reference: master/internal/rw_coordinator.go)
  std::string cluster_id;
  std::atomic<long long> started_at{0};

  // log ring served at /logs (reference: master/pkg/logger LogBuffer)
  void log(const std::string& level, const std::string& msg);
  // usage telemetry (reference: master/internal/telemetry — Segment track calls): one JSON line per event, opt-in
  void report_telemetry(const std::string& event, const Json& properties);
  std::vector<Json> recent_logs(size_t limit, long long after_id = -1);   // entries carry a running "id"; after_id: only newer ones

  // services used by actors
  long long create_experiment(const Json& resolved_config, const Json& model_definition, long long parent_id, bool activate,
                              const std::string& owner);
  ActorRef experiment_actor(long long id);
  void route_container_event(const ContainerEvent& ev);
  // Container placement: "agent-local" runs under the master's own ContainerManager; any other agent receives the
  // spec through its command queue (long-polled by det-agent) and reports events / logs back over REST.
  void start_container(const std::string& agent_id, const ContainerSpec& spec);
  void signal_container(const std::string& container_id, int sig);
  Json take_agent_commands(const std::string& agent_id, int wait_ms);
  void route_log(const LogLine& l);
  void register_container(const std::string& container_id, const ActorRef& owner, long long trial_id);
  void unregister_container(const std::string& container_id);
  void spawn_checkpoint_gc(long long experiment_id);
  void write_context_files(const std::string& dir, const Json& files);
  void apply_owner_identity(const std::string& owner, std::map<std::string, std::string>& env);
  std::string command_context_dir(long long command_id) const { return cfg.data_dir + "/commands/" + std::to_string(command_id) + "/context"; }
  std::string model_def_dir(long long experiment_id) const { return cfg.data_dir + "/experiments/" + std::to_string(experiment_id) + "/model_def"; }
  bool check_auth(const HttpRequest& req, std::string* user);

 private:
  void register_routes();
  void restore_experiments();
  std::mutex log_mu_;
  std::deque<Json> log_ring_;
  long long log_seq_ = 0;
  std::mutex container_mu_;
  std::map<std::string, std::pair<ActorRef, long long>> container_owner_;
};

std::unique_ptr<Actor> make_resource_pool(Master* m);
std::unique_ptr<Actor> make_trial_logger(Master* m);
std::unique_ptr<Actor> make_experiment(Master* m, long long id, bool restoring);
std::unique_ptr<Actor> make_command(Master* m, long long id, const Json& spec);
std::unique_ptr<Actor> make_root();
std::unique_ptr<Actor> make_provisioner(Master* m, const Json& cfg);

}  // namespace det
