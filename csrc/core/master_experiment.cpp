// master_experiment.cpp — experiment actor (owns the Searcher, spawns one trial actor per Create, routes operations,
// tracks the best validation, persists searcher events for replay-based restore, pause/activate/cancel/kill, GC on
// completion) and trial actor (slot request, harness launch, long-poll workload exchange, sequencer, graceful vs
// forced termination, restart-with-rollback up to max_restarts).
// Parity: reference master/internal/experiment.go:81-592 and trial.go:173-1161.
#include <algorithm>
#include <cmath>

#include "task_spec.hpp"
#include "master.hpp"

namespace det {

namespace {

Json obj(std::initializer_list<std::pair<const char*, Json>> kv) {
  Json j = Json::object();
  for (auto& p : kv) j[p.first] = p.second;
  return j;
}
const char* reason_name(ExitedReason r) { return r == ExitedReason::UserCanceled ? "USER_CANCELED" : r == ExitedReason::InvalidHP ? "INVALID_HP" : "ERRORED"; }
ExitedReason reason_from(const std::string& s) { return s == "USER_CANCELED" ? ExitedReason::UserCanceled : s == "INVALID_HP" ? ExitedReason::InvalidHP : ExitedReason::Errored; }

// =====================================================================================================================
class TrialActor : public Actor {
 public:
  TrialActor(Master* m, long long exp_id, const Json& config, const Operation& create, const Json& first_checkpoint, const std::string& exp_state,
             long long restored_trial_id)
      : m_(m), exp_id_(exp_id), config_(config), create_(create), exp_state_(exp_state),
        seq_(config, static_cast<int>(exp_id), create, first_checkpoint), trial_id_(restored_trial_id) {
    task_id_ = "trial-" + std::to_string(exp_id) + "-" + create.request_id.substr(0, 8);
    max_restarts_ = static_cast<int>(config["max_restarts"].as_int(5));
    if (trial_id_ > 0) seq_.set_trial_id(static_cast<int>(trial_id_));
  }

  // replay support: feed a stored op / completed workload without side effects
  void restore_op(const Operation& op) {
    if (op.kind == Operation::Close) close_requested_ = true; else seq_.operation_requested(op);
  }
  void restore_completed(const Json& j, bool is_best) {
    CompletedMessage msg = parse_completed(j);
    Operation op; Json metrics;
    try { seq_.workload_completed(msg, is_best, &op, &metrics); } catch (...) {}
  }
  void restore_finish() { seq_.roll_back(); }

  void receive(Context& ctx, const Message& msg) override {
    const std::string& t = msg.type;
    if (t == "operation") {
      Operation op = Operation::from_json(msg.body);
      m_->store.append("trial_ops", key(), msg.body);
      if (op.kind == Operation::Close) close_requested_ = true; else seq_.operation_requested(op);
      progress(ctx);
    } else if (t == "experiment_state") {
      exp_state_ = msg.body.as_string();
      progress(ctx);
    } else if (t == "resources_allocated") {
      allocation_requested_ = false;
      if (!wants_to_run()) { release(ctx); progress(ctx); return; }
      launch(ctx, msg.body["allocations"]);
    } else if (t == "release_resources") {  // preemption by the scheduler
      preempt_requested_ = true;
      progress(ctx);
    } else if (t == "next_workload") {
      // a harness left over from an earlier run of this trial (e.g. it outlived a master restart) is told to exit
      const std::string cid = msg.body["container_id"].str_or("");
      if (!cid.empty() && running_ && std::find(containers_.begin(), containers_.end(), cid) == containers_.end()) {
        msg.respond(obj({{"type", "TERMINATE"}}));
        return;
      }
      harness_ready_ = true;      // the trial's chief is up and asking for work
      if (poll_.reply) poll_.respond(obj({{"type", "WAIT"}}));
      poll_ = msg;
      ++poll_seq_;
      ctx.system().notify_after(ctx.self(), 20000, "poll_timeout", Json(static_cast<long long>(poll_seq_)));
      progress(ctx);
    } else if (t == "poll_timeout") {
      if (poll_.reply && msg.body.as_int() == static_cast<long long>(poll_seq_)) { poll_.respond(obj({{"type", "WAIT"}})); poll_ = Message(); }
    } else if (t == "workload_completed") {
      on_workload_completed(ctx, msg.body);
      msg.respond(Json(true));
    } else if (t == "container_event") {
      on_container_event(ctx, msg.body);
    } else if (t == "all_ready_timeout") {
      // reference trial.go:408-418 (allReadyTimeout): containers that never produce a working harness — a gang member
      // that hangs at rendezvous, a start-up hook that never returns — are killed; the exit takes the restart path
      if (running_ && !harness_ready_ && msg.body.as_int() == run_id_) {
        m_->log("warn", task_id_ + ": containers were not ready within " + std::to_string(m_->cfg.all_ready_timeout_ms / 1000) + " s; killing them");
        exit_failure_ = "trial containers did not become ready in time";
        for (auto& c : containers_) m_->signal_container(c, SIGKILL);
      }
    } else if (t == "terminate_timeout") {
      if (running_ && awaiting_exit_ && msg.body.as_int() == run_id_) {
        m_->log("warn", task_id_ + ": did not exit after TERMINATE; killing");
        for (auto& c : containers_) m_->signal_container(c, SIGKILL);
      }
    } else if (t == "kill") {
      killed_ = true;
      if (running_) for (auto& c : containers_) m_->signal_container(c, SIGKILL);
      else finish(ctx, "CANCELED");
      msg.respond(Json(true));
    } else if (t == "agent_lost") {
      if (running_) on_exit(ctx, 137, "agent lost");
    } else if (t == "describe") {
      Json d = seq_.snapshot_json();
      d["trial_id"] = trial_id_; d["running"] = running_; d["request_id"] = create_.request_id; d["restarts"] = restarts_;
      msg.respond(d);
    } else {
      msg.respond(Json());
    }
  }

  void post_stop(Context&) override {
    if (poll_.reply) poll_.respond(obj({{"type", "TERMINATE"}}));
    for (auto& c : containers_) { m_->signal_container(c, SIGKILL); m_->unregister_container(c); }
  }

 private:
  std::string key() const { return std::to_string(exp_id_) + ":" + create_.request_id; }
  bool wants_to_run() const { return exp_state_ == "ACTIVE" && !killed_ && !finished_ && !seq_.up_to_date(); }

  static CompletedMessage parse_completed(const Json& j) {
    CompletedMessage msg;
    msg.workload = Workload::from_json(j["workload"]);
    if (j["exited_reason"].is_string()) { msg.has_exited_reason = true; msg.exited_reason = reason_from(j["exited_reason"].as_string()); }
    msg.metrics = j["metrics"];
    msg.validation_metrics = j["metrics"];
    msg.checkpoint_metrics = j["metrics"];
    return msg;
  }

  void release(Context& ctx) {
    ctx.tell(m_->rm, "release", obj({{"task_id", task_id_}}));
    allocated_ = false;
  }

  void progress(Context& ctx) {
    if (finished_) return;
    if (!running_) {
      if (wants_to_run()) {
        if (!allocation_requested_ && !allocated_) {
          allocation_requested_ = true;
          preempt_requested_ = false;
          const Json& res = config_["resources"];
          m_->system.tell(m_->rm, "allocate",
                          obj({{"task_id", task_id_}, {"name", "Trial (Experiment " + std::to_string(exp_id_) + ")"},
                               {"group", "experiment-" + std::to_string(exp_id_)}, {"slots_needed", res["slots_per_trial"].as_int(1)},
                               {"label", res["agent_label"].str_or("")}}),
                          ctx.self());
        }
      } else if (allocation_requested_ && !allocated_) {
        allocation_requested_ = false;  // paused / finished while waiting in the queue
        release(ctx);
      }
      if (seq_.up_to_date() && close_requested_ && !allocation_requested_) finish(ctx, "COMPLETED");
      else if (is_stopping_state(exp_state_) && !allocation_requested_) finish(ctx, exp_state_ == "STOPPING_COMPLETED" ? "COMPLETED" : "CANCELED");
      return;
    }
    if (!poll_.reply || inflight_) return;
    const bool terminate = seq_.up_to_date() || exp_state_ != "ACTIVE" || preempt_requested_ || killed_;
    if (terminate) {
      Workload ck;
      if (!killed_ && !is_kill_state() && seq_.preclose_checkpoint_workload(&ck)) {
        send_workload(ck);
        return;
      }
      poll_.respond(obj({{"type", "TERMINATE"}, {"workload", seq_.terminate_workload().to_json()}}));
      poll_ = Message();
      awaiting_exit_ = true;
      ctx.system().notify_after(ctx.self(), m_->cfg.terminate_timeout_ms, "terminate_timeout", Json(static_cast<long long>(run_id_)));
      return;
    }
    send_workload(seq_.workload());
  }
  bool is_kill_state() const { return exp_state_ == "STOPPING_ERROR"; }

  void send_workload(const Workload& w) {
    inflight_ = true;
    inflight_workload_ = w;
    poll_.respond(obj({{"type", "RUN_WORKLOAD"}, {"workload", w.to_json()}}));
    poll_ = Message();
  }

  void launch(Context& ctx, const Json& allocations) {
    allocated_ = true;
    if (trial_id_ == 0) {
      trial_id_ = m_->store.next_id("trials");
      seq_.set_trial_id(static_cast<int>(trial_id_));
      m_->store.put("trials", trial_id_, obj({{"id", trial_id_}, {"experiment_id", exp_id_}, {"request_id", create_.request_id},
                                               {"state", "ACTIVE"}, {"hparams", create_.hparams}, {"seed", static_cast<long long>(create_.trial_seed)},
                                               {"start_time", now_rfc3339()}, {"end_time", Json()}, {"restarts", 0},
                                               {"warm_start_checkpoint", seq_.latest_checkpoint()}}));
      m_->system.tell(m_->experiment_actor(exp_id_), "trial_created",
                      obj({{"request_id", create_.request_id}, {"trial_id", trial_id_}, {"create", create_.to_json()}}), ctx.self());
    }
    ++run_id_;
    containers_.clear();
    const int n = static_cast<int>(allocations.size());
    int world_slots = 0, rank_offset = 0;
    for (auto& a : allocations.arr()) world_slots += static_cast<int>(a["devices"].size());
    std::string chief_addr = "127.0.0.1";
    if (n > 1) {   // the chief is the rank-0 container; a single-container trial keeps everything on loopback
      const Json& c0 = allocations.at(0);
      if (c0["agent"].str_or("") == "agent-local") chief_addr = m_->cfg.host == "0.0.0.0" || m_->cfg.host.empty() ? "127.0.0.1" : m_->cfg.host;
      else if (c0["addresses"].is_array() && c0["addresses"].size() > 0) chief_addr = c0["addresses"].at(0).str_or("127.0.0.1");
    }
    for (auto& a : allocations.arr()) {
      ContainerSpec spec;
      spec.container_id = a["container_id"].as_string();
      spec.task_id = task_id_;
      spec.description = "trial " + std::to_string(trial_id_);
      spec.argv = {m_->cfg.python, "-m", "determined_b200.exec.harness"};
      spec.cwd = m_->model_def_dir(exp_id_);
      for (auto& d : a["devices"].arr()) spec.devices.push_back(Device{static_cast<int>(d["id"].as_int()), d["brand"].as_string(), d["uuid"].as_string(), d["type"].as_string()});
      auto& e = spec.env;
      e["DET_MASTER_ADDR"] = m_->cfg.host == "0.0.0.0" ? "127.0.0.1" : m_->cfg.host;
      e["DET_MASTER_PORT"] = std::to_string(m_->http.port());
      e["DET_MASTER_HOST"] = e["DET_MASTER_ADDR"];
      if (m_->http.tls_enabled()) {   // the task pins the master's own certificate (reference tasks/task.go: master cert archive)
        e["DET_USE_TLS"] = "true";
        e["DET_MASTER_CERT_FILE"] = m_->cfg.tls_cert;
        if (!m_->cfg.tls_cert_name.empty()) e["DET_MASTER_CERT_NAME"] = m_->cfg.tls_cert_name;
      }
      e["DET_TRIAL_UNIQUE_PORT_OFFSET"] = std::to_string(trial_id_ % 16);
      e["DET_EXPERIMENT_ID"] = std::to_string(exp_id_);
      e["DET_TRIAL_ID"] = std::to_string(trial_id_);
      e["DET_TRIAL_SEED"] = std::to_string(create_.trial_seed);
      e["DET_TRIAL_RUN_ID"] = std::to_string(run_id_);
      e["DET_EXPERIMENT_CONFIG"] = config_.dump();
      e["DET_HPARAMS"] = create_.hparams.dump();
      Workload first = seq_.up_to_date() ? seq_.terminate_workload() : seq_.workload();
      first.total_batches_processed = seq_.total_batches_processed();
      e["DET_INITIAL_WORKLOAD"] = first.to_json().dump();
      e["DET_LATEST_CHECKPOINT"] = seq_.latest_checkpoint().dump();
      e["DET_WORKLOAD_MANAGER_TYPE"] = "TRIAL_WORKLOAD_MANAGER";
      e["DET_RENDEZVOUS_PORTS"] = std::to_string(1734 + (trial_id_ % 16)) + "," + std::to_string(1750 + (trial_id_ % 16));
      e["DET_TRIAL_RUNNER_NETWORK_INTERFACE"] = "lo";
      {
        // master-wide task_container_defaults (reference master/pkg/tasks/task.go TrialEnvVars): the interface and port ranges
        // the collectives' bootstrap may use — handed to NCCL and gloo in their own variables as well — and the shared-memory
        // size a trial gets unless its resources.shm_size says otherwise
        const Json& tcd = m_->cfg.task_container_defaults;
        const std::string nic = tcd["dtrain_network_interface"].str_or("");
        if (!nic.empty()) {
          e["DET_TRIAL_RUNNER_NETWORK_INTERFACE"] = nic;
          e["NCCL_SOCKET_IFNAME"] = nic;
          e["GLOO_SOCKET_IFNAME"] = nic;
        }
        if (!tcd["nccl_port_range"].str_or("").empty()) e["NCCL_PORT_RANGE"] = tcd["nccl_port_range"].as_string();
        if (!tcd["gloo_port_range"].str_or("").empty()) e["GLOO_PORT_RANGE"] = tcd["gloo_port_range"].as_string();
        if (tcd["shm_size_bytes"].as_int(0) > 0) e["DET_SHM_SIZE"] = std::to_string(tcd["shm_size_bytes"].as_int());
      }
      e["DET_CLUSTER_ID"] = m_->cluster_id;
      e["DET_AGENT_ID"] = a["agent"].as_string();
      e["DET_CONTAINER_RANK"] = std::to_string(a["rank"].as_int());
      e["DET_NUM_CONTAINERS"] = std::to_string(n);
      // Rendezvous of a trial that spans containers (reference: RENDEZVOUS_INFO pushed by trial.go:813-908 — addresses and
      // ranks of every container): every worker joins ONE process group hosted by the chief container; a worker's global
      // rank is its local rank plus the slots of the containers ranked before its own.
      e["DET_CHIEF_ADDR"] = chief_addr;
      e["DET_CHIEF_PORT"] = std::to_string(20000 + (trial_id_ * 7 + run_id_) % 20000);
      e["DET_RANK_OFFSET"] = std::to_string(rank_offset);
      e["DET_WORLD_SLOTS"] = std::to_string(world_slots);
      rank_offset += static_cast<int>(a["devices"].size());
      e["DET_DEBUG"] = config_["debug"].as_bool(false) ? "true" : "false";
      const char* pp = getenv("PYTHONPATH");
      e["PYTHONPATH"] = spec.cwd + ":" + m_->cfg.repo_root + (pp ? std::string(":") + pp : "");
      {
        const std::string slot_type = (!spec.devices.empty() && spec.devices[0].type == "gpu" && spec.devices[0].brand != "Artificial") ? "gpu" : "cpu";
        std::string warning;
        const std::string bad = apply_task_config(config_, slot_type, trial_id_ % 16, true, e, spec.argv, &warning);
        if (!bad.empty()) m_->log("error", task_id_ + ": " + bad);
        if (!warning.empty()) m_->log("warn", task_id_ + ": " + warning);
      }
      m_->apply_owner_identity(m_->store.get("experiments", exp_id_)["owner"].str_or(""), e);
      m_->register_container(spec.container_id, ctx.self(), trial_id_);
      containers_.push_back(spec.container_id);
      m_->start_container(a["agent"].as_string(), spec);
    }
    running_ = true;
    awaiting_exit_ = false;
    inflight_ = false;
    harness_ready_ = false;
    live_containers_ = static_cast<int>(containers_.size());
    exit_failure_.clear();
    ctx.system().notify_after(ctx.self(), m_->cfg.all_ready_timeout_ms, "all_ready_timeout", Json(static_cast<long long>(run_id_)));
  }

  void on_workload_completed(Context& ctx, const Json& body) {
    if (!inflight_) { m_->log("warn", task_id_ + ": unexpected workload completion"); return; }
    inflight_ = false;
    CompletedMessage msg = parse_completed(body);
    const Workload& w = msg.workload;
    const std::string tkey = std::to_string(trial_id_);
    double units = 0;
    bool is_best = false;
    if (w.kind == Workload::RunStep) {
      units = units_from_batches(w.num_batches, seq_.unit_context());
      m_->store.append("steps", tkey, obj({{"trial_id", trial_id_}, {"step_id", w.step_id}, {"num_batches", w.num_batches},
                                            {"total_batches", w.total_batches_processed + w.num_batches}, {"state", msg.has_exited_reason && msg.exited_reason == ExitedReason::Errored ? "ERROR" : "COMPLETED"},
                                            {"start_time", body["start_time"]}, {"end_time", body["end_time"]}, {"metrics", body["metrics"]}}));
    } else if (w.kind == Workload::ComputeValidationMetrics) {
      m_->store.append("validations", tkey, obj({{"trial_id", trial_id_}, {"step_id", w.step_id}, {"total_batches", w.total_batches_processed},
                                                  {"state", "COMPLETED"}, {"end_time", body["end_time"]}, {"metrics", body["metrics"]}}));
    } else if (w.kind == Workload::CheckpointModel && !msg.has_exited_reason) {
      const std::string uuid = body["metrics"]["uuid"].str_or("");
      m_->store.put("checkpoints", uuid, obj({{"uuid", uuid}, {"trial_id", trial_id_}, {"experiment_id", exp_id_}, {"request_id", create_.request_id},
                                               {"step_id", w.step_id}, {"total_batches", w.total_batches_processed}, {"state", "COMPLETED"},
                                               {"end_time", body["end_time"]}, {"resources", body["metrics"]["resources"]},
                                               {"framework", body["metrics"]["framework"]}, {"format", body["metrics"]["format"]},
                                               {"metadata", Json::object()}, {"determined_version", "0.13.10.dev0+b200.1"}}));
    }
    if (!msg.has_exited_reason || msg.exited_reason != ExitedReason::Errored) {
      Json r = m_->system.ask(m_->experiment_actor(exp_id_), "trial_completed_workload",
                              obj({{"trial_id", trial_id_}, {"kind", Workload::kind_name(w.kind)}, {"units", units}, {"metrics", body["metrics"]},
                                   {"message", body}}), 5000);
      is_best = r["is_best"].as_bool(false);
    }
    m_->store.append("trial_workloads", key(), obj({{"message", body}, {"is_best", is_best}}));
    Operation op; Json metrics;
    bool done = false;
    try {
      done = seq_.workload_completed(msg, is_best, &op, &metrics);
    } catch (const std::exception& e) {
      m_->log("error", task_id_ + ": " + e.what());
    }
    auto exp = m_->experiment_actor(exp_id_);
    if (done) m_->system.tell(exp, "trial_completed_operation", obj({{"trial_id", trial_id_}, {"op", op.to_json()}, {"metrics", metrics}}), ctx.self());
    while (seq_.complete_cached_checkpoints(&op, &metrics))
      m_->system.tell(exp, "trial_completed_operation", obj({{"trial_id", trial_id_}, {"op", op.to_json()}, {"metrics", metrics}}), ctx.self());
    if (msg.has_exited_reason) {
      if (msg.exited_reason == ExitedReason::Errored) {
        exit_failure_ = "workload reported an error";
      } else if (!exited_early_reported_) {
        exited_early_reported_ = true;
        m_->system.tell(exp, "trial_exited_early", obj({{"trial_id", trial_id_}, {"reason", reason_name(msg.exited_reason)}}), ctx.self());
      }
    }
    progress(ctx);
  }

  void on_container_event(Context& ctx, const Json& ev) {
    if (ev["state"].as_string() != "TERMINATED") return;
    const std::string cid = ev["container_id"].as_string();
    if (std::find(containers_.begin(), containers_.end(), cid) == containers_.end()) return;
    m_->unregister_container(cid);
    if (!ev["failure"].as_string().empty() && exit_failure_.empty()) exit_failure_ = ev["failure"].as_string();
    if (--live_containers_ > 0) {
      // any rank failing takes the whole gang down (reference trial.go:944-949)
      if (!ev["failure"].as_string().empty()) for (auto& c : containers_) m_->signal_container(c, SIGKILL);
      return;
    }
    on_exit(ctx, static_cast<int>(ev["exit_code"].as_int()), exit_failure_);
  }

  void on_exit(Context& ctx, int exit_code, const std::string& failure) {
    running_ = false;
    containers_.clear();
    if (poll_.reply) { poll_.respond(obj({{"type", "TERMINATE"}})); poll_ = Message(); }
    release(ctx);
    const bool graceful = awaiting_exit_ && failure.empty();
    awaiting_exit_ = false;
    inflight_ = false;
    if (killed_) { finish(ctx, "CANCELED"); return; }
    if (graceful) { progress(ctx); return; }
    // unexpected exit: restart from the last checkpoint, up to max_restarts
    ++restarts_;
    m_->log("warn", task_id_ + " exited unexpectedly (code " + std::to_string(exit_code) + ": " + failure + "), restart " +
                        std::to_string(restarts_) + "/" + std::to_string(max_restarts_));
    m_->store.update("trials", trial_id_, obj({{"restarts", restarts_}}));
    if (restarts_ <= max_restarts_) {
      const int step = seq_.roll_back();
      roll_back_store(step);
      progress(ctx);
      return;
    }
    m_->system.tell(m_->experiment_actor(exp_id_), "trial_exited_early", obj({{"trial_id", trial_id_}, {"reason", "ERRORED"}}), ctx.self());
    finish(ctx, "ERROR");
  }

  void roll_back_store(int step_id) {
    const std::string tkey = std::to_string(trial_id_);
    auto steps = m_->store.series("steps", tkey);
    size_t keep = 0;
    for (auto& s : steps) if (s["step_id"].as_int() <= step_id) ++keep;
    m_->store.truncate_series("steps", tkey, keep);
    auto vals = m_->store.series("validations", tkey);
    keep = 0;
    for (auto& v : vals) if (v["step_id"].as_int() <= step_id) ++keep;
    m_->store.truncate_series("validations", tkey, keep);
  }

  void finish(Context& ctx, const std::string& state) {
    if (finished_) return;
    finished_ = true;
    if (allocation_requested_ || allocated_) release(ctx);
    if (trial_id_ > 0) m_->store.update("trials", trial_id_, obj({{"state", state}, {"end_time", now_rfc3339()}}));
    m_->system.tell(m_->experiment_actor(exp_id_), "trial_closed", obj({{"request_id", create_.request_id}, {"trial_id", trial_id_}, {"state", state}}), ctx.self());
    ctx.stop_self();
  }

  Master* m_;
  long long exp_id_;
  Json config_;
  Operation create_;
  std::string exp_state_;
  TrialWorkloadSequencer seq_;
  long long trial_id_ = 0;
  std::string task_id_;
  int max_restarts_ = 5, restarts_ = 0, run_id_ = 0, live_containers_ = 0;
  bool allocation_requested_ = false, allocated_ = false, running_ = false, inflight_ = false, awaiting_exit_ = false;
  bool harness_ready_ = false;     // this run's harness has asked for a workload (the all-ready timeout watches it)
  bool close_requested_ = false, preempt_requested_ = false, killed_ = false, finished_ = false, exited_early_reported_ = false;
  Workload inflight_workload_;
  Message poll_;
  unsigned long long poll_seq_ = 0;
  std::vector<std::string> containers_;
  std::string exit_failure_;
};

// =====================================================================================================================
class ExperimentActor : public Actor {
 public:
  ExperimentActor(Master* m, long long id, bool restoring) : m_(m), id_(id), restoring_(restoring) {}

  void pre_start(Context& ctx) override {
    row_ = m_->store.get("experiments", id_);
    config_ = row_["config"];
    state_ = row_["state"].str_or("PAUSED");
    scfg_ = SearcherConfig::from_json(config_["searcher"]);
    searcher_.reset(new Searcher(static_cast<uint32_t>(config_["reproducibility"]["experiment_seed"].as_int(0)), new_search_method(scfg_),
                                 hyperparameters_from_json(config_["hyperparameters"])));
    const Json& res = config_["resources"];
    m_->system.tell(m_->rm, "set_group", obj({{"group", "experiment-" + std::to_string(id_)}, {"max_slots", res["max_slots"]},
                                               {"weight", res["weight"].as_double(1.0)}, {"priority", res["priority"]}}), ctx.self());
    if (row_.contains("best_validation") && row_["best_validation"].is_number()) { has_best_ = true; best_ = row_["best_validation"].as_double(); }
    if (restoring_) restore(ctx);
    else {
      m_->store.append("searcher_events", std::to_string(id_), obj({{"type", "initial"}}));
      process(ctx, searcher_->initial_operations());
    }
  }

  void receive(Context& ctx, const Message& msg) override {
    const std::string& t = msg.type;
    try {
      if (t == "trial_created") {
        record(obj({{"type", "trial_created"}, {"create", msg.body["create"]}, {"trial_id", msg.body["trial_id"]}}));
        process(ctx, searcher_->trial_created(Operation::from_json(msg.body["create"]), static_cast<int>(msg.body["trial_id"].as_int())));
      } else if (t == "trial_completed_workload") {
        bool is_best = false;
        const std::string kind = msg.body["kind"].as_string();
        if (kind == "RUN_STEP") {
          record(obj({{"type", "workload_completed"}, {"units", msg.body["units"]}}));
          searcher_->workload_completed(msg.body["message"], msg.body["units"].as_double());
        } else if (kind == "COMPUTE_VALIDATION_METRICS") {
          is_best = is_best_validation(msg.body["metrics"]);
        }
        m_->store.update("experiments", id_, obj({{"progress", searcher_->progress()}, {"best_validation", has_best_ ? Json(best_) : Json()}}));
        msg.respond(obj({{"is_best", is_best}}));
      } else if (t == "trial_completed_operation") {
        record(obj({{"type", "operation_completed"}, {"trial_id", msg.body["trial_id"]}, {"op", msg.body["op"]}, {"metrics", msg.body["metrics"]}}));
        process(ctx, searcher_->operation_completed(static_cast<int>(msg.body["trial_id"].as_int()), Operation::from_json(msg.body["op"]), msg.body["metrics"]));
      } else if (t == "trial_exited_early") {
        record(obj({{"type", "trial_exited_early"}, {"trial_id", msg.body["trial_id"]}, {"reason", msg.body["reason"]}}));
        process(ctx, searcher_->trial_exited_early(static_cast<int>(msg.body["trial_id"].as_int()), reason_from(msg.body["reason"].as_string())));
      } else if (t == "trial_closed") {
        const std::string rid = msg.body["request_id"].as_string();
        trials_.erase(rid);
        if (msg.body["trial_id"].as_int() > 0 || created_.count(rid)) {
          record(obj({{"type", "trial_closed"}, {"request_id", rid}}));
          int tid;
          if (searcher_->trial_id_of(rid, &tid)) process(ctx, searcher_->trial_closed(rid));
        }
        maybe_finish(ctx);
      } else if (t == "set_state") {
        set_state(ctx, msg.body.as_string());
        msg.respond(obj({{"state", state_}}));
      } else if (t == "kill") {
        set_state(ctx, "STOPPING_CANCELED");
        for (auto& kv : trials_) ctx.tell(kv.second, "kill");
        maybe_finish(ctx);
        msg.respond(obj({{"state", state_}}));
      } else if (t == "set_resources") {
        Json& res = config_["resources"];
        for (auto& kv : msg.body.obj()) res[kv.first] = kv.second;
        m_->store.update("experiments", id_, obj({{"config", config_}}));
        m_->system.tell(m_->rm, "set_group", obj({{"group", "experiment-" + std::to_string(id_)}, {"max_slots", res["max_slots"]},
                                                   {"weight", res["weight"].as_double(1.0)}, {"priority", res["priority"]}}), ctx.self());
        msg.respond(Json(true));
      } else if (t == "describe") {
        Json d = Json::object();
        d["state"] = state_; d["progress"] = searcher_->progress(); d["active_trials"] = static_cast<long long>(trials_.size());
        msg.respond(d);
      } else {
        msg.respond(Json());
      }
    } catch (const std::exception& e) {
      m_->log("error", "experiment " + std::to_string(id_) + ": " + e.what());
      msg.respond(obj({{"error", e.what()}}));
      set_state(ctx, "STOPPING_ERROR");
      maybe_finish(ctx);
    }
  }

  void child_failed(Context& ctx, const ActorRef& child, const std::string& error) override {
    m_->log("error", "trial actor " + child->address() + " failed: " + error);
    for (auto it = trials_.begin(); it != trials_.end(); ++it) if (it->second == child) { trials_.erase(it); break; }
    set_state(ctx, "STOPPING_ERROR");
    maybe_finish(ctx);
  }

 private:
  void record(const Json& ev) {
    if (!replaying_) m_->store.append("searcher_events", std::to_string(id_), ev);
  }

  bool is_best_validation(const Json& metrics) {
    double v;
    try { v = metric_from(metrics, scfg_.metric); } catch (...) { return false; }
    if (!scfg_.smaller_is_better) v = -v;
    if (!has_best_ || v < best_) { has_best_ = true; best_ = v; return true; }
    return false;
  }

  Json first_checkpoint_for(const Operation& create) {
    if (create.has_checkpoint) {  // PBT clone: latest checkpoint of the source trial
      Json best;
      for (auto& c : m_->store.list("checkpoints", [&](const Json& r) { return r["experiment_id"].as_int() == id_ && r["request_id"].as_string() == create.checkpoint_request_id && r["state"].as_string() == "COMPLETED"; }))
        if (best.is_null() || c["total_batches"].as_int() >= best["total_batches"].as_int()) best = c;
      return best;
    }
    if (!scfg_.source_checkpoint_uuid.empty()) return m_->store.get("checkpoints", scfg_.source_checkpoint_uuid);
    if (scfg_.source_trial_id >= 0) {
      Json best;
      for (auto& c : m_->store.list("checkpoints", [&](const Json& r) { return r["trial_id"].as_int() == scfg_.source_trial_id && r["state"].as_string() == "COMPLETED"; }))
        if (best.is_null() || c["total_batches"].as_int() >= best["total_batches"].as_int()) best = c;
      return best;
    }
    return Json();
  }

  void process(Context& ctx, const Ops& ops) {
    for (auto& op : ops) {
      switch (op.kind) {
        case Operation::Create: {
          created_.insert(op.request_id);
          if (replaying_) { pending_creates_[op.request_id] = op; break; }
          auto* actor = new TrialActor(m_, id_, config_, op, first_checkpoint_for(op), state_, 0);
          trials_[op.request_id] = ctx.actor_of(op.request_id, std::unique_ptr<Actor>(actor));
          break;
        }
        case Operation::Train: case Operation::Validate: case Operation::Checkpoint: case Operation::Close: {
          if (replaying_) { replay_ops_[op.request_id].push_back(op); break; }
          auto it = trials_.find(op.request_id);
          if (it != trials_.end()) ctx.tell(it->second, "operation", op.to_json());
          break;
        }
        case Operation::Shutdown:
          shutdown_requested_ = true;
          if (!replaying_) set_state(ctx, op.failure ? "STOPPING_ERROR" : "STOPPING_COMPLETED");
          break;
      }
    }
    if (!replaying_) m_->store.update("experiments", id_, obj({{"progress", searcher_->progress()}}));
  }

  void set_state(Context& ctx, const std::string& next) {
    if (next == state_) return;
    if (!experiment_transition_ok(state_, next)) { m_->log("warn", "experiment " + std::to_string(id_) + ": illegal transition " + state_ + " -> " + next); return; }
    state_ = next;
    Json upd = obj({{"state", state_}});
    if (is_terminal_state(state_)) { upd["end_time"] = now_rfc3339(); upd["progress"] = state_ == "COMPLETED" ? Json(1.0) : Json(searcher_->progress()); }
    m_->store.update("experiments", id_, upd);
    for (auto& kv : trials_) ctx.tell(kv.second, "experiment_state", Json(state_));
  }

  void maybe_finish(Context& ctx) {
    if (!is_stopping_state(state_) || !trials_.empty()) return;
    const std::string terminal = state_ == "STOPPING_COMPLETED" ? "COMPLETED" : state_ == "STOPPING_CANCELED" ? "CANCELED" : "ERROR";
    set_state(ctx, terminal);
    m_->log("info", "experiment " + std::to_string(id_) + " finished: " + terminal);
    m_->report_telemetry("experiment_state_changed", obj({{"id", id_}, {"state", terminal}, {"num_trials", static_cast<long long>(m_->store.list("trials", [this](const Json& r) { return r["experiment_id"].as_int() == id_; }).size())}}));
    m_->spawn_checkpoint_gc(id_);
    ctx.stop_self();
  }

  // Replay the stored searcher events into a fresh searcher (no DB writes, no launches), then re-create the trial
  // actors that were still open, each rolled back to its last checkpoint. (reference experiment.go:238-294)
  void restore(Context& ctx) {
    replaying_ = true;
    std::map<int, RequestID> by_trial_id;
    std::set<RequestID> closed;
    for (auto& ev : m_->store.series("searcher_events", std::to_string(id_))) {
      const std::string type = ev["type"].as_string();
      try {
        if (type == "initial") process(ctx, searcher_->initial_operations());
        else if (type == "trial_created") {
          Operation c = Operation::from_json(ev["create"]);
          by_trial_id[static_cast<int>(ev["trial_id"].as_int())] = c.request_id;
          process(ctx, searcher_->trial_created(c, static_cast<int>(ev["trial_id"].as_int())));
        } else if (type == "workload_completed") searcher_->workload_completed(Json::object(), ev["units"].as_double());
        else if (type == "operation_completed") process(ctx, searcher_->operation_completed(static_cast<int>(ev["trial_id"].as_int()), Operation::from_json(ev["op"]), ev["metrics"]));
        else if (type == "trial_exited_early") process(ctx, searcher_->trial_exited_early(static_cast<int>(ev["trial_id"].as_int()), reason_from(ev["reason"].as_string())));
        else if (type == "trial_closed") { closed.insert(ev["request_id"].as_string()); process(ctx, searcher_->trial_closed(ev["request_id"].as_string())); }
      } catch (const std::exception& e) {
        m_->log("error", "experiment " + std::to_string(id_) + " replay: " + e.what());
      }
    }
    replaying_ = false;
    for (auto& kv : pending_creates_) {
      if (closed.count(kv.first)) continue;
      long long trial_id = 0;
      for (auto& t : by_trial_id) if (t.second == kv.first) trial_id = t.first;
      auto* actor = new TrialActor(m_, id_, config_, kv.second, first_checkpoint_for(kv.second), state_, trial_id);
      for (auto& op : replay_ops_[kv.first]) actor->restore_op(op);
      for (auto& w : m_->store.series("trial_workloads", std::to_string(id_) + ":" + kv.first)) actor->restore_completed(w["message"], w["is_best"].as_bool(false));
      actor->restore_finish();
      auto ref = ctx.actor_of(kv.first, std::unique_ptr<Actor>(actor));
      trials_[kv.first] = ref;
      ctx.tell(ref, "experiment_state", Json(state_));
    }
    pending_creates_.clear();
    replay_ops_.clear();
    if (shutdown_requested_) set_state(ctx, "STOPPING_COMPLETED");
    maybe_finish(ctx);
    m_->log("info", "restored experiment " + std::to_string(id_) + " with " + std::to_string(trials_.size()) + " open trials");
  }

  Master* m_;
  long long id_;
  bool restoring_, replaying_ = false, shutdown_requested_ = false;
  Json row_, config_;
  std::string state_;
  SearcherConfig scfg_;
  std::unique_ptr<Searcher> searcher_;
  std::map<RequestID, ActorRef> trials_;
  std::set<RequestID> created_;
  std::map<RequestID, Operation> pending_creates_;
  std::map<RequestID, std::vector<Operation>> replay_ops_;
  bool has_best_ = false;
  double best_ = 0;
};

}  // namespace

std::unique_ptr<Actor> make_experiment(Master* m, long long id, bool restoring) { return std::unique_ptr<Actor>(new ExperimentActor(m, id, restoring)); }

}  // namespace det
