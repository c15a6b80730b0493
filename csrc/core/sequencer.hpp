// sequencer.hpp — trial workload sequencer: turns searcher operations (Train(len) / Validate / Checkpoint) into
// the concrete workloads a trial runs. Slices training into scheduling_unit-batch steps; injects validations for
// min_validation_period and checkpoints for min_checkpoint_period and checkpoint_policy (best/all/none);
// always checkpoints before a searcher Validate; pre-close and graceful-stop checkpoints; optional initial
// validation; snapshots its state at every checkpoint so a failed trial can roll back; tolerates checkpoint
// completions that arrive out of order. Parity: reference master/internal/trial_workload_sequencer.go:21-467.
#pragma once
#include <climits>
#include <map>
#include <string>
#include <vector>

#include "json.hpp"
#include "model.hpp"
#include "searcher.hpp"

namespace det {

struct Workload {
  enum Kind { RunStep, ComputeValidationMetrics, CheckpointModel, Terminate } kind = RunStep;
  int experiment_id = 0, trial_id = 0, step_id = 0, num_batches = 0, total_batches_processed = 0;

  bool operator==(const Workload& o) const {
    return kind == o.kind && experiment_id == o.experiment_id && trial_id == o.trial_id && step_id == o.step_id &&
           num_batches == o.num_batches && total_batches_processed == o.total_batches_processed;
  }
  bool operator<(const Workload& o) const {
    auto key = [](const Workload& w) { return std::make_tuple(static_cast<int>(w.kind), w.experiment_id, w.trial_id, w.step_id, w.num_batches, w.total_batches_processed); };
    return key(*this) < key(o);
  }
  static const char* kind_name(Kind k) {
    switch (k) {
      case RunStep: return "RUN_STEP";
      case ComputeValidationMetrics: return "COMPUTE_VALIDATION_METRICS";
      case CheckpointModel: return "CHECKPOINT_MODEL";
      default: return "TERMINATE";
    }
  }
  Json to_json() const {
    Json j = Json::object();
    j["kind"] = kind_name(kind);
    j["experiment_id"] = experiment_id; j["trial_id"] = trial_id; j["step_id"] = step_id;
    j["num_batches"] = num_batches; j["total_batches_processed"] = total_batches_processed;
    return j;
  }
  static Workload from_json(const Json& j) {
    Workload w;
    const std::string k = j["kind"].as_string();
    w.kind = k == "RUN_STEP" ? RunStep : k == "COMPUTE_VALIDATION_METRICS" ? ComputeValidationMetrics : k == "CHECKPOINT_MODEL" ? CheckpointModel : Terminate;
    w.experiment_id = static_cast<int>(j["experiment_id"].as_int()); w.trial_id = static_cast<int>(j["trial_id"].as_int());
    w.step_id = static_cast<int>(j["step_id"].as_int()); w.num_batches = static_cast<int>(j["num_batches"].as_int());
    w.total_batches_processed = static_cast<int>(j["total_batches_processed"].as_int());
    return w;
  }
};

struct CompletedMessage {
  Workload workload;
  bool has_exited_reason = false;
  ExitedReason exited_reason = ExitedReason::Errored;
  Json metrics;             // RUN_STEP metrics
  Json validation_metrics;  // {"num_inputs":..,"validation_metrics":{..}}
  Json checkpoint_metrics;  // {"uuid":..,"resources":{..},"framework":..,"format":..}
};

class TrialWorkloadSequencer {
 public:
  TrialWorkloadSequencer(const Json& experiment_config, int experiment_id, const Operation& create, const Json& first_checkpoint);

  void set_trial_id(int id) { trial_id_ = id; trial_id_valid_ = true; }
  const Json& latest_checkpoint() const { return st_.latest_checkpoint; }
  void operation_requested(const Operation& op);
  bool up_to_date() const;
  Workload workload() const;
  // Returns true and fills *op / *metrics when a searcher operation finished as a result of this completion.
  bool workload_completed(const CompletedMessage& msg, bool is_best_validation, Operation* op, Json* metrics);
  bool complete_cached_checkpoints(Operation* op, Json* metrics);
  bool preclose_checkpoint_workload(Workload* out) const;
  Workload terminate_workload() const;
  int roll_back();  // restore the snapshot taken at the last checkpoint; returns the step id to resume from
  int total_batches_processed() const { return st_.total_batches_processed; }
  const UnitContext& unit_context() const { return unit_ctx_; }
  Json snapshot_json() const;

 private:
  struct State {
    int batches_towards_current_op = 0, batches_since_last_val = 0, batches_since_last_ckpt = 0, total_batches_processed = 0;
    bool need_initial_validation = false, need_post_validation_ckpt = false, exiting_early = false, graceful_stop = false;
    int cur_op_idx = 0, cur_step_id = 0;
    Json latest_checkpoint;
    std::map<Workload, CompletedMessage> cached_checkpoints;
  };
  Workload make(Workload::Kind k, int step_id, int num_batches = 0) const;
  bool min_validation_needed() const;
  bool min_checkpoint_needed() const;
  int batches_until_val_needed() const;
  int batches_until_ckpt_needed() const;
  bool post_graceful_stop_ckpt_needed() const { return st_.graceful_stop && st_.batches_since_last_ckpt != 0; }
  bool post_validation_ckpt_needed() const { return st_.need_post_validation_ckpt && st_.batches_since_last_ckpt != 0; }
  void snapshot() { snapshot_ = st_; }

  std::vector<Operation> ops_;
  State st_, snapshot_;
  std::string checkpoint_policy_;
  Length min_validation_period_, min_checkpoint_period_;
  UnitContext unit_ctx_;
  int scheduling_unit_ = 100;
  int experiment_id_ = 0;
  int trial_id_ = 0;
  bool trial_id_valid_ = false;
};

}  // namespace det
