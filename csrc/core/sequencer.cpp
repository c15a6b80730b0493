// sequencer.cpp — see sequencer.hpp.
#include "sequencer.hpp"

#include <algorithm>
#include <tuple>

namespace det {

static int global_batch_size_of(const Json& hparams) {
  const Json& g = hparams["global_batch_size"];
  if (!g.is_number()) throw ConfigError("hyperparameter global_batch_size must resolve to a number");
  return static_cast<int>(g.as_int());
}

TrialWorkloadSequencer::TrialWorkloadSequencer(const Json& cfg, int experiment_id, const Operation& create, const Json& first_checkpoint)
    : experiment_id_(experiment_id) {
  st_.need_initial_validation = cfg["perform_initial_validation"].as_bool(false);
  st_.latest_checkpoint = first_checkpoint;
  snapshot_ = st_;
  checkpoint_policy_ = cfg["checkpoint_policy"].str_or("best");
  min_validation_period_ = cfg.contains("min_validation_period") ? Length::from_json(cfg["min_validation_period"]) : Length{Unit::Batches, 0};
  min_checkpoint_period_ = cfg.contains("min_checkpoint_period") ? Length::from_json(cfg["min_checkpoint_period"]) : Length{Unit::Batches, 0};
  SearcherConfig sc = SearcherConfig::from_json(cfg["searcher"]);
  unit_ctx_.default_unit = sc.unit();
  unit_ctx_.global_batch_size = global_batch_size_of(create.hparams);
  unit_ctx_.records_per_epoch = static_cast<int>(cfg["records_per_epoch"].as_int(0));
  scheduling_unit_ = static_cast<int>(cfg["scheduling_unit"].as_int(100));
}

void TrialWorkloadSequencer::operation_requested(const Operation& op) {
  if (!op.runnable()) throw std::runtime_error("illegal workload for trialWorkloadSequencer: " + op.str());
  ops_.push_back(op);
}

bool TrialWorkloadSequencer::up_to_date() const {
  return static_cast<int>(ops_.size()) == st_.cur_op_idx || (st_.exiting_early && !post_graceful_stop_ckpt_needed());
}

Workload TrialWorkloadSequencer::make(Workload::Kind k, int step_id, int num_batches) const {
  Workload w;
  w.kind = k; w.experiment_id = experiment_id_; w.trial_id = trial_id_; w.step_id = step_id;
  w.num_batches = num_batches; w.total_batches_processed = st_.total_batches_processed;
  return w;
}

bool TrialWorkloadSequencer::min_validation_needed() const {
  return min_validation_period_.units != 0 && min_validation_period_.equal_within_batch(st_.batches_since_last_val, unit_ctx_);
}
bool TrialWorkloadSequencer::min_checkpoint_needed() const {
  return min_checkpoint_period_.units != 0 && min_checkpoint_period_.equal_within_batch(st_.batches_since_last_ckpt, unit_ctx_);
}
int TrialWorkloadSequencer::batches_until_val_needed() const {
  return min_validation_period_.units == 0 ? INT_MAX : min_validation_period_.to_nearest_batch(unit_ctx_) - st_.batches_since_last_val;
}
int TrialWorkloadSequencer::batches_until_ckpt_needed() const {
  return min_checkpoint_period_.units == 0 ? INT_MAX : min_checkpoint_period_.to_nearest_batch(unit_ctx_) - st_.batches_since_last_ckpt;
}

Workload TrialWorkloadSequencer::workload() const {
  if (up_to_date()) throw std::runtime_error("cannot call sequencer.workload() with sequencer.up_to_date() == true");
  if (!trial_id_valid_) throw std::runtime_error("cannot call sequencer.workload() before sequencer.set_trial_id()");
  if (st_.need_initial_validation) return make(Workload::ComputeValidationMetrics, st_.cur_step_id);
  if (post_graceful_stop_ckpt_needed()) return make(Workload::CheckpointModel, st_.cur_step_id);
  if (post_validation_ckpt_needed()) return make(Workload::CheckpointModel, st_.cur_step_id);
  if (min_validation_needed()) return make(Workload::ComputeValidationMetrics, st_.cur_step_id);
  if (min_checkpoint_needed()) return make(Workload::CheckpointModel, st_.cur_step_id);
  const Operation& op = ops_[static_cast<size_t>(st_.cur_op_idx)];
  switch (op.kind) {
    case Operation::Validate:
      if (st_.batches_since_last_ckpt != 0) return make(Workload::CheckpointModel, st_.cur_step_id);
      return make(Workload::ComputeValidationMetrics, st_.cur_step_id);
    case Operation::Checkpoint:
      return make(Workload::CheckpointModel, st_.cur_step_id);
    case Operation::Train: {
      const int left = op.length.to_nearest_batch(unit_ctx_) - st_.batches_towards_current_op;
      const int n = std::max(std::min(std::min(left, batches_until_val_needed()), std::min(batches_until_ckpt_needed(), scheduling_unit_)), 1);
      return make(Workload::RunStep, st_.cur_step_id + 1, n);
    }
    default:
      throw std::runtime_error("unexpected op type determining workload");
  }
}

bool TrialWorkloadSequencer::workload_completed(const CompletedMessage& msg, bool is_best_validation, Operation* out_op, Json* out_metrics) {
  if (up_to_date()) {
    if (msg.workload.kind != Workload::CheckpointModel)
      throw std::runtime_error(std::string("illegal non-checkpoint workload completed message received: ") + Workload::kind_name(msg.workload.kind));
  } else {
    Workload w = workload();
    if (!(msg.workload == w) && msg.workload.kind != Workload::CheckpointModel)
      throw std::runtime_error(std::string("illegal completed message received: expected checkpoint or ") + Workload::kind_name(w.kind) +
                               ", got " + Workload::kind_name(msg.workload.kind));
  }
  if (msg.has_exited_reason) {
    st_.exiting_early = true;
    if (msg.exited_reason == ExitedReason::UserCanceled || msg.exited_reason == ExitedReason::InvalidHP) st_.graceful_stop = true;
    else return false;
  }
  switch (msg.workload.kind) {
    case Workload::RunStep: {
      ++st_.cur_step_id;
      const int n = msg.workload.num_batches;
      st_.total_batches_processed += n; st_.batches_towards_current_op += n;
      st_.batches_since_last_val += n; st_.batches_since_last_ckpt += n;
      if (st_.cur_op_idx < static_cast<int>(ops_.size())) {
        const Operation& op = ops_[static_cast<size_t>(st_.cur_op_idx)];
        if (op.kind == Operation::Train && op.length.equal_within_batch(st_.batches_towards_current_op, unit_ctx_)) {
          ++st_.cur_op_idx;
          st_.batches_towards_current_op = 0;
          *out_op = op;
          *out_metrics = msg.metrics;
          return true;
        }
      }
      return false;
    }
    case Workload::ComputeValidationMetrics: {
      st_.batches_since_last_val = 0;
      st_.need_initial_validation = false;
      if (st_.batches_since_last_ckpt != 0) {
        if (checkpoint_policy_ == "all") st_.need_post_validation_ckpt = true;
        else if (checkpoint_policy_ == "best" && is_best_validation) st_.need_post_validation_ckpt = true;
      }
      if (st_.cur_op_idx < static_cast<int>(ops_.size()) && ops_[static_cast<size_t>(st_.cur_op_idx)].kind == Operation::Validate) {
        *out_op = ops_[static_cast<size_t>(st_.cur_op_idx)];
        ++st_.cur_op_idx;
        if (st_.batches_since_last_ckpt == 0) snapshot();
        *out_metrics = msg.validation_metrics;
        return true;
      }
      if (st_.batches_since_last_ckpt == 0) snapshot();
      return false;
    }
    case Workload::CheckpointModel: {
      st_.batches_since_last_ckpt = 0;
      st_.need_post_validation_ckpt = false;
      st_.latest_checkpoint = msg.checkpoint_metrics;
      bool done = false;
      if (!up_to_date() && st_.cur_op_idx < static_cast<int>(ops_.size()) &&
          ops_[static_cast<size_t>(st_.cur_op_idx)].kind == Operation::Checkpoint) {
        *out_op = ops_[static_cast<size_t>(st_.cur_op_idx)];
        ++st_.cur_op_idx;
        *out_metrics = msg.checkpoint_metrics;
        done = true;
      } else {
        st_.cached_checkpoints[msg.workload] = msg;
      }
      snapshot();
      return done;
    }
    default:
      throw std::runtime_error("invalid operation for trialWorkloadSequencer");
  }
}

bool TrialWorkloadSequencer::complete_cached_checkpoints(Operation* op, Json* metrics) {
  if (up_to_date()) return false;
  Workload w = workload();
  auto it = st_.cached_checkpoints.find(w);
  if (it == st_.cached_checkpoints.end()) return false;
  CompletedMessage msg = it->second;
  st_.cached_checkpoints.erase(it);
  return workload_completed(msg, false, op, metrics);
}

bool TrialWorkloadSequencer::preclose_checkpoint_workload(Workload* out) const {
  if (st_.batches_since_last_ckpt == 0 || !trial_id_valid_) return false;
  *out = make(Workload::CheckpointModel, st_.cur_step_id);
  return true;
}

Workload TrialWorkloadSequencer::terminate_workload() const {
  Workload w = make(Workload::Terminate, st_.cur_step_id);
  w.total_batches_processed = 0;
  return w;
}

int TrialWorkloadSequencer::roll_back() {
  st_ = snapshot_;
  return st_.cur_step_id;
}

Json TrialWorkloadSequencer::snapshot_json() const {
  Json j = Json::object();
  j["cur_op_idx"] = st_.cur_op_idx; j["cur_step_id"] = st_.cur_step_id;
  j["total_batches_processed"] = st_.total_batches_processed;
  j["batches_since_last_ckpt"] = st_.batches_since_last_ckpt; j["batches_since_last_val"] = st_.batches_since_last_val;
  j["latest_checkpoint"] = st_.latest_checkpoint;
  j["exiting_early"] = st_.exiting_early; j["graceful_stop"] = st_.graceful_stop;
  return j;
}

}  // namespace det
