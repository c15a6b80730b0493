// model.hpp — experiment-configuration model: Length / units, hyperparameter specs, searcher configs with
// defaults and validation, experiment-config defaults / merge / validation, state machines.
// Parity: reference master/pkg/model/{length.go:13-196, hyperparameters_config.go:17-155,
// searcher_config.go:17-332, experiment_config.go:22-118, defaults.go:33-128, experiment.go:104-176}.
#pragma once
#include <algorithm>
#include <cmath>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>

#include "json.hpp"

namespace det {

struct ConfigError : std::runtime_error {
  using std::runtime_error::runtime_error;
};

// ---- Length ---------------------------------------------------------------------------------------------
enum class Unit { Records, Batches, Epochs, Unspecified };

inline const char* unit_name(Unit u) {
  switch (u) {
    case Unit::Records: return "records";
    case Unit::Batches: return "batches";
    case Unit::Epochs: return "epochs";
    default: return "unspecified";
  }
}

struct UnitContext {
  Unit default_unit = Unit::Batches;
  int global_batch_size = 1;
  int records_per_epoch = 0;
};

struct Length {
  Unit unit = Unit::Unspecified;
  int units = 0;

  static Length from_json(const Json& j) {
    if (!j.is_object() || j.size() != 1) throw ConfigError("a length must be {records|batches|epochs: N}, got " + j.dump());
    const auto& kv = *j.obj().begin();
    Length l;
    if (kv.first == "records") l.unit = Unit::Records;
    else if (kv.first == "batches") l.unit = Unit::Batches;
    else if (kv.first == "epochs") l.unit = Unit::Epochs;
    else throw ConfigError("invalid length unit '" + kv.first + "'");
    if (!kv.second.is_number()) throw ConfigError("length value must be a number");
    l.units = static_cast<int>(kv.second.as_int());
    return l;
  }
  Json to_json() const {
    Json j = Json::object();
    j[unit_name(unit)] = units;
    return j;
  }
  Length mult(int k) const { return Length{unit, units * k}; }
  Length div(int k) const { return Length{unit, units / k}; }
  std::string str() const { return std::to_string(units) + " " + unit_name(unit); }

  int to_nearest_batch(const UnitContext& c) const {
    switch (unit) {
      case Unit::Records: return units / c.global_batch_size;
      case Unit::Batches: return units;
      case Unit::Epochs: return static_cast<int>((static_cast<long long>(units) * c.records_per_epoch) / c.global_batch_size);
      default: throw ConfigError("length without a unit");
    }
  }
  bool equal_within_batch(int batches, const UnitContext& c) const {
    switch (unit) {
      case Unit::Records: return std::llabs(static_cast<long long>(units) - static_cast<long long>(batches) * c.global_batch_size) < c.global_batch_size;
      case Unit::Batches: return units == batches;
      case Unit::Epochs:
        return std::llabs(static_cast<long long>(units) * c.records_per_epoch - static_cast<long long>(batches) * c.global_batch_size) < c.global_batch_size;
      default: throw ConfigError("length without a unit");
    }
  }
};

inline double units_from_batches(int batches, const UnitContext& c) {
  switch (c.default_unit) {
    case Unit::Records: return static_cast<double>(batches) * c.global_batch_size;
    case Unit::Batches: return batches;
    case Unit::Epochs: return static_cast<double>(batches) * c.global_batch_size / c.records_per_epoch;
    default: throw ConfigError("unit context without a unit");
  }
}

// ---- Hyperparameters --------------------------------------------------------------------------------------
struct Hyperparameter {
  enum Kind { Const, Int, Double, Log, Categorical } kind = Const;
  Json val;                 // const
  double minval = 0, maxval = 0, base = 10;
  int count = -1;           // grid points (-1 = unset)
  std::vector<Json> vals;   // categorical

  static Hyperparameter from_json(const Json& j) {
    Hyperparameter h;
    if (!j.is_object() || !j.contains("type")) {  // bare value => const
      h.kind = Const;
      h.val = j;
      return h;
    }
    const std::string t = j["type"].as_string();
    if (j.contains("count")) h.count = static_cast<int>(j["count"].as_int());
    if (t == "const") { h.kind = Const; h.val = j["val"]; }
    else if (t == "int") { h.kind = Int; h.minval = j["minval"].as_double(); h.maxval = j["maxval"].as_double(); }
    else if (t == "double") { h.kind = Double; h.minval = j["minval"].as_double(); h.maxval = j["maxval"].as_double(); }
    else if (t == "log") { h.kind = Log; h.minval = j["minval"].as_double(); h.maxval = j["maxval"].as_double(); h.base = j["base"].as_double(10.0); }
    else if (t == "categorical") { h.kind = Categorical; h.vals = j["vals"].arr(); }
    else throw ConfigError("unknown hyperparameter type '" + t + "'");
    return h;
  }
  void validate(const std::string& name, std::vector<std::string>& errs) const {
    if ((kind == Int || kind == Double || kind == Log) && !(maxval > minval || (kind != Int && maxval == minval)))
      if (!(maxval > minval)) errs.push_back("hyperparameter " + name + ": maxval must be > minval");
    if (kind == Log && !(base > 0)) errs.push_back("hyperparameter " + name + ": base must be > 0");
    if (kind == Categorical && vals.empty()) errs.push_back("hyperparameter " + name + ": vals must not be empty");
    if (count != -1 && count < 1) errs.push_back("hyperparameter " + name + ": count must be >= 1");
  }
};

// name -> spec, iterated in sorted-name order (std::map)
using Hyperparameters = std::map<std::string, Hyperparameter>;

inline Hyperparameters hyperparameters_from_json(const Json& j) {
  Hyperparameters h;
  for (auto& kv : j.obj()) h[kv.first] = Hyperparameter::from_json(kv.second);
  return h;
}

constexpr int kMaxAllowedTrials = 2000;

// ---- Searcher config ------------------------------------------------------------------------------------------
struct SearcherConfig {
  std::string name;  // single | random | grid | sync_halving | async_halving | adaptive | adaptive_simple | adaptive_asha | pbt
  std::string metric;
  bool smaller_is_better = true;
  Length max_length, budget, length_per_round;
  int max_trials = 0, num_rungs = 0, max_rungs = 5, max_concurrent_trials = 0, population_size = 0, num_rounds = 0;
  double divisor = 4.0, truncate_fraction = 0.2, resample_probability = 0.2, perturb_factor = 0.2;
  bool train_stragglers = true;
  std::string mode = "standard";
  std::vector<int> bracket_rungs;
  long long source_trial_id = -1;
  std::string source_checkpoint_uuid;

  Unit unit() const { return name == "pbt" ? length_per_round.unit : max_length.unit; }

  static SearcherConfig from_json(const Json& j) {
    SearcherConfig c;
    c.name = j["name"].as_string();
    static const std::set<std::string> known = {"single", "random", "grid", "sync_halving", "async_halving", "adaptive",
                                                "adaptive_simple", "adaptive_asha", "pbt"};
    if (!known.count(c.name)) throw ConfigError("searcher.name must be one of single, random, grid, sync_halving, async_halving, adaptive, adaptive_simple, adaptive_asha, pbt (got '" + c.name + "')");
    static const std::set<std::string> allowed = {
        "name", "metric", "smaller_is_better", "max_length", "budget", "length_per_round", "max_trials", "num_rungs",
        "max_rungs", "max_concurrent_trials", "population_size", "num_rounds", "divisor", "train_stragglers", "mode",
        "bracket_rungs", "source_trial_id", "source_checkpoint_uuid", "replace_function", "explore_function", "max_steps",
        "step_budget", "steps_per_round"};
    for (auto& kv : j.obj())
      if (!allowed.count(kv.first)) throw ConfigError("unknown field in searcher config: '" + kv.first + "'");
    c.metric = j["metric"].as_string();
    if (j.contains("smaller_is_better")) c.smaller_is_better = j["smaller_is_better"].as_bool(true);
    if (j.contains("max_length")) c.max_length = Length::from_json(j["max_length"]);
    if (j.contains("budget")) c.budget = Length::from_json(j["budget"]);
    if (j.contains("length_per_round")) c.length_per_round = Length::from_json(j["length_per_round"]);
    c.max_trials = static_cast<int>(j["max_trials"].as_int(0));
    c.num_rungs = static_cast<int>(j["num_rungs"].as_int(0));
    c.max_rungs = static_cast<int>(j["max_rungs"].as_int(5));
    c.max_concurrent_trials = static_cast<int>(j["max_concurrent_trials"].as_int(0));
    c.population_size = static_cast<int>(j["population_size"].as_int(0));
    c.num_rounds = static_cast<int>(j["num_rounds"].as_int(0));
    c.divisor = j["divisor"].as_double(4.0);
    if (j.contains("train_stragglers")) c.train_stragglers = j["train_stragglers"].as_bool(true);
    c.mode = j["mode"].str_or("standard");
    for (auto& v : j["bracket_rungs"].arr()) c.bracket_rungs.push_back(static_cast<int>(v.as_int()));
    if (j.contains("source_trial_id") && !j["source_trial_id"].is_null()) c.source_trial_id = j["source_trial_id"].as_int();
    c.source_checkpoint_uuid = j["source_checkpoint_uuid"].str_or("");
    if (j.contains("replace_function")) c.truncate_fraction = j["replace_function"]["truncate_fraction"].as_double(0.2);
    if (j.contains("explore_function")) {
      c.resample_probability = j["explore_function"]["resample_probability"].as_double(0.2);
      c.perturb_factor = j["explore_function"]["perturb_factor"].as_double(0.2);
    }
    return c;
  }

  void validate(std::vector<std::string>& errs) const {
    auto need = [&](bool ok, const std::string& msg) { if (!ok) errs.push_back("searcher: " + msg); };
    need(!metric.empty(), "metric must be specified");
    const bool modes_ok = mode == "aggressive" || mode == "standard" || mode == "conservative";
    if (name == "single") need(max_length.units > 0, "max_length must be > 0");
    if (name == "random") { need(max_length.units > 0, "max_length must be > 0"); need(max_trials > 0, "max_trials must be > 0"); }
    if (name == "grid") need(max_length.units > 0, "max_length must be > 0");
    if (name == "sync_halving") {
      need(max_length.units > 0, "max_length must be > 0"); need(budget.units > 0, "budget must be > 0");
      need(divisor > 1.0, "divisor must be > 1.0"); need(num_rungs > 0, "num_rungs must be > 0");
      need(max_length.unit == budget.unit, "max_length and budget must be specified in terms of the same unit");
    }
    if (name == "async_halving") {
      need(max_length.units > 0, "max_length must be > 0"); need(max_trials > 0, "max_trials must be > 0");
      need(divisor > 1.0, "divisor must be > 1.0"); need(num_rungs > 0, "num_rungs must be > 0");
      need(max_concurrent_trials >= 0, "max_concurrent_trials must be >= 0");
    }
    if (name == "adaptive") {
      need(budget.units > max_length.units, "budget must be > max_length"); need(max_length.units > 0, "max_length must be > 0");
      need(budget.units > 0, "budget must be > 0"); need(divisor > 1.0, "divisor must be > 1.0"); need(modes_ok, "invalid adaptive mode");
      need(max_rungs > 0, "max_rungs must be > 0");
      need(max_length.unit == budget.unit, "max_length and budget must be specified in terms of the same unit");
    }
    if (name == "adaptive_simple") {
      need(max_length.units > 0, "max_length must be > 0"); need(max_trials > 0, "max_trials must be > 0");
      need(max_trials <= kMaxAllowedTrials, "max_trials must be <= " + std::to_string(kMaxAllowedTrials));
      need(divisor > 1.0, "divisor must be > 1.0"); need(modes_ok, "invalid adaptive mode"); need(max_rungs > 0, "max_rungs must be > 0");
    }
    if (name == "adaptive_asha") {
      need(max_length.units > 0, "max_length must be > 0"); need(max_trials > 0, "max_trials must be > 0");
      need(divisor > 1.0, "divisor must be > 1.0"); need(modes_ok, "invalid adaptive mode"); need(max_rungs > 0, "max_rungs must be > 0");
      need(max_concurrent_trials >= 0, "max_concurrent_trials must be >= 0");
    }
    if (name == "pbt") {
      need(population_size > 0, "population_size must be > 0"); need(num_rounds > 0, "num_rounds must be > 0");
      need(length_per_round.units > 0, "length_per_round must be > 0");
      need(truncate_fraction >= 0 && truncate_fraction <= 0.5, "truncate_fraction must be in [0, 0.5]");
      need(resample_probability >= 0 && resample_probability <= 1, "resample_probability must be in [0, 1]");
      need(perturb_factor >= 0 && perturb_factor <= 1, "perturb_factor must be in [0, 1]");
    }
  }
};

// ---- Experiment config: defaults, merge, validation ----------------------------------------------------------------
inline void json_merge(Json& base, const Json& over) {
  if (!over.is_object() || !base.is_object()) { base = over.deep_copy(); return; }
  // Tagged unions replace wholesale instead of merging field by field: a length {epochs: 1} must not inherit the
  // default's {batches: 0}, and a storage/searcher block of another "type"/"name" must not keep the old one's fields.
  auto is_length = [](const Json& j) {
    if (!j.is_object() || j.obj().empty()) return false;
    for (auto& kv : j.obj())
      if (kv.first != "records" && kv.first != "batches" && kv.first != "epochs") return false;
    return true;
  };
  auto other_variant = [](const Json& a, const Json& b) {
    for (const char* tag : {"type", "name"})
      if (a.contains(tag) && b.contains(tag) && a[tag].is_string() && b[tag].is_string() && a[tag].as_string() != b[tag].as_string())
        return true;
    return false;
  };
  for (auto& kv : over.obj()) {
    if (kv.second.is_object() && base.contains(kv.first) && base[kv.first].is_object() && !is_length(kv.second) &&
        !other_variant(base[kv.first], kv.second))
      json_merge(base[kv.first], kv.second);
    else
      base[kv.first] = kv.second.deep_copy();
  }
}

inline Json default_experiment_config() {
  return Json::parse(R"({
    "description": "Experiment",
    "labels": [],
    "data": {},
    "checkpoint_storage": {"type": "shared_fs", "host_path": "/tmp/determined-checkpoints", "save_experiment_best": 0, "save_trial_best": 1, "save_trial_latest": 1},
    "checkpoint_policy": "best",
    "perform_initial_validation": false,
    "min_checkpoint_period": {"batches": 0},
    "min_validation_period": {"batches": 0},
    "hyperparameters": {},
    "searcher": {"smaller_is_better": true},
    "resources": {"slots_per_trial": 1, "weight": 1, "native_parallel": false, "priority": 42, "max_slots": null, "shm_size": null, "agent_label": "", "resource_pool": "default"},
    "optimizations": {"aggregation_frequency": 1, "average_aggregated_gradients": true, "average_training_metrics": false, "gradient_compression": false, "mixed_precision": "O0", "tensor_fusion_threshold": 64, "tensor_fusion_cycle_time": 5, "auto_tune_tensor_fusion": false},
    "records_per_epoch": 0,
    "scheduling_unit": 100,
    "bind_mounts": [],
    "environment": {"image": {"cpu": "local", "gpu": "local"}, "environment_variables": {}, "ports": {}, "force_pull_image": false},
    "reproducibility": {},
    "max_restarts": 5,
    "debug": false,
    "data_layer": {"type": "shared_fs"}
  })");
}

inline const std::set<std::string>& experiment_config_keys() {
  static const std::set<std::string> k = {
      "description", "labels", "data", "checkpoint_storage", "tensorboard_storage", "perform_initial_validation",
      "min_checkpoint_period", "min_validation_period", "checkpoint_policy", "hyperparameters", "searcher", "resources",
      "optimizations", "records_per_epoch", "scheduling_unit", "bind_mounts", "environment", "reproducibility",
      "max_restarts", "debug", "internal", "entrypoint", "data_layer", "batches_per_step", "name", "workspace", "project"};
  return k;
}

// defaults <- master checkpoint storage <- template <- user config; unknown top-level keys are rejected.
inline Json resolve_experiment_config(const Json& user, const Json* master_storage = nullptr, const Json* tmpl = nullptr,
                                      unsigned default_seed = 0) {
  if (!user.is_object()) throw ConfigError("experiment config must be a mapping");
  for (auto& kv : user.obj())
    if (!experiment_config_keys().count(kv.first)) throw ConfigError("unknown field in experiment config: '" + kv.first + "'");
  Json cfg = default_experiment_config();
  if (master_storage && master_storage->is_object()) cfg["checkpoint_storage"] = master_storage->deep_copy();
  if (tmpl && tmpl->is_object()) json_merge(cfg, *tmpl);
  // a user-supplied checkpoint_storage of a different type replaces the default wholesale (keeps save_* defaults); one
  // that names no type (only save_* policies, say) is merged into the default location
  if (user.contains("checkpoint_storage") && user["checkpoint_storage"].is_object() &&
      user["checkpoint_storage"].contains("type") &&
      user["checkpoint_storage"]["type"].str_or("") != cfg["checkpoint_storage"]["type"].str_or("")) {
    Json keep = Json::object();
    for (const char* k : {"save_experiment_best", "save_trial_best", "save_trial_latest"}) keep[k] = cfg["checkpoint_storage"][k];
    cfg["checkpoint_storage"] = keep;
  }
  json_merge(cfg, user);
  if (cfg.contains("batches_per_step") && !user.contains("scheduling_unit")) cfg["scheduling_unit"] = cfg["batches_per_step"];
  if (!cfg["reproducibility"].contains("experiment_seed")) cfg["reproducibility"]["experiment_seed"] = static_cast<long long>(default_seed);
  return cfg;
}

inline std::vector<std::string> validate_experiment_config(const Json& cfg) {
  std::vector<std::string> errs;
  auto need = [&](bool ok, const std::string& m) { if (!ok) errs.push_back(m); };
  need(cfg["entrypoint"].is_string() || cfg["internal"]["native"].is_object(), "entrypoint must be specified");
  SearcherConfig sc;
  try {
    sc = SearcherConfig::from_json(cfg["searcher"]);
    sc.validate(errs);
  } catch (const std::exception& e) {
    errs.push_back(e.what());
  }
  Hyperparameters hp;
  try {
    hp = hyperparameters_from_json(cfg["hyperparameters"]);
  } catch (const std::exception& e) {
    errs.push_back(e.what());
  }
  need(hp.count("global_batch_size") > 0, "hyperparameters must include global_batch_size");
  if (hp.count("global_batch_size")) {
    const auto& g = hp["global_batch_size"];
    need(g.kind == Hyperparameter::Const || g.kind == Hyperparameter::Int || g.kind == Hyperparameter::Categorical,
         "global_batch_size must be a const, int or categorical hyperparameter");
    if (g.kind == Hyperparameter::Const) need(g.val.is_number() && g.val.as_int() > 0, "global_batch_size must be a positive number");
  }
  for (auto& kv : hp) kv.second.validate(kv.first, errs);
  if (sc.name == "grid") {
    long long total = 1;
    for (auto& kv : hp) {
      const auto& h = kv.second;
      if (h.kind == Hyperparameter::Int || h.kind == Hyperparameter::Double || h.kind == Hyperparameter::Log) {
        need(h.count >= 1, "grid search requires `count` for hyperparameter " + kv.first);
        long long c = std::max(h.count, 1);
        if (h.kind == Hyperparameter::Int) c = std::min<long long>(c, static_cast<long long>(h.maxval - h.minval) + 1);
        total *= c;
      } else if (h.kind == Hyperparameter::Categorical) {
        total *= static_cast<long long>(h.vals.size());
      }
    }
    need(total <= kMaxAllowedTrials, "number of trials for grid search must be <= " + std::to_string(kMaxAllowedTrials));
  }
  const std::string policy = cfg["checkpoint_policy"].str_or("best");
  need(policy == "best" || policy == "all" || policy == "none", "checkpoint_policy must be one of best, all, none");
  const Json& res = cfg["resources"];
  need(res["slots_per_trial"].as_int(1) >= 0, "resources.slots_per_trial must be >= 0");
  need(res["weight"].as_double(1) >= 0, "resources.weight must be >= 0");
  if (!res["max_slots"].is_null()) need(res["max_slots"].as_int() >= 0, "resources.max_slots must be >= 0");
  if (!res["priority"].is_null()) need(res["priority"].as_int(42) >= 1 && res["priority"].as_int(42) <= 99, "resources.priority must be in [1, 99]");
  const Json& opt = cfg["optimizations"];
  need(opt["aggregation_frequency"].as_int(1) > 0, "optimizations.aggregation_frequency must be > 0");
  const std::string mp = opt["mixed_precision"].str_or("O0");
  need(mp == "O0" || mp == "O1" || mp == "O2" || mp == "O3", "optimizations.mixed_precision must be one of O0..O3");
  need(cfg["scheduling_unit"].as_int(100) > 0, "scheduling_unit must be > 0");
  need(cfg["max_restarts"].as_int(5) >= 0, "max_restarts must be >= 0");
  need(cfg["records_per_epoch"].as_int(0) >= 0, "records_per_epoch must be >= 0");
  for (const char* k : {"min_validation_period", "min_checkpoint_period"}) {
    try {
      Length l = Length::from_json(cfg[k]);
      need(l.units >= 0, std::string(k) + " must be >= 0");
      if (l.unit == Unit::Epochs && l.units > 0) need(cfg["records_per_epoch"].as_int(0) > 0, "must specify records_per_epoch when " + std::string(k) + " is in epochs");
    } catch (const std::exception& e) {
      errs.push_back(std::string(k) + ": " + e.what());
    }
  }
  if (sc.unit() == Unit::Epochs) need(cfg["records_per_epoch"].as_int(0) > 0, "must specify records_per_epoch when the searcher is configured in epochs");
  const std::string st = cfg["checkpoint_storage"]["type"].str_or("");
  need(st == "shared_fs" || st == "s3" || st == "gcs" || st == "hdfs", "checkpoint_storage.type must be one of shared_fs, s3, gcs, hdfs");
  if (st == "shared_fs") need(cfg["checkpoint_storage"]["host_path"].is_string(), "checkpoint_storage.host_path must be specified");
  for (const char* k : {"save_experiment_best", "save_trial_best", "save_trial_latest"})
    need(cfg["checkpoint_storage"][k].as_int(0) >= 0, std::string("checkpoint_storage.") + k + " must be >= 0");
  return errs;
}

// ---- State machines -------------------------------------------------------------------------------------------------
inline bool experiment_transition_ok(const std::string& from, const std::string& to) {
  static const std::map<std::string, std::set<std::string>> t = {
      {"ACTIVE", {"PAUSED", "STOPPING_CANCELED", "STOPPING_COMPLETED", "STOPPING_ERROR"}},
      {"PAUSED", {"ACTIVE", "STOPPING_CANCELED", "STOPPING_COMPLETED", "STOPPING_ERROR"}},
      {"STOPPING_CANCELED", {"CANCELED", "STOPPING_ERROR"}},
      {"STOPPING_COMPLETED", {"COMPLETED", "STOPPING_ERROR"}},
      {"STOPPING_ERROR", {"ERROR"}},
      {"CANCELED", {}}, {"COMPLETED", {}}, {"ERROR", {}}};
  auto it = t.find(from);
  return it != t.end() && it->second.count(to) > 0;
}
inline bool trial_transition_ok(const std::string& from, const std::string& to) {
  static const std::map<std::string, std::set<std::string>> t = {
      {"ACTIVE", {"CANCELED", "COMPLETED", "ERROR"}}, {"CANCELED", {}}, {"COMPLETED", {}}, {"ERROR", {}}};
  auto it = t.find(from);
  return it != t.end() && it->second.count(to) > 0;
}
inline bool checkpoint_transition_ok(const std::string& from, const std::string& to) {
  static const std::map<std::string, std::set<std::string>> t = {
      {"ACTIVE", {"COMPLETED", "ERROR"}}, {"COMPLETED", {"DELETED"}}, {"DELETED", {}}, {"ERROR", {"DELETED"}}};
  auto it = t.find(from);
  return it != t.end() && it->second.count(to) > 0;
}
inline bool is_terminal_state(const std::string& s) { return s == "CANCELED" || s == "COMPLETED" || s == "ERROR"; }
inline bool is_stopping_state(const std::string& s) { return s.rfind("STOPPING_", 0) == 0; }

}  // namespace det
