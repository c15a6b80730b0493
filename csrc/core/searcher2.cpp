// searcher2.cpp — async_halving (ASHA), adaptive_asha, pbt, the Searcher facade + event log, the simulator.
#include <algorithm>
#include <cmath>

#include "searcher.hpp"
#include "searcher_internal.hpp"

namespace det {

namespace {

// ---- ASHA ------------------------------------------------------------------------------------------------------
class AsyncHalvingSearch : public SearchMethod {
 public:
  explicit AsyncHalvingSearch(const SearcherConfig& c) : cfg_(c), max_trials_(c.max_trials) {
    for (int id = 0; id < c.num_rungs; ++id) {
      const double rate = std::pow(c.divisor, static_cast<double>(c.num_rungs - id - 1));
      Rung r;
      r.units_needed = Length{c.unit(), std::max(static_cast<int>(static_cast<double>(c.max_length.units) / rate), 1)};
      rungs_.push_back(r);
    }
  }
  Ops initial_operations(SearchContext ctx) override {
    int conc;
    if (cfg_.max_concurrent_trials > 0) conc = std::min(cfg_.max_concurrent_trials, cfg_.max_trials);
    else conc = std::max(std::min(static_cast<int>(std::pow(cfg_.divisor, static_cast<double>(cfg_.num_rungs - 1))), cfg_.max_trials), 1);
    Ops ops;
    for (int t = 0; t < conc; ++t) append_new_trial(ctx, ops);
    return ops;
  }
  Ops trial_created(SearchContext, const RequestID& id) override {
    ++rungs_[0].outstanding_trials;
    trial_rungs_[id] = 0;
    return {};
  }
  Ops trial_closed(SearchContext, const RequestID& id) override {
    ++trials_completed_;
    closed_.insert(id);
    return {};
  }
  Ops validation_completed(SearchContext ctx, const RequestID& id, const Operation&, const Json& metrics) override {
    double m = metric_from(metrics, cfg_.metric);
    if (!cfg_.smaller_is_better) m = -m;
    return promote(ctx, id, m);
  }
  double progress(double) override {
    const int all = static_cast<int>(rungs_[0].metrics.size());
    double p = static_cast<double>(all) / (1.2 * static_cast<double>(max_trials_));
    if (all == max_trials_) p = std::max(static_cast<double>(trials_completed_) / static_cast<double>(max_trials_), p);
    return p;
  }
  Ops trial_exited_early(SearchContext ctx, const RequestID& id, ExitedReason) override {
    early_exit_.insert(id);
    closed_.insert(id);
    return promote(ctx, id, kExitedMetricValue);
  }
  Unit unit() const override { return cfg_.unit(); }

 private:
  void append_new_trial(SearchContext ctx, Ops& ops) {
    Operation c = Operation::make_create(*ctx.rand, sample_all(*ctx.hparams, *ctx.rand));
    trial_rungs_[c.request_id] = 0;
    ops.push_back(c);
    ops.push_back(Operation::make_train(c.request_id, rungs_[0].units_needed));
    ops.push_back(Operation::make_validate(c.request_id));
  }
  Ops promote(SearchContext ctx, const RequestID& id, double metric) {
    const int ri = trial_rungs_[id];
    Rung& rung = rungs_[static_cast<size_t>(ri)];
    --rung.outstanding_trials;
    bool added_train = false;
    Ops ops;
    if (ri == cfg_.num_rungs - 1) {
      rung.metrics.push_back(TrialMetric{id, metric, false});
      if (!early_exit_.count(id)) {
        ops.push_back(Operation::make_close(id));
        closed_.insert(id);
      }
    } else {
      Rung& next = rungs_[static_cast<size_t>(ri + 1)];
      for (auto& pid : rung.promotions_async(id, metric, cfg_.divisor)) {
        trial_rungs_[pid] = ri + 1;
        ++next.outstanding_trials;
        if (!early_exit_.count(pid)) {
          const int units = std::max(next.units_needed.units - rung.units_needed.units, 1);
          ops.push_back(Operation::make_train(pid, Length{unit(), units}));
          ops.push_back(Operation::make_validate(pid));
          added_train = true;
        } else {
          return promote(ctx, pid, kExitedMetricValue);
        }
      }
    }
    if (!added_train && static_cast<int>(trial_rungs_.size()) < max_trials_) append_new_trial(ctx, ops);
    if (static_cast<int>(rungs_[0].metrics.size()) == max_trials_) {
      Ops extra = close_out_rungs();
      ops.insert(ops.end(), extra.begin(), extra.end());
    }
    return ops;
  }
  Ops close_out_rungs() {
    Ops ops;
    for (auto& rung : rungs_) {
      if (rung.outstanding_trials > 0) break;
      for (auto& tm : rung.metrics) {
        if (!tm.promoted && !closed_.count(tm.request_id) && !early_exit_.count(tm.request_id)) {
          ops.push_back(Operation::make_close(tm.request_id));
          closed_.insert(tm.request_id);
        }
      }
    }
    return ops;
  }
  SearcherConfig cfg_;
  std::vector<Rung> rungs_;
  std::map<RequestID, int> trial_rungs_;
  std::set<RequestID> early_exit_, closed_;
  int max_trials_;
  int trials_completed_ = 0;
};

// ---- PBT --------------------------------------------------------------------------------------------------------
class PBTSearch : public SearchMethod {
 public:
  explicit PBTSearch(const SearcherConfig& c) : cfg_(c) {}
  Ops initial_operations(SearchContext ctx) override {
    Ops ops;
    for (int t = 0; t < cfg_.population_size; ++t) {
      Operation c = Operation::make_create(*ctx.rand, sample_all(*ctx.hparams, *ctx.rand));
      trial_params_[c.request_id] = c.hparams;
      ops.push_back(c);
      ops.push_back(Operation::make_train(c.request_id, cfg_.length_per_round));
      ops.push_back(Operation::make_validate(c.request_id));
    }
    return ops;
  }
  Ops validation_completed(SearchContext ctx, const RequestID& id, const Operation&, const Json& metrics) override {
    const double m = metric_from(metrics, cfg_.metric);
    metrics_[id] = cfg_.smaller_is_better ? m : -m;
    return run_new_trials(ctx, id);
  }
  Ops checkpoint_completed(SearchContext, const RequestID& id, const Operation&, const Json&) override {
    Ops ops;
    auto it = waiting_.find(id);
    if (it != waiting_.end()) { ops = it->second; waiting_.erase(it); }
    return ops;
  }
  double progress(double units) override {
    return units / static_cast<double>(cfg_.length_per_round.mult(cfg_.population_size).mult(cfg_.num_rounds).units);
  }
  Ops trial_exited_early(SearchContext ctx, const RequestID& id, ExitedReason) override {
    early_exit_.insert(id);
    metrics_[id] = kExitedMetricValue;
    return run_new_trials(ctx, id);
  }
  Unit unit() const override { return cfg_.length_per_round.unit; }

 private:
  Ops run_new_trials(SearchContext ctx, const RequestID& id) {
    Ops ops;
    ++trial_rounds_[id];
    if (static_cast<int>(metrics_.size()) < cfg_.population_size) return ops;
    ++rounds_completed_;
    if (rounds_completed_ >= cfg_.num_rounds) {
      for (auto& kv : metrics_)
        if (!early_exit_.count(kv.first)) ops.push_back(Operation::make_close(kv.first));
      return ops;
    }
    const int num_truncate = static_cast<int>(cfg_.truncate_fraction * static_cast<double>(cfg_.population_size));
    std::vector<RequestID> ids;
    for (auto& kv : metrics_) ids.push_back(kv.first);
    std::sort(ids.begin(), ids.end(), [&](const RequestID& a, const RequestID& b) {
      const double ma = metrics_[a], mb = metrics_[b];
      if (ma != mb) return ma < mb;
      return a < b;
    });
    metrics_.clear();
    for (size_t i = ids.size() - static_cast<size_t>(num_truncate); i < ids.size(); ++i)
      if (!early_exit_.count(ids[i])) ops.push_back(Operation::make_close(ids[i]));
    for (int i = 0; i < num_truncate; ++i) {
      const RequestID& rid = ids[static_cast<size_t>(i)];
      if (early_exit_.count(rid)) continue;
      Operation ck = Operation::make_checkpoint(rid);
      ops.push_back(ck);
      Json new_params = explore(ctx, trial_params_[rid]);
      Operation c = Operation::make_create_from_checkpoint(*ctx.rand, new_params, rid);
      trial_params_[c.request_id] = new_params;
      waiting_[rid] = {c, Operation::make_train(c.request_id, cfg_.length_per_round), Operation::make_validate(c.request_id)};
    }
    for (size_t i = 0; i < ids.size() - static_cast<size_t>(num_truncate); ++i) {
      if (!early_exit_.count(ids[i])) {
        ops.push_back(Operation::make_train(ids[i], cfg_.length_per_round));
        ops.push_back(Operation::make_validate(ids[i]));
      } else {
        metrics_[ids[i]] = kExitedMetricValue;
      }
    }
    return ops;
  }
  Json explore(SearchContext ctx, const Json& old) {
    Json params = Json::object();
    for (auto& kv : *ctx.hparams) {
      const Hyperparameter& h = kv.second;
      if (ctx.rand->unit_interval() < cfg_.resample_probability) {
        params[kv.first] = sample_one(h, *ctx.rand);
        continue;
      }
      Json val = old[kv.first];
      const bool decrease = ctx.rand->unit_interval() < 0.5;
      const double mult = decrease ? 1 - cfg_.perturb_factor : 1 + cfg_.perturb_factor;
      auto clampd = [](double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); };
      if (h.kind == Hyperparameter::Int) {
        const double scaled = static_cast<double>(val.as_int()) * mult;
        const int iv = static_cast<int>(decrease ? std::floor(scaled) : std::ceil(scaled));
        val = Json(static_cast<int>(clampd(iv, h.minval, h.maxval)));
      } else if (h.kind == Hyperparameter::Double) {
        val = Json(clampd(val.as_double() * mult, h.minval, h.maxval));
      } else if (h.kind == Hyperparameter::Log) {
        val = Json(clampd(val.as_double() * mult, std::pow(h.base, h.minval), std::pow(h.base, h.maxval)));
      }
      params[kv.first] = val;
    }
    return params;
  }
  SearcherConfig cfg_;
  int rounds_completed_ = 0;
  std::map<RequestID, double> metrics_;
  std::map<RequestID, int> trial_rounds_;
  std::map<RequestID, Json> trial_params_;
  std::map<RequestID, Ops> waiting_;  // keyed by the trial whose checkpoint is awaited
  std::set<RequestID> early_exit_;
};

}  // namespace

std::unique_ptr<SearchMethod> new_async_halving(const SearcherConfig& c) { return std::unique_ptr<SearchMethod>(new AsyncHalvingSearch(c)); }
std::unique_ptr<SearchMethod> new_pbt(const SearcherConfig& c) { return std::unique_ptr<SearchMethod>(new PBTSearch(c)); }

std::unique_ptr<SearchMethod> new_adaptive_asha(const SearcherConfig& cfg) {
  SearcherConfig c = cfg;
  std::vector<int> brackets = c.bracket_rungs;
  if (brackets.empty()) {
    c.max_rungs = std::min(c.max_rungs, static_cast<int>(std::log(static_cast<double>(c.max_length.units)) / std::log(c.divisor)) + 1);
    c.max_rungs = std::min(c.max_rungs, static_cast<int>(std::log(static_cast<double>(c.max_trials)) / std::log(c.divisor)) + 1);
    brackets = adaptive_brackets(c.mode, c.max_rungs);
  }
  std::sort(brackets.rbegin(), brackets.rend());
  const size_t nb = brackets.size();
  // trials per bracket, weighted divisor^(r-1)/r
  std::vector<double> weight(nb);
  double total = 0;
  for (size_t i = 0; i < nb; ++i) { weight[i] = std::pow(c.divisor, static_cast<double>(brackets[i] - 1)) / static_cast<double>(brackets[i]); total += weight[i]; }
  std::vector<int> trials(nb);
  int allocated = 0;
  for (size_t i = 0; i < nb; ++i) { trials[i] = std::max(static_cast<int>(weight[i] / total * static_cast<double>(c.max_trials)), 1); allocated += trials[i]; }
  trials[0] += std::max(c.max_trials - allocated, 0);
  // concurrency per bracket
  std::vector<int> conc(nb);
  int min_trials, remainder = 0;
  if (c.max_concurrent_trials == 0) {
    min_trials = std::max(trials[nb - 1], static_cast<int>(c.divisor));
  } else {
    const int mct = std::max(c.max_concurrent_trials, static_cast<int>(nb));
    min_trials = mct / static_cast<int>(nb);
    remainder = mct % static_cast<int>(nb);
  }
  for (size_t i = 0; i < nb; ++i) conc[i] = min_trials + (static_cast<int>(i) < remainder ? 1 : 0);
  std::vector<std::unique_ptr<SearchMethod>> subs;
  for (size_t i = 0; i < nb; ++i) {
    SearcherConfig s = c;
    s.name = "async_halving";
    s.num_rungs = brackets[i];
    s.max_trials = trials[i];
    s.max_concurrent_trials = conc[i];
    subs.push_back(new_async_halving(s));
  }
  return new_tournament(std::move(subs));
}

// ---- facade -----------------------------------------------------------------------------------------------------------
Searcher::Searcher(uint32_t seed, std::unique_ptr<SearchMethod> method, Hyperparameters hparams)
    : rand_(seed), hparams_(std::move(hparams)), method_(std::move(method)) {}

void Searcher::ops_created(const Ops& ops) {
  for (auto& o : ops) {
    if (o.kind == Operation::Create) ++trials_requested_;
    if (o.kind == Operation::Shutdown) shutdown_ = true;
  }
}
Ops Searcher::initial_operations() {
  Ops ops = method_->initial_operations(ctx());
  ops_created(ops);
  return ops;
}
Ops Searcher::trial_created(const Operation& create, int trial_id) {
  Json ev = Json::object();
  ev["type"] = "trial_created";
  ev["create"] = create.to_json();
  ev["trial_id"] = trial_id;
  uncommitted_.push_back(ev);
  trial_ids_[create.request_id] = trial_id;
  request_ids_[trial_id] = create.request_id;
  Ops ops = method_->trial_created(ctx(), create.request_id);
  ops_created(ops);
  return ops;
}
Ops Searcher::trial_exited_early(int trial_id, ExitedReason reason) {
  auto it = request_ids_.find(trial_id);
  if (it == request_ids_.end()) throw std::runtime_error("unexpected trial ID sent to searcher: " + std::to_string(trial_id));
  early_exits_.insert(it->second);
  Ops ops = method_->trial_exited_early(ctx(), it->second, reason);
  ops_created(ops);
  return ops;
}
void Searcher::workload_completed(const Json& msg, double units) {
  total_units_completed_ += units;
  Json ev = Json::object();
  ev["type"] = "workload_completed";
  ev["message"] = msg;
  ev["units"] = units;
  uncommitted_.push_back(ev);
}
Ops Searcher::operation_completed(int trial_id, const Operation& op, const Json& metrics) {
  auto it = request_ids_.find(trial_id);
  if (it == request_ids_.end()) throw std::runtime_error("unexpected trial ID sent to searcher: " + std::to_string(trial_id));
  Ops ops;
  switch (op.kind) {
    case Operation::Train: ops = method_->train_completed(ctx(), it->second, op); break;
    case Operation::Checkpoint: ops = method_->checkpoint_completed(ctx(), it->second, op, metrics); break;
    case Operation::Validate: ops = method_->validation_completed(ctx(), it->second, op, metrics); break;
    default: throw std::runtime_error("unexpected op: " + op.str());
  }
  ops_created(ops);
  return ops;
}
Ops Searcher::trial_closed(const RequestID& id) {
  Json ev = Json::object();
  ev["type"] = "trial_closed";
  ev["request_id"] = id;
  uncommitted_.push_back(ev);
  ++trials_closed_;
  Ops ops = method_->trial_closed(ctx(), id);
  ops_created(ops);
  if (trials_requested_ == trials_closed_) {
    Operation sd = Operation::make_shutdown(static_cast<int>(early_exits_.size()) >= trials_requested_);
    ops_created({sd});
    ops.push_back(sd);
  }
  return ops;
}
double Searcher::progress() {
  const double p = method_->progress(total_units_completed_);
  return (std::isnan(p) || std::isinf(p)) ? 0.0 : p;
}
bool Searcher::trial_id_of(const RequestID& id, int* out) const {
  auto it = trial_ids_.find(id);
  if (it == trial_ids_.end()) return false;
  *out = it->second;
  return true;
}
bool Searcher::request_id_of(int trial_id, RequestID* out) const {
  auto it = request_ids_.find(trial_id);
  if (it == request_ids_.end()) return false;
  *out = it->second;
  return true;
}
std::vector<Json> Searcher::uncommitted_events() {
  std::vector<Json> out;
  out.swap(uncommitted_);
  return out;
}

// ---- simulator ----------------------------------------------------------------------------------------------------------
Json SimulationResult::summary() const {
  std::map<std::string, int> counts;
  for (auto& kv : results) {
    std::string key;
    for (auto& op : kv.second) {
      if (!key.empty()) key += " ";
      if (op.kind == Operation::Train) key += std::to_string(op.length.units) + (op.length.unit == Unit::Records ? "R" : op.length.unit == Unit::Batches ? "B" : "E");
      else if (op.kind == Operation::Validate) key += "V";
      else if (op.kind == Operation::Checkpoint) key += "C";
    }
    ++counts[key];
  }
  Json j = Json::object();
  for (auto& kv : counts) j[kv.first] = kv.second;
  return j;
}

static bool handle_ops(std::map<RequestID, Ops>& pending, std::vector<RequestID>& order, const Ops& ops) {
  for (auto& op : ops) {
    if (op.kind == Operation::Create) { order.push_back(op.request_id); pending[op.request_id] = {op}; }
    else if (op.kind == Operation::Shutdown) return true;
    else pending[op.request_id].push_back(op);
  }
  return false;
}

SimulationResult simulate(Searcher& s, long long seed, const ValidationFunction& val, bool random_order, const std::string& metric_name) {
  SimulationResult sim;
  sim.seed = seed;
  std::mt19937_64 rng(static_cast<uint64_t>(seed));
  std::map<RequestID, Ops> pending;
  std::map<RequestID, int> trial_ids, op_idx;
  std::vector<RequestID> order;
  Ops ops = s.initial_operations();
  double last = s.progress();
  if (last != 0.0) throw std::runtime_error("initial searcher progress started at " + std::to_string(last));
  bool shutdown = handle_ops(pending, order, ops);
  int next_trial = 1;
  while (!shutdown) {
    RequestID pick;
    bool found = false;
    if (!random_order) {
      for (auto& id : order) {
        auto it = pending.find(id);
        if (it != pending.end() && !it->second.empty()) { pick = id; found = true; break; }
      }
    } else {
      std::vector<RequestID> cands;
      for (auto& kv : pending) if (!kv.second.empty()) cands.push_back(kv.first);  // map iteration = sorted ids
      if (!cands.empty()) { pick = cands[rng() % cands.size()]; found = true; }
    }
    if (!found) throw std::runtime_error("tried to pick a trial when no trial had pending operations");
    Operation op = pending[pick].front();
    pending[pick].erase(pending[pick].begin());
    if (op.kind == Operation::Create) {
      sim.results[pick] = {};
      trial_ids[pick] = next_trial;
      Ops r = s.trial_created(op, next_trial);
      op_idx[pick] = 0;
      shutdown = handle_ops(pending, order, r);
      ++next_trial;
    } else if (op.runnable()) {
      Json metrics = Json::object();
      if (op.kind == Operation::Validate) {
        Json vm = Json::object();
        vm[metric_name] = val(rng, trial_ids[pick], op_idx[pick]);
        metrics["num_inputs"] = 1;
        metrics["validation_metrics"] = vm;
      } else if (op.kind == Operation::Checkpoint) {
        metrics["uuid"] = "00000000-0000-4000-8000-" + std::to_string(100000000000ULL + (rng() % 899999999999ULL));
        metrics["resources"] = Json::object();
      }
      sim.results[pick].push_back(op);
      if (op.kind == Operation::Train) {
        Json msg = Json::object();
        msg["kind"] = "RUN_STEP";
        msg["trial_id"] = trial_ids[pick];
        msg["step_id"] = op_idx[pick];
        s.workload_completed(msg, static_cast<double>(op.length.units));
      }
      Ops r = s.operation_completed(trial_ids[pick], op, metrics);
      ++op_idx[pick];
      shutdown = handle_ops(pending, order, r);
    } else if (op.kind == Operation::Close) {
      pending.erase(pick);
      Ops r = s.trial_closed(pick);
      shutdown = handle_ops(pending, order, r);
    } else {
      throw std::runtime_error("unexpected searcher operation " + op.str());
    }
    if (shutdown) {
      for (auto& kv : pending)
        if (!kv.second.empty()) throw std::runtime_error("searcher shutdown prematurely");
      break;
    }
    const double p = s.progress();
    if (p < last) throw std::runtime_error("searcher progress dropped from " + std::to_string(last) + " to " + std::to_string(p));
    last = p;
  }
  last = s.progress();
  if (std::fabs(last - 1.0) > 1e-9) throw std::runtime_error("searcher progress did not end at 100%: " + std::to_string(last * 100) + "%");
  if (sim.results.size() != order.size()) throw std::runtime_error("more trials created than completed");
  return sim;
}

}  // namespace det
