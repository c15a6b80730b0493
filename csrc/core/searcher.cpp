// searcher.cpp — operations, hyperparameter sampling, single/random/grid/sync_halving/adaptive/adaptive_simple
// search methods and the tournament combinator. (ASHA, adaptive ASHA, PBT, facade, simulator: searcher2.cpp.)
#include "searcher.hpp"
#include "searcher_internal.hpp"

#include <algorithm>
#include <cmath>

namespace det {

// ---- Operation ---------------------------------------------------------------------------------------------
Operation Operation::make_create(NpRand& rng, const Json& hparams) {
  Operation o;
  o.kind = Create;
  o.request_id = rng.uuid4();                                   // 16 bytes from the stream first ...
  o.trial_seed = static_cast<uint32_t>(rng.int64n(1LL << 31));  // ... then the seed
  o.hparams = hparams;
  return o;
}
Operation Operation::make_create_from_checkpoint(NpRand& rng, const Json& hparams, const RequestID& ckpt_of) {
  Operation o = make_create(rng, hparams);
  o.has_checkpoint = true;
  o.checkpoint_request_id = ckpt_of;
  return o;
}
std::string Operation::str() const {
  switch (kind) {
    case Create: return "{Create " + request_id + ", seed " + std::to_string(trial_seed) + (has_checkpoint ? ", checkpoint " + checkpoint_request_id : "") + "}";
    case Train: return "{Train " + request_id + ", " + length.str() + "}";
    case Validate: return "{Validate " + request_id + "}";
    case Checkpoint: return "{Checkpoint " + request_id + "}";
    case Close: return "{Close " + request_id + "}";
    default: return "{Shutdown}";
  }
}
Json Operation::to_json() const {
  Json j = Json::object();
  static const char* names[] = {"create", "train", "validate", "checkpoint", "close", "shutdown"};
  j["kind"] = names[kind];
  if (kind != Shutdown) j["request_id"] = request_id;
  if (kind == Create) {
    j["trial_seed"] = static_cast<long long>(trial_seed);
    j["hparams"] = hparams;
    j["checkpoint"] = has_checkpoint ? Json(checkpoint_request_id) : Json();
  }
  if (kind == Train) j["length"] = length.to_json();
  if (kind == Shutdown) j["failure"] = failure;
  return j;
}
Operation Operation::from_json(const Json& j) {
  Operation o;
  const std::string k = j["kind"].as_string();
  o.kind = k == "create" ? Create : k == "train" ? Train : k == "validate" ? Validate : k == "checkpoint" ? Checkpoint : k == "close" ? Close : Shutdown;
  o.request_id = j["request_id"].str_or("");
  if (o.kind == Create) {
    o.trial_seed = static_cast<uint32_t>(j["trial_seed"].as_int());
    o.hparams = j["hparams"];
    if (j["checkpoint"].is_string()) { o.has_checkpoint = true; o.checkpoint_request_id = j["checkpoint"].as_string(); }
  }
  if (o.kind == Train) o.length = Length::from_json(j["length"]);
  if (o.kind == Shutdown) o.failure = j["failure"].as_bool(false);
  return o;
}

// ---- sampling ------------------------------------------------------------------------------------------------
Json sample_one(const Hyperparameter& h, NpRand& rng) {
  switch (h.kind) {
    case Hyperparameter::Const: return h.val;
    case Hyperparameter::Int: {
      const int lo = static_cast<int>(h.minval), hi = static_cast<int>(h.maxval);
      return Json(lo + rng.intn(hi - lo));  // max exclusive
    }
    case Hyperparameter::Double: return Json(rng.uniform(h.minval, h.maxval));
    case Hyperparameter::Log: return Json(std::pow(h.base, rng.uniform(h.minval, h.maxval)));
    case Hyperparameter::Categorical: return h.vals[static_cast<size_t>(rng.intn(static_cast<int>(h.vals.size())))];
  }
  throw ConfigError("unexpected hyperparameter type");
}
Json sample_all(const Hyperparameters& h, NpRand& rng) {
  Json out = Json::object();
  for (auto& kv : h) out[kv.first] = sample_one(kv.second, rng);  // sorted-name order
  return out;
}
static std::vector<Json> grid_axis(const Hyperparameter& h) {
  std::vector<Json> vals;
  switch (h.kind) {
    case Hyperparameter::Const: vals.push_back(h.val); break;
    case Hyperparameter::Int: {
      const int lo = static_cast<int>(h.minval), hi = static_cast<int>(h.maxval);
      const int count = std::min(std::max(h.count, 1), hi - lo + 1);
      if (count == 1) vals.push_back(Json(static_cast<int>(std::round((lo + hi) / 2.0))));
      else for (int i = 0; i < count; ++i) vals.push_back(Json(static_cast<int>(std::round(lo + static_cast<double>(i) * (hi - lo) / (count - 1)))));
      break;
    }
    case Hyperparameter::Double: {
      const int count = std::max(h.count, 1);
      if (count == 1) vals.push_back(Json((h.minval + h.maxval) / 2.0));
      else for (int i = 0; i < count; ++i) vals.push_back(Json(h.minval + i * (h.maxval - h.minval) / (count - 1)));
      break;
    }
    case Hyperparameter::Log: {
      const int count = std::max(h.count, 1);
      if (count == 1) vals.push_back(Json(std::pow(h.base, (h.minval + h.maxval) / 2.0)));
      else for (int i = 0; i < count; ++i) vals.push_back(Json(std::pow(h.base, h.minval + i * (h.maxval - h.minval) / (count - 1))));
      break;
    }
    case Hyperparameter::Categorical: vals = h.vals; break;
  }
  return vals;
}
std::vector<Json> hyperparameter_grid(const Hyperparameters& h) {
  // Cartesian product with the FIRST (alphabetically) name varying slowest.
  std::vector<std::string> names;
  std::vector<std::vector<Json>> axes;
  for (auto& kv : h) { names.push_back(kv.first); axes.push_back(grid_axis(kv.second)); }
  std::vector<Json> out;
  if (names.empty()) return out;
  std::vector<size_t> idx(names.size(), 0);
  while (true) {
    Json point = Json::object();
    for (size_t i = 0; i < names.size(); ++i) point[names[i]] = axes[i][idx[i]];
    out.push_back(point);
    int d = static_cast<int>(names.size()) - 1;
    while (d >= 0 && ++idx[static_cast<size_t>(d)] == axes[static_cast<size_t>(d)].size()) { idx[static_cast<size_t>(d)] = 0; --d; }
    if (d < 0) break;
  }
  return out;
}

double metric_from(const Json& vm, const std::string& name) {
  const Json& metrics = vm.contains("validation_metrics") ? vm["validation_metrics"] : (vm.contains("metrics") ? vm["metrics"] : vm);
  const Json& v = metrics[name];
  if (!v.is_number()) throw std::runtime_error("selected metric " + name + " is not a scalar value: " + v.dump());
  return v.as_double();
}

// ---- single / random / grid -------------------------------------------------------------------------------------
namespace {

class RandomSearch : public SearchMethod {
 public:
  RandomSearch(int max_trials, Length max_length) : max_trials_(max_trials), max_length_(max_length) {}
  Ops initial_operations(SearchContext ctx) override {
    Ops ops;
    for (int t = 0; t < max_trials_; ++t) {
      Operation c = Operation::make_create(*ctx.rand, sample_all(*ctx.hparams, *ctx.rand));
      ops.push_back(c);
      ops.push_back(Operation::make_train(c.request_id, max_length_));
      ops.push_back(Operation::make_validate(c.request_id));
      ops.push_back(Operation::make_close(c.request_id));
    }
    return ops;
  }
  double progress(double units) override { return units / static_cast<double>(max_length_.mult(max_trials_).units); }
  Ops trial_exited_early(SearchContext, const RequestID&, ExitedReason) override { return {}; }
  Unit unit() const override { return max_length_.unit; }

 private:
  int max_trials_;
  Length max_length_;
};

class GridSearch : public SearchMethod {
 public:
  explicit GridSearch(Length max_length) : max_length_(max_length) {}
  Ops initial_operations(SearchContext ctx) override {
    Ops ops;
    auto grid = hyperparameter_grid(*ctx.hparams);
    trials_ = static_cast<int>(grid.size());
    for (auto& point : grid) {
      Operation c = Operation::make_create(*ctx.rand, point);
      ops.push_back(c);
      ops.push_back(Operation::make_train(c.request_id, max_length_));
      ops.push_back(Operation::make_validate(c.request_id));
      ops.push_back(Operation::make_close(c.request_id));
    }
    return ops;
  }
  double progress(double units) override { return units / static_cast<double>(max_length_.mult(trials_).units); }
  Ops trial_exited_early(SearchContext, const RequestID&, ExitedReason) override { return {}; }
  Unit unit() const override { return max_length_.unit; }

 private:
  Length max_length_;
  int trials_ = 0;
};

}  // namespace


namespace {

class SyncHalvingSearch : public SearchMethod {
 public:
  // regular constructor: trials per rung scaled to the budget
  explicit SyncHalvingSearch(const SearcherConfig& c) : cfg_(c) {
    int expected = 0;
    for (int id = 0; id < c.num_rungs; ++id) {
      const double compound = std::pow(c.divisor, static_cast<double>(c.num_rungs - id - 1));
      Rung r;
      r.units_needed = Length{c.unit(), std::max(static_cast<int>(static_cast<double>(c.max_length.units) / compound), 1)};
      r.start_trials = std::max(static_cast<int>(compound), 1);
      if (id == 0) expected += r.units_needed.units * r.start_trials;
      else expected += (r.units_needed.units - rungs_[static_cast<size_t>(id - 1)].units_needed.units) * r.start_trials;
      rungs_.push_back(r);
    }
    const double multiplier = static_cast<double>(c.budget.units) / static_cast<double>(expected);
    expected = 0;
    for (int id = 0; id < c.num_rungs; ++id) {
      Rung& cur = rungs_[static_cast<size_t>(id)];
      cur.start_trials = static_cast<int>(multiplier * static_cast<double>(cur.start_trials));
      if (id == 0) {
        expected += cur.units_needed.units * cur.start_trials;
      } else {
        Rung& prev = rungs_[static_cast<size_t>(id - 1)];
        cur.units_needed.units = std::max(cur.units_needed.units, prev.units_needed.units);
        cur.start_trials = std::max(std::min(cur.start_trials, prev.start_trials), 1);
        prev.promote_trials = cur.start_trials;
        expected += (cur.units_needed.units - prev.units_needed.units) * cur.start_trials;
      }
    }
    expected_units_ = expected;
  }
  // "simple" constructor: a fixed number of starting trials (adaptive_simple)
  SyncHalvingSearch(const SearcherConfig& c, int trials) : cfg_(c) {
    int expected = 0;
    for (int id = 0; id < c.num_rungs; ++id) {
      int units = std::max(static_cast<int>(static_cast<double>(c.max_length.units) / std::pow(c.divisor, static_cast<double>(c.num_rungs - id - 1))), 1);
      int start = std::max(static_cast<int>(static_cast<double>(trials) / std::pow(c.divisor, static_cast<double>(id))), 1);
      if (id != 0) {
        Rung& prev = rungs_[static_cast<size_t>(id - 1)];
        units = std::max(units, prev.units_needed.units);
        start = std::max(start, prev.promote_trials);
        prev.promote_trials = start;
        expected += (units - prev.units_needed.units) * start;
      } else {
        expected += units * start;
      }
      Rung r;
      r.units_needed = Length{c.unit(), units};
      r.start_trials = start;
      rungs_.push_back(r);
    }
    expected_units_ = expected;
  }

  Ops initial_operations(SearchContext ctx) override {
    Ops ops;
    for (int t = 0; t < rungs_[0].start_trials; ++t) {
      Operation c = Operation::make_create(*ctx.rand, sample_all(*ctx.hparams, *ctx.rand));
      ops.push_back(c);
      ops.push_back(Operation::make_train(c.request_id, rungs_[0].units_needed));
      ops.push_back(Operation::make_validate(c.request_id));
    }
    return ops;
  }
  Ops validation_completed(SearchContext ctx, const RequestID& id, const Operation&, const Json& metrics) override {
    double m = metric_from(metrics, cfg_.metric);
    if (!cfg_.smaller_is_better) m = -m;
    return promote(ctx, id, m);
  }
  double progress(double units) override { return std::min(1.0, units / static_cast<double>(expected_units_)); }
  Ops trial_exited_early(SearchContext ctx, const RequestID& id, ExitedReason) override {
    early_exit_.insert(id);
    return promote(ctx, id, kExitedMetricValue);
  }
  Unit unit() const override { return cfg_.unit(); }

 private:
  Ops promote(SearchContext ctx, const RequestID& id, double metric) {
    const int ri = trial_rungs_[id];
    Rung& rung = rungs_[static_cast<size_t>(ri)];
    if (ri == cfg_.num_rungs - 1) {
      ++trials_completed_;
      if (!early_exit_.count(id)) return {Operation::make_close(id)};
      return {};
    }
    Ops ops;
    auto to_promote = rung.promotions_sync(id, metric);
    if (!to_promote.empty()) {
      for (auto& pid : to_promote) {
        trial_rungs_[pid] = ri + 1;
        if (!early_exit_.count(pid)) {
          const int units = std::max(rungs_[static_cast<size_t>(ri + 1)].units_needed.units - rung.units_needed.units, 1);
          ops.push_back(Operation::make_train(pid, Length{unit(), units}));
          ops.push_back(Operation::make_validate(pid));
        } else {
          return promote(ctx, pid, kExitedMetricValue);
        }
      }
      if (rung.start_trials < static_cast<int>(rung.metrics.size()))
        throw std::runtime_error("number of trials exceeded initial trials for rung");
      if (static_cast<int>(rung.metrics.size()) == rung.start_trials) {
        for (size_t i = static_cast<size_t>(rung.promote_trials); i < rung.metrics.size(); ++i) {
          ++trials_completed_;
          if (!early_exit_.count(rung.metrics[i].request_id)) ops.push_back(Operation::make_close(rung.metrics[i].request_id));
        }
      }
    }
    return ops;
  }
  SearcherConfig cfg_;
  std::vector<Rung> rungs_;
  std::map<RequestID, int> trial_rungs_;
  std::set<RequestID> early_exit_;
  int trials_completed_ = 0;
  int expected_units_ = 0;
};

// ---- tournament -------------------------------------------------------------------------------------------------------
class TournamentSearch : public SearchMethod {
 public:
  explicit TournamentSearch(std::vector<std::unique_ptr<SearchMethod>> subs) : subs_(std::move(subs)), units_(subs_.size(), 0.0) {}
  Ops initial_operations(SearchContext ctx) override {
    Ops all;
    for (size_t i = 0; i < subs_.size(); ++i) {
      Ops ops = subs_[i]->initial_operations(ctx);
      mark(i, ops);
      all.insert(all.end(), ops.begin(), ops.end());
    }
    return all;
  }
  Ops trial_created(SearchContext ctx, const RequestID& id) override { size_t i = table_.at(id); return mark(i, subs_[i]->trial_created(ctx, id)); }
  Ops train_completed(SearchContext ctx, const RequestID& id, const Operation& op) override {
    size_t i = table_.at(id);
    units_[i] += static_cast<double>(op.length.units);
    return mark(i, subs_[i]->train_completed(ctx, id, op));
  }
  Ops checkpoint_completed(SearchContext ctx, const RequestID& id, const Operation& op, const Json& m) override {
    size_t i = table_.at(id);
    return mark(i, subs_[i]->checkpoint_completed(ctx, id, op, m));
  }
  Ops validation_completed(SearchContext ctx, const RequestID& id, const Operation& op, const Json& m) override {
    size_t i = table_.at(id);
    return mark(i, subs_[i]->validation_completed(ctx, id, op, m));
  }
  Ops trial_closed(SearchContext ctx, const RequestID& id) override { size_t i = table_.at(id); return mark(i, subs_[i]->trial_closed(ctx, id)); }
  Ops trial_exited_early(SearchContext ctx, const RequestID& id, ExitedReason r) override {
    size_t i = table_.at(id);
    return mark(i, subs_[i]->trial_exited_early(ctx, id, r));
  }
  double progress(double) override {
    double sum = 0;
    for (size_t i = 0; i < subs_.size(); ++i) sum += subs_[i]->progress(units_[i]);
    return sum / static_cast<double>(subs_.size());
  }
  Unit unit() const override { return subs_[0]->unit(); }

 private:
  Ops mark(size_t i, Ops ops) {
    for (auto& o : ops)
      if (o.kind == Operation::Create) table_[o.request_id] = i;
    return ops;
  }
  std::vector<std::unique_ptr<SearchMethod>> subs_;
  std::vector<double> units_;
  std::map<RequestID, size_t> table_;
};

std::vector<int> bracket_rungs_for_mode(const std::string& mode, int max_rungs) {
  std::vector<int> b;
  if (mode == "conservative") for (int i = 1; i <= max_rungs; ++i) b.push_back(i);
  else if (mode == "standard") for (int i = (max_rungs - 1) / 2 + 1; i <= max_rungs; ++i) b.push_back(i);
  else if (mode == "aggressive") b.push_back(max_rungs);
  else throw ConfigError("unexpected adaptive mode: " + mode);
  return b;
}

}  // namespace

std::unique_ptr<SearchMethod> new_async_halving(const SearcherConfig& c);                       // searcher2.cpp
std::unique_ptr<SearchMethod> new_pbt(const SearcherConfig& c);                                 // searcher2.cpp
std::unique_ptr<SearchMethod> new_tournament(std::vector<std::unique_ptr<SearchMethod>> subs) {
  return std::unique_ptr<SearchMethod>(new TournamentSearch(std::move(subs)));
}
std::vector<int> adaptive_brackets(const std::string& mode, int max_rungs) { return bracket_rungs_for_mode(mode, max_rungs); }
std::unique_ptr<SearchMethod> new_adaptive_asha(const SearcherConfig& c);                       // searcher2.cpp

std::unique_ptr<SearchMethod> new_search_method(const SearcherConfig& c) {
  if (c.name == "single") return std::unique_ptr<SearchMethod>(new RandomSearch(1, c.max_length));
  if (c.name == "random") return std::unique_ptr<SearchMethod>(new RandomSearch(c.max_trials, c.max_length));
  if (c.name == "grid") return std::unique_ptr<SearchMethod>(new GridSearch(c.max_length));
  if (c.name == "sync_halving") return std::unique_ptr<SearchMethod>(new SyncHalvingSearch(c));
  if (c.name == "adaptive") {
    std::vector<int> brackets = c.bracket_rungs.empty() ? bracket_rungs_for_mode(c.mode, c.max_rungs) : c.bracket_rungs;
    std::sort(brackets.rbegin(), brackets.rend());
    std::vector<std::unique_ptr<SearchMethod>> subs;
    for (int rungs : brackets) {
      SearcherConfig s = c;
      s.name = "sync_halving";
      s.num_rungs = rungs;
      s.budget = c.budget.div(static_cast<int>(brackets.size()));
      subs.emplace_back(new SyncHalvingSearch(s));
    }
    return new_tournament(std::move(subs));
  }
  if (c.name == "adaptive_simple") {
    std::vector<int> brackets = bracket_rungs_for_mode(c.mode, c.max_rungs);
    std::sort(brackets.rbegin(), brackets.rend());
    std::vector<std::unique_ptr<SearchMethod>> subs;
    const int nb = static_cast<int>(brackets.size());
    for (int i = 0; i < nb; ++i) {
      SearcherConfig s = c;
      s.name = "sync_halving";
      s.num_rungs = brackets[static_cast<size_t>(i)];
      s.train_stragglers = true;
      int count = c.max_trials / nb + (i < c.max_trials % nb ? 1 : 0);
      subs.emplace_back(new SyncHalvingSearch(s, std::max(count, 1)));
    }
    return new_tournament(std::move(subs));
  }
  if (c.name == "async_halving") return new_async_halving(c);
  if (c.name == "adaptive_asha") return new_adaptive_asha(c);
  if (c.name == "pbt") return new_pbt(c);
  throw ConfigError("no searcher type specified");
}

}  // namespace det
