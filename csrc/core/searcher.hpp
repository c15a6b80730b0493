// searcher.hpp — hyperparameter search: operations, the SearchMethod interface, the nine search methods,
// the tournament combinator, the Searcher facade with its event log, and the simulator.
// Parity: reference master/pkg/searcher/*.go (operations.go:20-227, hyperparameters.go:24-74, random.go,
// grid.go, sha.go, asha.go, adaptive*.go, pbt.go, tournament.go, searcher.go:13-162, event_log.go, simulate.go).
#pragma once
#include <functional>
#include <limits>
#include <map>
#include <memory>
#include <random>
#include <set>
#include <string>
#include <vector>

#include "json.hpp"
#include "model.hpp"
#include "nprand.hpp"

namespace det {

using RequestID = std::string;  // canonical lowercase UUIDv4 text (orders like the raw bytes)

struct Operation {
  enum Kind { Create, Train, Validate, Checkpoint, Close, Shutdown } kind = Shutdown;
  RequestID request_id;
  // Create
  uint32_t trial_seed = 0;
  Json hparams;
  bool has_checkpoint = false;
  RequestID checkpoint_request_id;  // Create-from-checkpoint: the trial whose checkpoint to start from
  // Train
  Length length;
  // Shutdown
  bool failure = false;

  bool runnable() const { return kind == Train || kind == Validate || kind == Checkpoint; }
  std::string str() const;
  Json to_json() const;
  static Operation from_json(const Json& j);

  static Operation make_create(NpRand& rng, const Json& hparams);
  static Operation make_create_from_checkpoint(NpRand& rng, const Json& hparams, const RequestID& ckpt_of);
  static Operation make_train(const RequestID& id, Length l) { Operation o; o.kind = Train; o.request_id = id; o.length = l; return o; }
  static Operation make_validate(const RequestID& id) { Operation o; o.kind = Validate; o.request_id = id; return o; }
  static Operation make_checkpoint(const RequestID& id) { Operation o; o.kind = Checkpoint; o.request_id = id; return o; }
  static Operation make_close(const RequestID& id) { Operation o; o.kind = Close; o.request_id = id; return o; }
  static Operation make_shutdown(bool failure = false) { Operation o; o.kind = Shutdown; o.failure = failure; return o; }
};
using Ops = std::vector<Operation>;

enum class ExitedReason { Errored, UserCanceled, InvalidHP };

struct SearchContext {
  NpRand* rand;
  const Hyperparameters* hparams;
};

Json sample_one(const Hyperparameter& h, NpRand& rng);
Json sample_all(const Hyperparameters& h, NpRand& rng);
std::vector<Json> hyperparameter_grid(const Hyperparameters& h);

class SearchMethod {
 public:
  virtual ~SearchMethod() = default;
  virtual Ops initial_operations(SearchContext ctx) = 0;
  virtual Ops trial_created(SearchContext, const RequestID&) { return {}; }
  virtual Ops train_completed(SearchContext, const RequestID&, const Operation&) { return {}; }
  virtual Ops checkpoint_completed(SearchContext, const RequestID&, const Operation&, const Json&) { return {}; }
  virtual Ops validation_completed(SearchContext, const RequestID&, const Operation&, const Json& /*metrics*/) { return {}; }
  virtual Ops trial_closed(SearchContext, const RequestID&) { return {}; }
  virtual double progress(double units_completed) = 0;
  virtual Ops trial_exited_early(SearchContext, const RequestID&, ExitedReason) { return {Operation::make_shutdown(true)}; }
  virtual Unit unit() const = 0;
};

std::unique_ptr<SearchMethod> new_search_method(const SearcherConfig& c);
double metric_from(const Json& validation_metrics, const std::string& name);  // throws when absent / non-scalar

// ---- facade + event log ---------------------------------------------------------------------------------------
class Searcher {
 public:
  Searcher(uint32_t seed, std::unique_ptr<SearchMethod> method, Hyperparameters hparams);
  Ops initial_operations();
  Ops trial_created(const Operation& create, int trial_id);
  Ops trial_exited_early(int trial_id, ExitedReason reason);
  void workload_completed(const Json& msg, double units_completed);
  Ops operation_completed(int trial_id, const Operation& op, const Json& metrics);
  Ops trial_closed(const RequestID& id);
  double progress();
  bool trial_id_of(const RequestID& id, int* out) const;
  bool request_id_of(int trial_id, RequestID* out) const;
  std::vector<Json> uncommitted_events();
  int trials_requested() const { return trials_requested_; }
  int trials_closed() const { return trials_closed_; }
  bool shutdown() const { return shutdown_; }

 private:
  void ops_created(const Ops& ops);
  SearchContext ctx() { return SearchContext{&rand_, &hparams_}; }
  NpRand rand_;
  Hyperparameters hparams_;
  std::unique_ptr<SearchMethod> method_;
  std::vector<Json> uncommitted_;
  std::set<RequestID> early_exits_;
  double total_units_completed_ = 0;
  bool shutdown_ = false;
  int trials_requested_ = 0, trials_closed_ = 0;
  std::map<RequestID, int> trial_ids_;
  std::map<int, RequestID> request_ids_;
};

// ---- simulation ---------------------------------------------------------------------------------------------------
using ValidationFunction = std::function<double(std::mt19937_64& rng, int trial_id, int idx)>;
struct SimulationResult {
  std::map<RequestID, std::vector<Operation>> results;  // per trial: the runnable ops it executed, in order
  long long seed = 0;
  Json summary() const;  // {"100B V 300B V": count, ...}
};
SimulationResult simulate(Searcher& s, long long seed, const ValidationFunction& val, bool random_order,
                          const std::string& metric_name);

}  // namespace det
