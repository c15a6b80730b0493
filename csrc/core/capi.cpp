// capi.cpp — C ABI of libdet_core.so (declared in include/det_core.h). Everything crosses the boundary as JSON
// text; returned strings are malloc'd and released with det_core_free(). Errors come back as {"error": "..."}.
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <random>

#include "json.hpp"
#include "lttb.hpp"
#include "model.hpp"
#include "nprand.hpp"
#include "cloud_providers.hpp"
#include "kubernetes.hpp"
#include "provisioner.hpp"
#include "scheduler.hpp"
#include "searcher.hpp"
#include "sequencer.hpp"

using namespace det;

namespace {
char* dup(const std::string& s) {
  char* p = static_cast<char*>(std::malloc(s.size() + 1));
  std::memcpy(p, s.c_str(), s.size() + 1);
  return p;
}
char* err(const std::string& what) {
  Json j = Json::object();
  j["error"] = what;
  return dup(j.dump());
}
template <class F>
char* guarded(F&& f) {
  try {
    return dup(f().dump());
  } catch (const std::exception& e) {
    return err(e.what());
  }
}
Json ops_json(const Ops& ops) {
  Json a = Json::array();
  for (auto& o : ops) a.push_back(o.to_json());
  return a;
}
ExitedReason reason_of(const std::string& s) {
  if (s == "USER_CANCELED") return ExitedReason::UserCanceled;
  if (s == "INVALID_HP") return ExitedReason::InvalidHP;
  return ExitedReason::Errored;
}
struct SearcherHandle {
  std::unique_ptr<Searcher> s;
  SearcherConfig cfg;
};
}  // namespace

extern "C" {

void det_core_free(char* p) { std::free(p); }
int det_core_abi_version() { return 1; }

// ---- nprand ------------------------------------------------------------------------------------------------------
void* det_core_nprand_new(unsigned seed) { return new NpRand(seed); }
void det_core_nprand_free(void* h) { delete static_cast<NpRand*>(h); }
unsigned det_core_nprand_bits32(void* h) { return static_cast<NpRand*>(h)->bits32(); }
double det_core_nprand_unit_interval(void* h) { return static_cast<NpRand*>(h)->unit_interval(); }
double det_core_nprand_uniform(void* h, double lo, double hi) { return static_cast<NpRand*>(h)->uniform(lo, hi); }
long long det_core_nprand_int64n(void* h, long long n) { return static_cast<NpRand*>(h)->int64n(n); }
long long det_core_nprand_int64(void* h, long long lo, long long hi) { return static_cast<NpRand*>(h)->int64_range(lo, hi); }
char* det_core_nprand_uuid4(void* h) { return dup(static_cast<NpRand*>(h)->uuid4()); }
void det_core_nprand_bytes(void* h, unsigned char* out, int n) { static_cast<NpRand*>(h)->read(out, static_cast<size_t>(n)); }

// ---- config ---------------------------------------------------------------------------------------------------------
// in: {"config": {...}, "master_checkpoint_storage": {...}?, "template": {...}?, "default_seed": n?}
// out: {"config": resolved, "errors": [..]}
char* det_core_resolve_config(const char* in) {
  return guarded([&] {
    Json j = Json::parse(in);
    Json storage = j["master_checkpoint_storage"], tmpl = j["template"];
    Json out = Json::object();
    Json cfg = resolve_experiment_config(j["config"], storage.is_object() ? &storage : nullptr, tmpl.is_object() ? &tmpl : nullptr,
                                         static_cast<unsigned>(j["default_seed"].as_int(0)));
    out["config"] = cfg;
    Json errs = Json::array();
    for (auto& e : validate_experiment_config(cfg)) errs.push_back(e);
    out["errors"] = errs;
    return out;
  });
}

// ---- searcher -----------------------------------------------------------------------------------------------------------
// in: {"searcher": {...}, "hyperparameters": {...}, "seed": n}
void* det_core_searcher_new(const char* in, char** error_out) {
  try {
    Json j = Json::parse(in);
    auto h = std::make_unique<SearcherHandle>();
    h->cfg = SearcherConfig::from_json(j["searcher"]);
    std::vector<std::string> errs;
    h->cfg.validate(errs);
    if (!errs.empty()) throw ConfigError(errs[0]);
    h->s = std::make_unique<Searcher>(static_cast<uint32_t>(j["seed"].as_int(0)), new_search_method(h->cfg),
                                      hyperparameters_from_json(j["hyperparameters"]));
    if (error_out) *error_out = nullptr;
    return h.release();
  } catch (const std::exception& e) {
    if (error_out) *error_out = dup(e.what());
    return nullptr;
  }
}
void det_core_searcher_free(void* h) { delete static_cast<SearcherHandle*>(h); }
char* det_core_searcher_initial_operations(void* h) {
  return guarded([&] { return ops_json(static_cast<SearcherHandle*>(h)->s->initial_operations()); });
}
char* det_core_searcher_trial_created(void* h, const char* create_json, int trial_id) {
  return guarded([&] { return ops_json(static_cast<SearcherHandle*>(h)->s->trial_created(Operation::from_json(Json::parse(create_json)), trial_id)); });
}
char* det_core_searcher_operation_completed(void* h, int trial_id, const char* op_json, const char* metrics_json) {
  return guarded([&] {
    return ops_json(static_cast<SearcherHandle*>(h)->s->operation_completed(trial_id, Operation::from_json(Json::parse(op_json)),
                                                                            Json::parse(metrics_json ? metrics_json : "{}")));
  });
}
char* det_core_searcher_trial_closed(void* h, const char* request_id) {
  return guarded([&] { return ops_json(static_cast<SearcherHandle*>(h)->s->trial_closed(request_id)); });
}
char* det_core_searcher_trial_exited_early(void* h, int trial_id, const char* reason) {
  return guarded([&] { return ops_json(static_cast<SearcherHandle*>(h)->s->trial_exited_early(trial_id, reason_of(reason ? reason : "ERRORED"))); });
}
void det_core_searcher_workload_completed(void* h, const char* msg_json, double units) {
  try {
    static_cast<SearcherHandle*>(h)->s->workload_completed(Json::parse(msg_json ? msg_json : "{}"), units);
  } catch (...) {
  }
}
double det_core_searcher_progress(void* h) { return static_cast<SearcherHandle*>(h)->s->progress(); }
char* det_core_searcher_events(void* h) {
  return guarded([&] {
    Json a = Json::array();
    for (auto& e : static_cast<SearcherHandle*>(h)->s->uncommitted_events()) a.push_back(e);
    return a;
  });
}

// in: {"searcher":..., "hyperparameters":..., "seed": n, "simulation_seed": n, "validation": "constant"|"random", "random_order": bool}
// out: {"results": {"100B V": 3, ...}, "trials": n, "seed": n}
char* det_core_simulate(const char* in) {
  return guarded([&] {
    Json j = Json::parse(in);
    SearcherConfig cfg = SearcherConfig::from_json(j["searcher"]);
    std::vector<std::string> errs;
    cfg.validate(errs);
    if (!errs.empty()) throw ConfigError(errs[0]);
    Searcher s(static_cast<uint32_t>(j["seed"].as_int(0)), new_search_method(cfg), hyperparameters_from_json(j["hyperparameters"]));
    const bool random_val = j["validation"].str_or("random") == "random";
    ValidationFunction vf = [random_val](std::mt19937_64& rng, int, int) {
      return random_val ? std::uniform_real_distribution<double>(0.0, 1.0)(rng) : 1.0;
    };
    SimulationResult r = simulate(s, j["simulation_seed"].as_int(0), vf, j["random_order"].as_bool(false), cfg.metric);
    Json out = Json::object();
    out["results"] = r.summary();
    out["trials"] = static_cast<long long>(r.results.size());
    out["seed"] = r.seed;
    Json per = Json::object();
    for (auto& kv : r.results) {
      Json a = Json::array();
      for (auto& op : kv.second) a.push_back(op.to_json());
      per[kv.first] = a;
    }
    out["per_trial"] = per;
    return out;
  });
}

// ---- sequencer --------------------------------------------------------------------------------------------------------------
// in: {"config": resolved experiment config, "experiment_id": n, "create": create-op json, "first_checkpoint": {...}|null}
void* det_core_sequencer_new(const char* in, char** error_out) {
  try {
    Json j = Json::parse(in);
    auto* s = new TrialWorkloadSequencer(j["config"], static_cast<int>(j["experiment_id"].as_int(1)), Operation::from_json(j["create"]), j["first_checkpoint"]);
    if (error_out) *error_out = nullptr;
    return s;
  } catch (const std::exception& e) {
    if (error_out) *error_out = dup(e.what());
    return nullptr;
  }
}
void det_core_sequencer_free(void* h) { delete static_cast<TrialWorkloadSequencer*>(h); }
void det_core_sequencer_set_trial_id(void* h, int id) { static_cast<TrialWorkloadSequencer*>(h)->set_trial_id(id); }
char* det_core_sequencer_operation_requested(void* h, const char* op_json) {
  return guarded([&] {
    static_cast<TrialWorkloadSequencer*>(h)->operation_requested(Operation::from_json(Json::parse(op_json)));
    return Json::object();
  });
}
int det_core_sequencer_up_to_date(void* h) { return static_cast<TrialWorkloadSequencer*>(h)->up_to_date() ? 1 : 0; }
char* det_core_sequencer_workload(void* h) {
  return guarded([&] { return static_cast<TrialWorkloadSequencer*>(h)->workload().to_json(); });
}
// msg: {"workload": {...}, "exited_reason": "..."?, "metrics": .., "validation_metrics": .., "checkpoint_metrics": ..}
// out: {"op": op|null, "metrics": ...}
char* det_core_sequencer_workload_completed(void* h, const char* msg_json, int is_best_validation) {
  return guarded([&] {
    Json m = Json::parse(msg_json);
    CompletedMessage msg;
    msg.workload = Workload::from_json(m["workload"]);
    if (m["exited_reason"].is_string()) { msg.has_exited_reason = true; msg.exited_reason = reason_of(m["exited_reason"].as_string()); }
    msg.metrics = m["metrics"]; msg.validation_metrics = m["validation_metrics"]; msg.checkpoint_metrics = m["checkpoint_metrics"];
    Operation op;
    Json metrics;
    Json out = Json::object();
    if (static_cast<TrialWorkloadSequencer*>(h)->workload_completed(msg, is_best_validation != 0, &op, &metrics)) {
      out["op"] = op.to_json();
      out["metrics"] = metrics;
    } else {
      out["op"] = Json();
    }
    return out;
  });
}
char* det_core_sequencer_complete_cached_checkpoints(void* h) {
  return guarded([&] {
    Operation op;
    Json metrics, out = Json::object();
    if (static_cast<TrialWorkloadSequencer*>(h)->complete_cached_checkpoints(&op, &metrics)) { out["op"] = op.to_json(); out["metrics"] = metrics; }
    else out["op"] = Json();
    return out;
  });
}
char* det_core_sequencer_preclose_checkpoint(void* h) {
  return guarded([&] {
    Workload w;
    if (static_cast<TrialWorkloadSequencer*>(h)->preclose_checkpoint_workload(&w)) return w.to_json();
    return Json();
  });
}
char* det_core_sequencer_terminate_workload(void* h) {
  return guarded([&] { return static_cast<TrialWorkloadSequencer*>(h)->terminate_workload().to_json(); });
}
int det_core_sequencer_roll_back(void* h) { return static_cast<TrialWorkloadSequencer*>(h)->roll_back(); }
char* det_core_sequencer_state(void* h) {
  return guarded([&] { return static_cast<TrialWorkloadSequencer*>(h)->snapshot_json(); });
}

// ---- scheduler / misc -----------------------------------------------------------------------------------------------------------
char* det_core_schedule(const char* in) {
  return guarded([&] { return schedule_json(Json::parse(in)); });
}
// in: {"config": {max_instances, slots_per_instance, max_idle_agent_period, max_agent_starting_period},
//      "snapshot": {instances, agents, demand, now}} -> {"launch": n, "terminate": [...], "reasons": {...}}
char* det_core_scale_decide(const char* in) {
  return guarded([&] {
    Json j = Json::parse(in);
    return scale_decide(ProvisionerConfig::from_json(j["config"]), j["snapshot"]);
  });
}
// in: {"points": [[x,y],...], "threshold": n} -> [[x,y],...]
char* det_core_lttb(const char* in) {
  return guarded([&] {
    Json j = Json::parse(in);
    std::vector<std::pair<double, double>> pts;
    for (auto& p : j["points"].arr()) pts.emplace_back(p.at(0).as_double(), p.at(1).as_double());
    Json out = Json::array();
    for (auto& p : lttb_downsample(pts, static_cast<int>(j["threshold"].as_int(0)))) {
      Json q = Json::array();
      q.push_back(p.first);
      q.push_back(p.second);
      out.push_back(q);
    }
    return out;
  });
}
// in: {"container": {container_id, task_id, description, argv, env, cwd, devices:[{id,type,...}]}, "config": {kubernetes RM config},
//      "cluster_id": s, "node": s, "host_network": b} -> the pod manifest the Kubernetes back-end would POST
char* det_core_pod_spec(const char* in) {
  return guarded([&] {
    Json j = Json::parse(in);
    const Json& c = j["container"];
    ContainerSpec spec;
    spec.container_id = c["container_id"].str_or("");
    spec.task_id = c["task_id"].str_or(spec.container_id);
    spec.description = c["description"].str_or("");
    spec.cwd = c["cwd"].str_or("");
    if (c["argv"].is_array()) for (auto& a : c["argv"].arr()) spec.argv.push_back(a.as_string());
    if (c["env"].is_object()) for (auto& kv : c["env"].obj()) spec.env[kv.first] = kv.second.as_string();
    if (c["devices"].is_array())
      for (auto& d : c["devices"].arr()) spec.devices.push_back(Device{static_cast<int>(d["id"].as_int()), d["brand"].str_or(""), d["uuid"].str_or(""), d["type"].str_or("gpu")});
    return make_pod_spec(spec, KubeConfig::from_json(j["config"]), j["cluster_id"].str_or("cluster"), j["node"].str_or(""), j["host_network"].as_bool(false));
  });
}
// in: {"provider": {type: aws|gcp, ...}, "op": "launch"|"list"|"terminate", "count": n, "ids": [...], "names": [...],
//      "cluster_id": s, "master_host": s, "master_port": n} -> {"argv": [...]} ; op "parse" + "output": <cli json> -> {"instances": [...]}
char* det_core_cloud_command(const char* in) {
  return guarded([&] {
    Json j = Json::parse(in);
    CloudProviderConfig c = cloud_config_from_json(j["provider"], j["cluster_id"].str_or("cluster"), j["master_host"].str_or("127.0.0.1"),
                                                   static_cast<int>(j["master_port"].as_int(8080)), static_cast<int>(j["slots_per_instance"].as_int(8)));
    const std::string op = j["op"].str_or("list");
    std::vector<std::string> ids;
    if (j["ids"].is_array()) for (auto& i : j["ids"].arr()) ids.push_back(i.as_string());
    Json out = Json::object();
    if (op == "parse") {
      Json arr = Json::array();
      for (auto& i : (c.type == "aws" ? aws_parse_list(j["output"]) : gcp_parse_list(j["output"]))) {
        Json o = Json::object();
        o["id"] = i.id; o["state"] = i.state;
        arr.push_back(o);
      }
      out["instances"] = arr;
      return out;
    }
    std::vector<std::string> argv;
    if (c.type == "aws") argv = op == "launch" ? aws_launch_argv(c, static_cast<int>(j["count"].as_int(1))) : (op == "terminate" ? aws_terminate_argv(c, ids) : aws_list_argv(c));
    else if (c.type == "gcp") argv = op == "launch" ? gcp_launch_argv(c, ids) : (op == "terminate" ? gcp_terminate_argv(c, ids) : gcp_list_argv(c));
    else throw ConfigError("provider.type must be aws or gcp");
    Json a = Json::array();
    for (auto& s : argv) a.push_back(Json(s));
    out["argv"] = a;
    return out;
  });
}
unsigned long long det_core_md5_low64(const char* s) { return md5_low64(s); }

}  // extern "C"
