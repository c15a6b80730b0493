// store.hpp — the master's persistence: an in-memory table store backed by an append-only JSON-lines journal
// (replayed at start-up) plus per-trial log files. Replaces the reference's Postgres layer
// (master/internal/db/postgres.go, static/migrations): tables experiments, trials, steps, validations,
// checkpoints, searcher_events, users, sessions, templates, models, model_versions, commands, cluster.
#pragma once
#include <sys/stat.h>

#include <fstream>
#include <functional>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "json.hpp"

namespace det {

inline void mkdirs(const std::string& path) {
  std::string cur;
  for (size_t i = 0; i <= path.size(); ++i) {
    if (i == path.size() || path[i] == '/') {
      if (!cur.empty()) ::mkdir(cur.c_str(), 0755);
    }
    if (i < path.size()) cur.push_back(path[i]);
  }
}

class Store {
 public:
  explicit Store(const std::string& dir) : dir_(dir) {
    mkdirs(dir_);
    mkdirs(dir_ + "/logs");
    journal_path_ = dir_ + "/journal.jsonl";
    replay();
    journal_.open(journal_path_, std::ios::app);
  }

  long long next_id(const std::string& table) {
    std::lock_guard<std::mutex> g(mu_);
    long long id = ++counters_[table];
    write({{"op", "counter"}, {"table", table}, {"value", id}});
    return id;
  }
  void put(const std::string& table, const std::string& id, const Json& row) {
    std::lock_guard<std::mutex> g(mu_);
    tables_[table][id] = row.deep_copy();
    write({{"op", "put"}, {"table", table}, {"id", id}, {"row", row}});
  }
  void put(const std::string& table, long long id, const Json& row) { put(table, std::to_string(id), row); }
  // Shallow-merge `fields` into the row.
  Json update(const std::string& table, const std::string& id, const Json& fields) {
    std::lock_guard<std::mutex> g(mu_);
    Json& row = tables_[table][id];
    if (!row.is_object()) row = Json::object();
    for (auto& kv : fields.obj()) row[kv.first] = kv.second.deep_copy();
    write({{"op", "put"}, {"table", table}, {"id", id}, {"row", row}});
    return row.deep_copy();
  }
  Json update(const std::string& table, long long id, const Json& fields) { return update(table, std::to_string(id), fields); }
  Json get(const std::string& table, const std::string& id) {
    std::lock_guard<std::mutex> g(mu_);
    auto t = tables_.find(table);
    if (t == tables_.end()) return Json();
    auto r = t->second.find(id);
    return r == t->second.end() ? Json() : r->second.deep_copy();
  }
  Json get(const std::string& table, long long id) { return get(table, std::to_string(id)); }
  bool exists(const std::string& table, const std::string& id) { return !get(table, id).is_null(); }
  void del(const std::string& table, const std::string& id) {
    std::lock_guard<std::mutex> g(mu_);
    tables_[table].erase(id);
    write({{"op", "del"}, {"table", table}, {"id", id}});
  }
  std::vector<Json> list(const std::string& table, const std::function<bool(const Json&)>& pred = nullptr) {
    std::lock_guard<std::mutex> g(mu_);
    std::vector<Json> out;
    auto t = tables_.find(table);
    if (t == tables_.end()) return out;
    for (auto& kv : t->second)
      if (!pred || pred(kv.second)) out.push_back(kv.second.deep_copy());
    return out;
  }
  // Append-only child rows (searcher events, metric series): kept in insertion order under `key`.
  void append(const std::string& table, const std::string& key, const Json& row) {
    std::lock_guard<std::mutex> g(mu_);
    series_[table][key].push_back(row.deep_copy());
    write({{"op", "append"}, {"table", table}, {"id", key}, {"row", row}});
  }
  std::vector<Json> series(const std::string& table, const std::string& key) {
    std::lock_guard<std::mutex> g(mu_);
    std::vector<Json> out;
    for (auto& r : series_[table][key]) out.push_back(r.deep_copy());
    return out;
  }
  void truncate_series(const std::string& table, const std::string& key, size_t keep) {
    std::lock_guard<std::mutex> g(mu_);
    auto& v = series_[table][key];
    if (v.size() > keep) v.resize(keep);
    write({{"op", "truncate"}, {"table", table}, {"id", key}, {"value", static_cast<long long>(keep)}});
  }

  // ---- trial logs: one file per trial ----
  // an external log store (Elasticsearch, elastic_logs.hpp) replaces the per-trial files when both hooks are set
  std::function<void(long long, const std::vector<Json>&)> log_append_hook;
  std::function<std::vector<Json>(long long, size_t, size_t)> log_read_hook;
  void append_logs(long long trial_id, const std::vector<Json>& lines) {
    if (log_append_hook) { log_append_hook(trial_id, lines); return; }
    std::lock_guard<std::mutex> g(log_mu_);
    std::ofstream f(dir_ + "/logs/trial-" + std::to_string(trial_id) + ".jsonl", std::ios::app);
    for (auto& l : lines) f << l.dump() << "\n";
  }
  std::vector<Json> read_logs(long long trial_id, size_t offset, size_t limit) {
    if (log_read_hook) return log_read_hook(trial_id, offset, limit);
    std::lock_guard<std::mutex> g(log_mu_);
    std::vector<Json> out;
    std::ifstream f(dir_ + "/logs/trial-" + std::to_string(trial_id) + ".jsonl");
    std::string line;
    size_t i = 0;
    while (std::getline(f, line)) {
      if (i++ < offset) continue;
      if (out.size() >= limit) break;
      try { Json j = Json::parse(line); j["id"] = static_cast<long long>(i - 1); out.push_back(j); } catch (...) {}   // id: position in the log
    }
    return out;
  }
  const std::string& dir() const { return dir_; }

 private:
  static Json obj(std::initializer_list<std::pair<const char*, Json>> kv) {
    Json j = Json::object();
    for (auto& p : kv) j[p.first] = p.second;
    return j;
  }
  void write(std::initializer_list<std::pair<const char*, Json>> kv) {
    if (replaying_ || !journal_.is_open()) return;
    journal_ << obj(kv).dump() << "\n";
    journal_.flush();
  }
  void replay() {
    replaying_ = true;
    std::ifstream f(journal_path_);
    std::string line;
    while (std::getline(f, line)) {
      try {
        Json j = Json::parse(line);
        const std::string op = j["op"].as_string(), table = j["table"].as_string(), id = j["id"].str_or("");
        if (op == "counter") counters_[table] = j["value"].as_int();
        else if (op == "put") tables_[table][id] = j["row"];
        else if (op == "del") tables_[table].erase(id);
        else if (op == "append") series_[table][id].push_back(j["row"]);
        else if (op == "truncate") { auto& v = series_[table][id]; size_t keep = static_cast<size_t>(j["value"].as_int()); if (v.size() > keep) v.resize(keep); }
      } catch (...) {
      }
    }
    replaying_ = false;
  }
  std::string dir_, journal_path_;
  std::mutex mu_, log_mu_;
  std::ofstream journal_;
  bool replaying_ = false;
  std::map<std::string, long long> counters_;
  std::map<std::string, std::map<std::string, Json>> tables_;
  std::map<std::string, std::map<std::string, std::vector<Json>>> series_;
};

}  // namespace det
