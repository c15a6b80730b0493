// elastic_logs.hpp — Elasticsearch as the trial-log store (master config `logging: {type: elastic, host, port, index}`).
//
// Reference: master/internal/elastic/elastic_trial_logs.go (bulk insert of log rows, filtered / paginated search ordered
// by timestamp) selected by `logging.type: elastic` in the master config; the default there is Postgres, here the
// per-trial JSONL files of the journal store. Rows are written with the `_bulk` API (one `index` action per line, index
// `determined-triallogs` unless configured) and read back with `_search`: term filter on trial_id, sorted by the sequence
// number the master stamps on every row, `from` / `size` pagination — the same position-based contract the file store
// gives the REST route (`GET /api/v1/trials/:id/logs?offset=&limit=`).
#pragma once
#include <algorithm>
#include <atomic>
#include <string>
#include <vector>

#include "http.hpp"

namespace det {

class ElasticLogs {
 public:
  ElasticLogs(std::string host, int port, std::string index) : host_(std::move(host)), port_(port), index_(std::move(index)) {}

  bool append(long long trial_id, const std::vector<Json>& lines, std::string* err) {
    std::string body;
    for (auto& l : lines) {
      Json row = l.deep_copy();
      row["trial_id"] = trial_id;
      row["seq"] = static_cast<long long>(seq_.fetch_add(1));
      body += "{\"index\":{\"_index\":\"" + index_ + "\"}}\n" + row.dump() + "\n";
    }
    if (body.empty()) return true;
    std::string out;
    const int st = http_request(host_, port_, "POST", "/_bulk?refresh=true", body, &out, 30);
    if (st < 200 || st >= 300) { if (err) *err = "elasticsearch _bulk answered " + std::to_string(st) + ": " + out.substr(0, 200); return false; }
    try {
      if (Json::parse(out)["errors"].as_bool(false)) { if (err) *err = "elasticsearch _bulk reported item errors"; return false; }
    } catch (...) {}
    return true;
  }

  std::vector<Json> read(long long trial_id, size_t offset, size_t limit, std::string* err) {
    std::vector<Json> rows;
    const size_t page = 10000;          // one _search window
    while (rows.size() < limit) {
      const size_t want = std::min(page, limit - rows.size());
      Json q = Json::object();
      q["from"] = static_cast<long long>(offset + rows.size());
      q["size"] = static_cast<long long>(want);
      Json term = Json::object();
      term["trial_id"] = trial_id;
      Json query = Json::object();
      query["term"] = term;
      q["query"] = query;
      Json sort_field = Json::object();
      sort_field["seq"] = "asc";
      Json sort = Json::array();
      sort.push_back(sort_field);
      q["sort"] = sort;
      std::string out;
      const int st = http_request(host_, port_, "POST", "/" + index_ + "/_search", q.dump(), &out, 30);
      if (st < 200 || st >= 300) { if (err) *err = "elasticsearch _search answered " + std::to_string(st); return rows; }
      size_t got = 0;
      try {
        Json r = Json::parse(out);
        for (auto& h : r["hits"]["hits"].arr()) { rows.push_back(h["_source"]); ++got; }
      } catch (const std::exception& e) {
        if (err) *err = std::string("elasticsearch _search: ") + e.what();
        return rows;
      }
      if (got < want) break;
    }
    return rows;
  }

 private:
  std::string host_;
  int port_;
  std::string index_;
  std::atomic<long long> seq_{static_cast<long long>(std::time(nullptr)) * 1000000LL};   // monotone across master restarts
};

}  // namespace det
