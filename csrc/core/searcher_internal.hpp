// searcher_internal.hpp — pieces shared by the search-method translation units (rungs, factory hooks).
#pragma once
#include <limits>
#include <memory>
#include <vector>

#include "searcher.hpp"

namespace det {

// ---- rungs (shared by SHA and ASHA) ---------------------------------------------------------------------------------
struct TrialMetric {
  RequestID request_id;
  double metric = 0;
  bool promoted = false;
};
struct Rung {
  Length units_needed;
  std::vector<TrialMetric> metrics;  // sorted ascending by metric
  int start_trials = 0, promote_trials = 0, outstanding_trials = 0;

  size_t insert_index(double metric) const {
    // first position whose metric is strictly greater (stable for ties)
    size_t lo = 0, hi = metrics.size();
    while (lo < hi) {
      size_t mid = (lo + hi) / 2;
      if (metrics[mid].metric > metric) hi = mid; else lo = mid + 1;
    }
    return lo;
  }
  std::vector<RequestID> promotions_sync(const RequestID& id, double metric) {
    const size_t at = insert_index(metric);
    metrics.insert(metrics.begin() + static_cast<long>(at), TrialMetric{id, metric, false});
    const int curr = static_cast<int>(metrics.size()) + promote_trials - start_trials;
    if (curr <= 0) return {};
    if (static_cast<int>(at) < curr) return {id};
    return {metrics[static_cast<size_t>(curr - 1)].request_id};
  }
  std::vector<RequestID> promotions_async(const RequestID& id, double metric, double divisor) {
    const int old_num = static_cast<int>(static_cast<double>(metrics.size()) / divisor);
    const int num = static_cast<int>(static_cast<double>(metrics.size() + 1) / divisor);
    const size_t at = insert_index(metric);
    const bool now = static_cast<int>(at) < num;
    metrics.insert(metrics.begin() + static_cast<long>(at), TrialMetric{id, metric, now});
    if (now) return {id};
    if (num != old_num && !metrics[static_cast<size_t>(old_num)].promoted) {
      metrics[static_cast<size_t>(old_num)].promoted = true;
      return {metrics[static_cast<size_t>(old_num)].request_id};
    }
    return {};
  }
};

constexpr double kExitedMetricValue = std::numeric_limits<double>::max();

std::unique_ptr<SearchMethod> new_async_halving(const SearcherConfig& c);
std::unique_ptr<SearchMethod> new_adaptive_asha(const SearcherConfig& c);
std::unique_ptr<SearchMethod> new_pbt(const SearcherConfig& c);
std::unique_ptr<SearchMethod> new_tournament(std::vector<std::unique_ptr<SearchMethod>> subs);
std::vector<int> adaptive_brackets(const std::string& mode, int max_rungs);

}  // namespace det
