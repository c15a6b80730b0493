// lttb.hpp — Largest-Triangle-Three-Buckets downsampling of a metric series for charts.
// Parity: reference master/internal/lttb/lttb.go:1-112 (keep first and last point; per bucket keep the point
// forming the largest triangle with the previously kept point and the next bucket's average).
#pragma once
#include <cmath>
#include <utility>
#include <vector>

namespace det {

inline std::vector<std::pair<double, double>> lttb_downsample(const std::vector<std::pair<double, double>>& data, int threshold) {
  const int n = static_cast<int>(data.size());
  if (threshold >= n || threshold <= 2) return data;
  std::vector<std::pair<double, double>> out;
  out.reserve(static_cast<size_t>(threshold));
  const double every = static_cast<double>(n - 2) / static_cast<double>(threshold - 2);
  int a = 0;
  out.push_back(data[0]);
  for (int i = 0; i < threshold - 2; ++i) {
    int avg_start = static_cast<int>(std::floor((i + 1) * every)) + 1;
    int avg_end = std::min(static_cast<int>(std::floor((i + 2) * every)) + 1, n);
    double avg_x = 0, avg_y = 0;
    const int avg_len = std::max(avg_end - avg_start, 1);
    for (int j = avg_start; j < avg_end; ++j) { avg_x += data[static_cast<size_t>(j)].first; avg_y += data[static_cast<size_t>(j)].second; }
    avg_x /= avg_len; avg_y /= avg_len;
    const int range_start = static_cast<int>(std::floor(i * every)) + 1;
    const int range_end = static_cast<int>(std::floor((i + 1) * every)) + 1;
    const double ax = data[static_cast<size_t>(a)].first, ay = data[static_cast<size_t>(a)].second;
    double best = -1;
    int best_idx = range_start;
    for (int j = range_start; j < range_end && j < n; ++j) {
      const double area = std::fabs((ax - avg_x) * (data[static_cast<size_t>(j)].second - ay) - (ax - data[static_cast<size_t>(j)].first) * (avg_y - ay)) * 0.5;
      if (area > best) { best = area; best_idx = j; }
    }
    out.push_back(data[static_cast<size_t>(best_idx)]);
    a = best_idx;
  }
  out.push_back(data[static_cast<size_t>(n - 1)]);
  return out;
}

}  // namespace det
