// Small statistics helpers (reference: include/tenzing/numeric.hpp:18-109,
// src/numeric.cpp:11-42).  `med` is the true median (the reference reads one
// element past the middle for even sizes).
#pragma once

#include <algorithm>
#include <cmath>
#include <cstddef>
#include <vector>

#include "tenzing/macro_at.hpp"

namespace tenzing {

template <typename T> double avg(const std::vector<T> &v) {
  double acc = 0;
  for (const T &e : v) acc += double(e);
  return acc / double(v.size());
}

template <typename T> double med(const std::vector<T> &v) {
  if (v.empty()) return 0;
  std::vector<T> s(v);
  std::sort(s.begin(), s.end());
  const size_t n = s.size();
  return (n % 2) ? double(s[n / 2]) : (double(s[n / 2 - 1]) + double(s[n / 2])) / 2.0;
}

// population variance
template <typename T> double var(const std::vector<T> &v) {
  const double m = avg(v);
  double acc = 0;
  for (const T &e : v) acc += (double(e) - m) * (double(e) - m);
  return acc / double(v.size());
}

template <typename T> double stddev(const std::vector<T> &v) { return std::sqrt(var(v)); }

// Pearson correlation, clamped to [-1, 1]; 0 for empty input or zero variance
template <typename T> double corr(const std::vector<T> &a, const std::vector<T> &b) {
  if (a.size() != b.size()) THROW_RUNTIME("corr: vectors must be the same size");
  if (a.empty()) return 0;
  const double am = avg(a), bm = avg(b), as = stddev(a), bs = stddev(b);
  if (as == 0 || bs == 0) return 0;
  double acc = 0;
  for (size_t i = 0; i < a.size(); ++i) acc += (double(a[i]) - am) * (double(b[i]) - bm);
  double c = acc / (double(a.size()) * as * bs);
  if (c < -1.01 || c > 1.01) THROW_RUNTIME("corr=" << c << " outside [-1,1]");
  return std::min(1.0, std::max(-1.0, c));
}

// prime factors, largest first; empty for 0 and 1
template <typename T> std::vector<T> prime_factors(T n) {
  std::vector<T> r;
  if (n <= 1) return r;
  for (T p = 2; p * p <= n; p += (p == 2 ? 1 : 2)) {
    while (n % p == 0) {
      r.push_back(p);
      n /= p;
    }
  }
  if (n > 1) r.push_back(n);
  std::sort(r.begin(), r.end(), [](T x, T y) { return y < x; });
  return r;
}

template <typename T> T round_up(T x, T step) { return (x + step - 1) / step * step; }

// sorted-vector percentile with the reference's index rule size*p/100
template <typename T> T percentile_sorted(const std::vector<T> &sorted, int p) {
  size_t i = sorted.size() * size_t(p) / 100;
  if (i >= sorted.size()) i = sorted.size() - 1;
  return sorted[i];
}

} // namespace tenzing
