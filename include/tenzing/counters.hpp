// Optional per-phase counters (reference: include/tenzing/counters.hpp:13-34 and
// tenzing-mcts/include/tenzing/mcts/counters.hpp:15-25).  Compiled in by
// default here (they cost one add per phase); define TENZING_DISABLE_COUNTERS
// to compile them out.
#pragma once

#include <chrono>
#include <cstdint>

namespace tenzing {
namespace counters {

template <typename T> struct ZeroInit {
  T v{};
  ZeroInit() = default;
  ZeroInit(T x) : v(x) {}
  operator T() const { return v; }
  template <typename U> ZeroInit &operator+=(U x) { v += T(x); return *this; }
};

inline double now() {
  using clk = std::chrono::steady_clock;
  return std::chrono::duration<double>(clk::now().time_since_epoch()).count();
}

// phases of one MCTS iteration
struct Mcts {
  ZeroInit<double> SELECT_TIME, EXPAND_TIME, ROLLOUT_TIME, REDUNDANT_SYNC_TIME, RMAP_TIME,
      BENCHMARK_TIME, BACKPROP_TIME;
  ZeroInit<uint64_t> ITERS;
};
struct Dfs {
  ZeroInit<double> ENUMERATE_TIME, DEDUP_TIME, BENCHMARK_TIME;
  ZeroInit<uint64_t> STATES_VISITED, STATES_PRUNED;
};
// counters of the calling thread (= of the rank, also when ranks are threads of one process)
Mcts &mcts();
Dfs &dfs();
void reset();

} // namespace counters
} // namespace tenzing

#ifndef TENZING_DISABLE_COUNTERS
#define TENZING_COUNTER_EXPR(expr) expr
#define TENZING_COUNTER_OP(group, key, op) ::tenzing::counters::group().key op
#else
#define TENZING_COUNTER_EXPR(expr)
#define TENZING_COUNTER_OP(group, key, op)
#endif
