// Exhaustive search: enumerate every distinct complete sequence, benchmark each.
//
// Reference: tenzing-dfs/include/tenzing/dfs/dfs.hpp:26-178, tenzing-dfs/src/dfs.cpp:11-104.
// The enumeration keeps a hash set of canonical state keys, so states reached
// through different decision orders (and sequences that differ only by a
// renaming of streams/events) are visited once; the reference removes
// equivalent sequences afterwards with an all-pairs comparison.
#pragma once

#include <cstdint>
#include <iostream>
#include <vector>

#include "tenzing/benchmarker.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/json.hpp"
#include "tenzing/platform.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {
namespace dfs {

struct Opts {
  int64_t maxSeqs = -1; // stop enumerating after this many sequences; negative = unlimited
  bool removeRedundantSyncs = false; // peephole-clean each sequence before de-duplication
  Benchmark::Opts benchOpts;
};
void to_json(Json &j, const Opts &opts);

struct SimResult {
  Sequence<BoundOp> seq;
  Benchmark::Result benchResult;
};

struct Result {
  std::vector<SimResult> simResults;
  Opts opts_;
  explicit Result(const Opts &opts = Opts()) : opts_(opts) {}
  // {"dfs__Opts":{"maxSeqs":N}} then one row per sequence
  void dump_csv(std::ostream &os) const;
  void dump_csv() const { dump_csv(std::cout); }
  std::string csv() const;
  // index of the fastest sequence by pct10 (-1 if empty)
  int best() const;
};

// stop flag shared between ranks
struct Stop {
  enum class Reason { predicate };
  Reason reason_ = Reason::predicate;
  bool value_ = false;
  Stop() = default;
  Stop(bool b, Reason r) : reason_(r), value_(b) {}
  explicit operator bool() const { return value_; }
  const char *c_str() const { return "predicate"; }
  void bcast(int root, Communicator &comm);
};

// every distinct complete sequence of g on plat (streams = plat.streams_)
std::vector<Sequence<BoundOp>> get_all_sequences(const Graph<OpBase> &g, const Platform &plat,
                                                 int64_t maxSeqs = -1,
                                                 bool removeRedundantSyncs = false);

// rank 0 enumerates; all ranks benchmark each sequence in lock step
Result explore(const Graph<OpBase> &g, Platform &plat, Benchmarker &benchmarker,
               const Opts &opts = Opts());

} // namespace dfs
} // namespace tenzing
