// A state of the sequential decision process: the (constrained) program graph
// plus the executable sequence chosen so far.
//
// Reference: include/tenzing/state.hpp:15-54, src/state.cpp:5-143.
// The graph is held through a shared pointer, so the common decision
// (ExecuteOp) copies only the sequence.
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "tenzing/decision.hpp"
#include "tenzing/event_synchronizer.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {
namespace SDP {

class State {
  std::shared_ptr<const Graph<OpBase>> graph_;
  Sequence<BoundOp> sequence_;
  mutable std::string keyCache_; // canonical_key() is asked for by frontier() and by the solvers

public:
  // initial state: sequence = [Start]
  explicit State(const Graph<OpBase> &graph);
  State(const Graph<OpBase> &graph, const Sequence<BoundOp> &sequence);
  State(std::shared_ptr<const Graph<OpBase>> graph, const Sequence<BoundOp> &sequence)
      : graph_(std::move(graph)), sequence_(sequence) {}

  const Graph<OpBase> &graph() const { return *graph_; }
  const std::shared_ptr<const Graph<OpBase>> &graph_ptr() const { return graph_; }
  const Sequence<BoundOp> &sequence() const { return sequence_; }

  // synchronisation ops `op` still needs before it may execute (empty = ready)
  std::vector<BoundOpPtr> get_syncs_before_op(const BoundOpPtr &op) const;
  // every decision available in this state
  std::vector<DecisionPtr> get_decisions(const Platform &plat, bool quiet = true) const;
  // the state a decision leads to
  State apply(const Decision &d) const;
  // all successor states (equivalent siblings removed)
  std::vector<State> frontier(const Platform &plat, bool quiet = true) const;
  // Finish has been executed
  bool is_terminal() const;
  // equal keys <=> equivalent states (up to renaming streams/events)
  std::string canonical_key() const;
};

Equivalence get_equivalence(const State &a, const State &b);

} // namespace SDP
using SDP::State;
} // namespace tenzing
