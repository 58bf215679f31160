// A state of the sequential decision process: the (constrained) program graph
// plus the executable sequence chosen so far.
//
// Reference: include/tenzing/state.hpp:15-54, src/state.cpp:5-143.
// The graph is held through a shared pointer, so the common decision
// (ExecuteOp) copies the sequence and a few small tables: what a state has
// already worked out about its path (digest, executed and synchronised
// vertices, the synchronizer's index) goes to its successors, which then
// only account for the op they add.
#pragma once

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "tenzing/decision.hpp"
#include "tenzing/event_synchronizer.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {
namespace SDP {

// 128-bit digest of a state's canonical form: equal for equivalent states, and different for
// different ones up to a collision chance of about 2^-100 for the millions of states a search
// visits.  What the solvers de-duplicate on; canonical_key() is the readable long form.
struct StateKey {
  uint64_t a = 0, b = 0;
  bool operator==(const StateKey &o) const { return a == o.a && b == o.b; }
  bool operator!=(const StateKey &o) const { return !(*this == o); }
};
struct StateKeyHasher {
  size_t operator()(const StateKey &k) const { return size_t(k.a ^ (k.b * 0x9E3779B97F4A7C15ull)); }
};

namespace detail {
// running digest of a token stream plus the stream/event renumbering it has seen so far
struct KeyDigest {
  uint64_t a = 0x243F6A8885A308D3ull, b = 0x13198A2E03707344ull;
  std::vector<int> streams, events; // id at position i was the i-th distinct one seen
  void mix(uint64_t v) {
    a = (a ^ v) * 0xFF51AFD7ED558CCDull;
    a ^= a >> 32;
    b = (b + v) * 0xC4CEB9FE1A85EC53ull;
    b ^= b >> 29;
  }
  void token(char tag, uint32_t v) { mix((uint64_t(uint8_t(tag)) << 32) | v); }
  static uint32_t rank_of(std::vector<int> &seen, int id) {
    for (size_t i = 0; i < seen.size(); ++i)
      if (seen[i] == id) return uint32_t(i);
    seen.push_back(id);
    return uint32_t(seen.size() - 1);
  }
  void stream(int id) { token('s', rank_of(streams, id)); }
  void event(int id) { token('e', rank_of(events, id)); }
  // kind/identity of the op, then its streams and events renumbered by first use
  void op(const OpBase &o);
};
// the graph half of the key, prepared once per graph object: vertices in a binding-independent
// order with their successor lists; stream/event ids are stored raw and renumbered at digest time
struct GraphTokens {
  std::vector<uint64_t> tokens; // (tag << 32) | value; tags 's' / 'e' carry raw ids
  bool hasResources = false;    // false: the digest of `tokens` does not depend on the sequence
  StateKey plain;               // that digest, when !hasResources
};
} // namespace detail

class State {
  std::shared_ptr<const Graph<OpBase>> graph_;
  Sequence<BoundOp> sequence_;
  mutable std::string keyCache_; // canonical_key(), built on request
  // canonical_hash(): the sequence half is extended op by op as states are derived from each
  // other, the graph half is prepared once per graph and shared by every state holding it
  mutable detail::KeyDigest seqDigest_;
  mutable size_t seqDigested_ = 0;
  mutable std::shared_ptr<const detail::GraphTokens> graphTokens_;
  // which vertices the sequence has executed, by vertex number; extended like the digest
  mutable std::vector<char> done_;
  mutable size_t doneCounted_ = 0;
  // what the synchronizer asks about the sequence, extended the same way
  mutable PathIndex pathIndex_;
  // frontier vertices already found synchronised with their predecessors (see get_decisions)
  mutable std::vector<char> synced_;

  // same sequence (and its digest so far), other graph
  State with_graph(Graph<OpBase> &&g) const;

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
  // digest of the same canonical form; what frontier() and the DFS de-duplicate on
  StateKey canonical_hash() const;
  // digest of the sequence alone (the sequence half of canonical_hash)
  StateKey sequence_hash() const;
};

// digest of a sequence's canonical form (same tokens as State::sequence_hash)
StateKey sequence_hash(const Sequence<BoundOp> &seq);

Equivalence get_equivalence(const State &a, const State &b);

} // namespace SDP
using SDP::State;
using SDP::StateKey;
using SDP::StateKeyHasher;
using SDP::sequence_hash;
} // namespace tenzing
