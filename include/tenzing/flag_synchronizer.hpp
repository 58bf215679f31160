// Cross-rank synchronisation through device-side flags, as seen by the scheduler.
//
// The reference synchronises ranks with MPI_Wait on the host; the EventSynchronizer never
// sees those dependencies.  Here a consumer kernel spins on a flag word that a producer
// kernel of ANOTHER rank publishes.  All ranks execute the same sequence, so the schedule is
// deadlock free iff every op that waits for a flag channel comes after an op that publishes
// that channel: by induction over sequence positions every publish a wait depends on has an
// earlier position on every rank.  A channel is the byte offset of the flag word inside the
// symmetric heap (the producer writes word X of the peer, the consumer polls its own word X).
// Where the word is indexed by the SOURCE rank (SpMV: rank s publishes word s on everybody), an
// op declares the words of all ranks: what the same op publishes on the other ranks is what
// arrives here.
//
// Ops declare their channels through HasFlags; FlagSynchronizer checks sequences, derives the
// publish -> wait graph edges, and can filter decisions so a search never offers a wait
// before its publish.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "tenzing/graph.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {

class HasFlags {
public:
  virtual ~HasFlags() = default;
  // heap offsets of the flag words this op publishes on its peers / polls locally
  virtual std::vector<uint64_t> flags_published() const = 0;
  virtual std::vector<uint64_t> flags_awaited() const = 0;
  // true for single kernels that publish first and wait afterwards (e.g. the one-launch SpMV):
  // their own publishes count as "earlier" for their own waits
  virtual bool publishes_before_waiting() const { return false; }
};

struct FlagSynchronizer {
  struct Violation {
    size_t position;      // index of the waiting op in the sequence
    std::string op;       // its name
    uint64_t channel;     // the flag nobody published before it
  };
  // every awaited channel has been published by an earlier op of the sequence
  static std::vector<Violation> check(const Sequence<BoundOp> &seq);
  static bool is_deadlock_free(const Sequence<BoundOp> &seq) { return check(seq).empty(); }
  // add an edge publisher -> waiter for every shared channel (idempotent); returns #edges added
  static int add_edges(Graph<OpBase> &g);
  // true iff every channel `op` waits for is published by some op already in `path`
  static bool is_ready(const BoundOp &op, const Sequence<BoundOp> &path);
};

} // namespace tenzing
