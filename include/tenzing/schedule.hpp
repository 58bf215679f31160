// Post-processing of complete sequences and the legacy "all orders" enumerators.
//
// Reference: include/tenzing/schedule.hpp:15-51, src/schedule.cpp:19-658.
// remove_redundant_syncs is a fix-point peephole pass over a sequence:
//   R1 drop a CudaEventRecord whose event nobody waits on later;
//   R2 drop a CudaStreamWaitEvent with no later GPU op in its stream;
//   R3 of two StreamSyncs of one stream with no GPU op of that stream between, drop the first;
//   R4 two records of the SAME point of a stream, both host-synced: the host
//      sync that comes second is redundant -> drop that record+sync pair;
//   R5 records r1 then r2 in one stream, sync(r2) before sync(r1): r1's pair is redundant;
//   R6 a second CudaEventSync of an event already synced; a CudaStreamWaitEvent
//      on the stream that recorded the event with no GPU op in between.
// None of the rules can invalidate a sequence produced by the EventSynchronizer.
#pragma once

#include <cstdint>
#include <vector>

#include "tenzing/graph.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {

class Schedule {
public:
  std::vector<BoundOpPtr> remaining; // ops not yet placed (enumerators)
  Sequence<BoundOp> order;           // ops placed so far

  void run(Platform &plat) {
    for (auto &op : order) op->run(plat);
  }
  int remove_redundant_syncs() { return remove_redundant_syncs(order); }
  // returns the number of ops removed
  static int remove_redundant_syncs(Sequence<BoundOp> &order);
  // true if a and b are the same schedule up to a renaming of streams/events
  static bool predicate(const Schedule &a, const Schedule &b);
  // strict weak order on op kind, then key (stable sort key for op lists)
  static bool by_op_typeid(const BoundOpPtr &a, const BoundOpPtr &b);
};

// every topological order of a graph whose vertices are all BoundOps
// (maxCount > 0 stops early)
std::vector<Schedule> make_schedules(const Graph<OpBase> &g, size_t maxCount = 0);
// n random topological orders; each step picks a ready op with probability
// proportional to the number of complete orders that continue through it
// (exact counting up to 62 vertices, uniform choice beyond)
std::vector<Schedule> make_schedules_random(const Graph<OpBase> &g, size_t n, uint64_t seed = 0);
// number of topological orders (linear extensions); saturates at UINT64_MAX
uint64_t count_schedules(const Graph<OpBase> &g);

} // namespace tenzing
