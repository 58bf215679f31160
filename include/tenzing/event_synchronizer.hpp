// Decides which CUDA synchronisation a candidate op still needs, given the
// graph's edges and the sequence built so far.
//
// Rules (reference: include/tenzing/event_synchronizer.hpp:7-330; table in
// SURVEY §3.3) for a graph edge p -> v, with p already in the sequence:
//   Start -> any, CPU -> CPU, CPU -> GPU        nothing
//   GPU(s) -> GPU(s)                            nothing (stream order)
//   GPU(s1) -> GPU(s2)   CudaEventRecord(e,s1) after p, then CudaStreamWaitEvent(s2,e)
//   GPU(s1) -> CPU       CudaEventRecord(e,s1) after p, then CudaEventSync(e)
// make_syncs returns, per predecessor, only the NEXT missing op of its chain
// (first the record, on a later call the wait), so where the record and the
// wait are placed relative to other ready ops is itself part of the search.
#pragma once

#include <unordered_map>
#include <vector>

#include "tenzing/graph.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {

// What the synchronizer asks about a path, answered from tables built in one pass: where an
// op sits, which events were recorded on which stream and who waited for them afterwards.
// State::get_decisions builds one per call and shares it between all frontier ops (the plain
// overloads below build a temporary one); without it every question is a scan of the path and
// a decision costs O(frontier x predecessors x path^2).
class PathIndex {
public:
  struct Record {
    int pos = 0;
    Event event;
    bool hostSynced = false;     // a CudaEventSync on the event follows
    std::vector<int> waiters;    // streams with a CudaStreamWaitEvent on the event afterwards
    std::shared_ptr<CudaEventRecord> op;
  };
  PathIndex() = default;
  explicit PathIndex(const Sequence<BoundOp> &path) { extend(path); }
  // index the ops of `path` from position size() on: a path that has only grown at its end
  // since the last call costs its new ops, which is what a search that derives states from
  // one another pays (SDP::State keeps its index and hands a copy to its successors)
  void extend(const Sequence<BoundOp> &path);
  size_t size() const { return n_; }
  int pos_of(const OpBase &op) const;          // by identity; -1 if absent
  int pos_of_unbound(const OpBase &op) const;  // ignoring the stream binding; -1 if absent
  const std::vector<Record> &records(const Stream &s) const;
  // smallest event id no op of the path uses (= Sequence::new_unique_event)
  Event new_unique_event() const { return Event(nextEvent_); }

private:
  // paths are tens to a few hundred ops and an index is built for every state the search
  // visits: flat sorted arrays (no node allocations) instead of hash maps
  std::vector<std::pair<uint32_t, int>> byUid_, byUnbound_;       // sorted by id, first position wins
  std::vector<std::pair<int, std::vector<Record>>> byStream_;    // a handful of streams
  int nextEvent_ = 0;
  size_t n_ = 0; // ops indexed so far
  struct Latest {
    int event, stream;
    size_t idx;
  };
  std::vector<Latest> recordOf_; // the latest record of each event: (stream, index in its stream's list)
  std::vector<char> used_;       // event ids in use
  void append(const BoundOpPtr &op, int pos, bool sorted);
  static int lookup(const std::vector<std::pair<uint32_t, int>> &v, uint32_t id);
  std::vector<Record> &records_mut(int stream);
};

struct EventSynchronizer {
  // true iff every graph predecessor of `bo` found in `path` is synchronised with it
  static bool is_synced(const BoundOpPtr &bo, const Graph<OpBase> &g,
                        const Sequence<BoundOp> &path);
  static bool is_synced(const BoundOpPtr &bo, const Graph<OpBase> &g,
                        const Sequence<BoundOp> &path, const PathIndex &idx);
  static std::vector<BoundOpPtr> make_syncs(const BoundOpPtr &bo, const Graph<OpBase> &g,
                                            const Sequence<BoundOp> &path, const PathIndex &idx,
                                            bool quiet = true);
  // next missing synchronisation ops (deduplicated); empty iff is_synced
  static std::vector<BoundOpPtr> make_syncs(const BoundOpPtr &bo, const Graph<OpBase> &g,
                                            const Sequence<BoundOp> &path, bool quiet = true);

  // building blocks, exposed like the reference's
  static bool is_synced_gpu_then_cpu(const std::shared_ptr<BoundGpuOp> &a,
                                     const Sequence<BoundOp> &path);
  static bool is_synced_gpu_then_gpu(const std::shared_ptr<BoundGpuOp> &a,
                                     const std::shared_ptr<BoundGpuOp> &b,
                                     const Sequence<BoundOp> &path);
  static BoundOpPtr make_sync_gpu_then_cpu(const std::shared_ptr<BoundGpuOp> &a,
                                           const std::shared_ptr<CpuOp> &b,
                                           const Sequence<BoundOp> &path);
  static BoundOpPtr make_sync_gpu_then_gpu(const std::shared_ptr<BoundGpuOp> &a,
                                           const std::shared_ptr<BoundGpuOp> &b,
                                           const Sequence<BoundOp> &path);
};

using Synchronizer = EventSynchronizer;

} // namespace tenzing
