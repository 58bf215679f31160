#include "tenzing/event_synchronizer.hpp"

#include <typeinfo>

namespace tenzing {

// ---------------------------------------------------------------- PathIndex
PathIndex::PathIndex(const Sequence<BoundOp> &path) {
  std::unordered_map<int, std::pair<int, size_t>> recordOf; // event -> (stream, index in byStream_)
  std::vector<char> used;
  auto mark = [&](const Event &e) {
    if (e.id() < 0) return;
    if (size_t(e.id()) >= used.size()) used.resize(size_t(e.id()) + 1, 0);
    used[size_t(e.id())] = 1;
  };
  for (size_t i = 0; i < path.size(); ++i) {
    const BoundOpPtr &op = path[i];
    byUid_.emplace(op->uid(), int(i));
    byUnbound_.emplace(op->unbound_uid(), int(i));
    if (auto he = std::dynamic_pointer_cast<HasEvent>(op))
      for (const Event &e : he->get_events()) mark(e);
    if (auto cer = std::dynamic_pointer_cast<CudaEventRecord>(op)) {
      auto &v = byStream_[cer->stream().id()];
      Record r;
      r.pos = int(i);
      r.event = cer->event();
      r.op = cer;
      recordOf[cer->event().id()] = {cer->stream().id(), v.size()}; // the latest record of the event
      v.push_back(std::move(r));
    } else if (auto ces = std::dynamic_pointer_cast<CudaEventSync>(op)) {
      auto it = recordOf.find(ces->event().id());
      if (it != recordOf.end()) byStream_[it->second.first][it->second.second].hostSynced = true;
    } else if (auto w = std::dynamic_pointer_cast<CudaStreamWaitEvent>(op)) {
      auto it = recordOf.find(w->event().id());
      if (it != recordOf.end())
        byStream_[it->second.first][it->second.second].waiters.push_back(w->stream().id());
    }
  }
  while (size_t(nextEvent_) < used.size() && used[size_t(nextEvent_)]) ++nextEvent_;
}

int PathIndex::pos_of(const OpBase &op) const {
  auto it = byUid_.find(op.uid());
  return it == byUid_.end() ? -1 : it->second;
}

int PathIndex::pos_of_unbound(const OpBase &op) const {
  auto it = byUnbound_.find(op.unbound_uid());
  return it == byUnbound_.end() ? -1 : it->second;
}

const std::vector<PathIndex::Record> &PathIndex::records(const Stream &s) const {
  static const std::vector<Record> none;
  auto it = byStream_.find(s.id());
  return it == byStream_.end() ? none : it->second;
}

namespace {

using Path = Sequence<BoundOp>;

int index_of(const PathIndex &idx, const OpBase &a) {
  const int i = idx.pos_of(a);
  if (i < 0) THROW_RUNTIME("couldn't find " << a.desc() << " in path");
  return i;
}

// first CudaEventRecord on `stream` at or after position `from`
std::shared_ptr<CudaEventRecord> first_record(const PathIndex &idx, int from, const Stream &stream) {
  for (const auto &r : idx.records(stream))
    if (r.pos >= from) return r.op;
  return nullptr;
}

// does some record on a's stream after a have a later CudaEventSync?
bool host_waited(const PathIndex &idx, int ai, const Stream &stream) {
  for (const auto &r : idx.records(stream))
    if (r.pos >= ai && r.hostSynced) return true;
  return false;
}

// does some record on a's stream after a have a later CudaStreamWaitEvent on `waiter`?
bool stream_waited(const PathIndex &idx, int ai, const Stream &stream, const Stream &waiter) {
  for (const auto &r : idx.records(stream)) {
    if (r.pos < ai) continue;
    for (int w : r.waiters)
      if (w == waiter.id()) return true;
  }
  return false;
}

enum class Need { None, GpuThenCpu, GpuThenGpu };

struct PredInfo {
  Need need = Need::None;
  std::shared_ptr<BoundGpuOp> pGpu;
};

PredInfo classify(const BoundOpPtr &pred, const BoundOpPtr &bo) {
  PredInfo r;
  // BoundGpuOp and Start are final classes: an exact type test (one pointer comparison) is
  // enough and this function runs once per (frontier op, predecessor) pair of every decision
  const bool pIsGpu = typeid(*pred) == typeid(BoundGpuOp);
  const bool bIsGpu = typeid(*bo) == typeid(BoundGpuOp);
  if (!pIsGpu) {
    if (typeid(*pred) == typeid(Start)) return r;
    if (dynamic_cast<const CpuOp *>(pred.get()) && (bIsGpu || dynamic_cast<const CpuOp *>(bo.get())))
      return r; // CPU -> anything: the host issues in order
    THROW_RUNTIME("unexpected op combination: " << pred->desc() << " -> " << bo->desc());
  }
  r.pGpu = std::static_pointer_cast<BoundGpuOp>(pred);
  if (bIsGpu) {
    // a flag-driven consumer only needs its GPU predecessors issued earlier, which the
    // sequence order already guarantees
    if (static_cast<const BoundGpuOp &>(*bo).unbound()->issue_order_only()) return r;
    r.need = Need::GpuThenGpu;
    return r;
  }
  if (dynamic_cast<const CpuOp *>(bo.get())) {
    r.need = Need::GpuThenCpu;
    return r;
  }
  THROW_RUNTIME("unexpected op combination: " << pred->desc() << " -> " << bo->desc());
}

// graph predecessors of bo, as the bound ops found in the path
std::vector<BoundOpPtr> preds_in_path(const BoundOpPtr &bo, const Graph<OpBase> &g, const Path &path,
                                      const PathIndex &idx) {
  auto found = g.preds_find_or_find_unbound(bo);
  if (!found.first) THROW_RUNTIME("couldn't find " << bo->desc() << " in graph");
  std::vector<BoundOpPtr> r;
  for (const auto &gp : found.second) {
    const int i = idx.pos_of_unbound(*gp);
    if (i < 0)
      THROW_RUNTIME("couldn't find " << gp->desc() << " in path (pred of " << bo->desc() << ")");
    r.push_back(path[size_t(i)]);
  }
  return r;
}

bool synced_gpu_cpu(const std::shared_ptr<BoundGpuOp> &a, const PathIndex &idx) {
  return host_waited(idx, index_of(idx, *a), a->stream());
}

bool synced_gpu_gpu(const std::shared_ptr<BoundGpuOp> &a, const std::shared_ptr<BoundGpuOp> &b,
                    const PathIndex &idx) {
  if (a->stream() == b->stream()) return true;
  return stream_waited(idx, index_of(idx, *a), a->stream(), b->stream());
}

BoundOpPtr sync_gpu_cpu(const std::shared_ptr<BoundGpuOp> &a, const std::shared_ptr<CpuOp> &b,
                        const PathIndex &idx) {
  const int ai = index_of(idx, *a);
  if (host_waited(idx, ai, a->stream())) return nullptr;
  if (auto cer = first_record(idx, ai, a->stream())) {
    auto ces = std::make_shared<CudaEventSync>(cer->event());
    ces->update_name({}, {b});
    return ces;
  }
  auto cer = std::make_shared<CudaEventRecord>(idx.new_unique_event(), a->stream());
  cer->update_name({a}, {});
  return cer;
}

BoundOpPtr sync_gpu_gpu(const std::shared_ptr<BoundGpuOp> &a, const std::shared_ptr<BoundGpuOp> &b,
                        const PathIndex &idx) {
  if (a->stream() == b->stream()) return nullptr;
  const int ai = index_of(idx, *a);
  if (stream_waited(idx, ai, a->stream(), b->stream())) return nullptr;
  if (auto cer = first_record(idx, ai, a->stream())) {
    auto w = std::make_shared<CudaStreamWaitEvent>(b->stream(), cer->event());
    w->update_name({a}, {});
    return w;
  }
  auto cer = std::make_shared<CudaEventRecord>(idx.new_unique_event(), a->stream());
  cer->update_name({}, {b});
  return cer;
}

} // namespace

bool EventSynchronizer::is_synced_gpu_then_cpu(const std::shared_ptr<BoundGpuOp> &a,
                                               const Path &path) {
  return synced_gpu_cpu(a, PathIndex(path));
}

bool EventSynchronizer::is_synced_gpu_then_gpu(const std::shared_ptr<BoundGpuOp> &a,
                                               const std::shared_ptr<BoundGpuOp> &b,
                                               const Path &path) {
  if (a->stream() == b->stream()) return true;
  return synced_gpu_gpu(a, b, PathIndex(path));
}

BoundOpPtr EventSynchronizer::make_sync_gpu_then_cpu(const std::shared_ptr<BoundGpuOp> &a,
                                                     const std::shared_ptr<CpuOp> &b,
                                                     const Path &path) {
  return sync_gpu_cpu(a, b, PathIndex(path));
}

BoundOpPtr EventSynchronizer::make_sync_gpu_then_gpu(const std::shared_ptr<BoundGpuOp> &a,
                                                     const std::shared_ptr<BoundGpuOp> &b,
                                                     const Path &path) {
  return sync_gpu_gpu(a, b, PathIndex(path));
}

bool EventSynchronizer::is_synced(const BoundOpPtr &bo, const Graph<OpBase> &g, const Path &path) {
  return is_synced(bo, g, path, PathIndex(path));
}

bool EventSynchronizer::is_synced(const BoundOpPtr &bo, const Graph<OpBase> &g, const Path &path,
                                  const PathIndex &idx) {
  for (const BoundOpPtr &pred : preds_in_path(bo, g, path, idx)) {
    const PredInfo info = classify(pred, bo);
    if (info.need == Need::GpuThenCpu) {
      if (!synced_gpu_cpu(info.pGpu, idx)) return false;
    } else if (info.need == Need::GpuThenGpu) {
      if (!synced_gpu_gpu(info.pGpu, std::static_pointer_cast<BoundGpuOp>(bo), idx)) return false;
    }
  }
  return true;
}

std::vector<BoundOpPtr> EventSynchronizer::make_syncs(const BoundOpPtr &bo, const Graph<OpBase> &g,
                                                      const Path &path, bool quiet) {
  return make_syncs(bo, g, path, PathIndex(path), quiet);
}

std::vector<BoundOpPtr> EventSynchronizer::make_syncs(const BoundOpPtr &bo, const Graph<OpBase> &g,
                                                      const Path &path, const PathIndex &idx,
                                                      bool quiet) {
  std::vector<BoundOpPtr> syncs;
  for (const BoundOpPtr &pred : preds_in_path(bo, g, path, idx)) {
    const PredInfo info = classify(pred, bo);
    BoundOpPtr s;
    if (info.need == Need::GpuThenCpu) {
      s = sync_gpu_cpu(info.pGpu, std::static_pointer_cast<CpuOp>(bo), idx);
    } else if (info.need == Need::GpuThenGpu) {
      s = sync_gpu_gpu(info.pGpu, std::static_pointer_cast<BoundGpuOp>(bo), idx);
    }
    if (s) {
      if (!quiet) TENZING_INFO("sync " << s->desc() << " for " << pred->desc() << " -> " << bo->desc());
      syncs.push_back(s);
    }
  }
  keep_uniques(syncs);
  return syncs;
}

} // namespace tenzing
