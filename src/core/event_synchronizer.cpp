#include "tenzing/event_synchronizer.hpp"

#include <algorithm>
#include <typeinfo>

namespace tenzing {

// ---------------------------------------------------------------- PathIndex
PathIndex::PathIndex(const Sequence<BoundOp> &path) {
  struct Latest { int event, stream; size_t idx; };
  std::vector<Latest> recordOf; // the latest record of each event: (stream, index in its stream's list)
  auto latest = [&recordOf](int event) -> Latest * {
    for (auto &r : recordOf)
      if (r.event == event) return &r;
    return nullptr;
  };
  std::vector<char> used;
  auto mark = [&](const Event &e) {
    if (e.id() < 0) return;
    if (size_t(e.id()) >= used.size()) used.resize(size_t(e.id()) + 1, 0);
    used[size_t(e.id())] = 1;
  };
  byUid_.reserve(path.size());
  byUnbound_.reserve(path.size());
  for (size_t i = 0; i < path.size(); ++i) {
    const BoundOpPtr &op = path[i];
    byUid_.emplace_back(op->uid(), int(i));
    byUnbound_.emplace_back(op->unbound_uid(), int(i));
    const uint16_t t = op->traits();
    if (!(t & op_trait::WithEvent)) continue;
    if (t & op_trait::EventRecord) {
      auto cer = std::static_pointer_cast<CudaEventRecord>(op);
      mark(cer->event());
      auto &v = records_mut(cer->stream().id());
      Record r;
      r.pos = int(i);
      r.event = cer->event();
      r.op = cer;
      if (Latest *l = latest(cer->event().id())) {
        l->stream = cer->stream().id();
        l->idx = v.size();
      } else {
        recordOf.push_back(Latest{cer->event().id(), cer->stream().id(), v.size()});
      }
      v.push_back(std::move(r));
    } else if (t & op_trait::EventSync) {
      const auto &ces = static_cast<const CudaEventSync &>(*op);
      mark(ces.event());
      if (const Latest *l = latest(ces.event().id())) records_mut(l->stream)[l->idx].hostSynced = true;
    } else if (t & op_trait::StreamWaitEvent) {
      const auto &w = static_cast<const CudaStreamWaitEvent &>(*op);
      mark(w.event());
      if (const Latest *l = latest(w.event().id()))
        records_mut(l->stream)[l->idx].waiters.push_back(w.stream().id());
    } else if (const auto *he = dynamic_cast<const HasEvent *>(op.get())) {
      for (const Event &e : he->get_events()) mark(e);
    }
  }
  // (id, position) order: among equal ids the earliest position comes first, the one lookups return
  std::sort(byUid_.begin(), byUid_.end());
  std::sort(byUnbound_.begin(), byUnbound_.end());
  while (size_t(nextEvent_) < used.size() && used[size_t(nextEvent_)]) ++nextEvent_;
}

std::vector<PathIndex::Record> &PathIndex::records_mut(int stream) {
  for (auto &kv : byStream_)
    if (kv.first == stream) return kv.second;
  byStream_.emplace_back(stream, std::vector<Record>{});
  return byStream_.back().second;
}

int PathIndex::lookup(const std::vector<std::pair<uint32_t, int>> &v, uint32_t id) {
  auto it = std::lower_bound(v.begin(), v.end(), id,
                             [](const std::pair<uint32_t, int> &x, uint32_t key) { return x.first < key; });
  return it == v.end() || it->first != id ? -1 : it->second;
}

int PathIndex::pos_of(const OpBase &op) const { return lookup(byUid_, op.uid()); }

int PathIndex::pos_of_unbound(const OpBase &op) const { return lookup(byUnbound_, op.unbound_uid()); }

const std::vector<PathIndex::Record> &PathIndex::records(const Stream &s) const {
  static const std::vector<Record> none;
  for (const auto &kv : byStream_)
    if (kv.first == s.id()) return kv.second;
  return none;
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
  // runs once per (frontier op, predecessor) pair of every decision: cached type bits, no casts
  const uint16_t pt = pred->traits(), bt = bo->traits();
  const bool bIsGpu = (bt & op_trait::BoundGpu) != 0;
  if (!(pt & op_trait::BoundGpu)) {
    if (pt & op_trait::IsStart) return r;
    if ((pt & op_trait::Cpu) && (bIsGpu || (bt & op_trait::Cpu))) return r; // CPU -> anything: the host issues in order
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
  if (bt & op_trait::Cpu) {
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
