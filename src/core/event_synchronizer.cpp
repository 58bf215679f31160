#include "tenzing/event_synchronizer.hpp"

#include <algorithm>
#include <typeinfo>

namespace tenzing {

// ---------------------------------------------------------------- PathIndex
void PathIndex::append(const BoundOpPtr &op, int pos, bool sorted) {
  auto put = [sorted](std::vector<std::pair<uint32_t, int>> &v, uint32_t id, int p) {
    // (id, position) order: among equal ids the earliest position comes first, the one lookups return
    if (sorted) v.insert(std::upper_bound(v.begin(), v.end(), std::make_pair(id, p)), std::make_pair(id, p));
    else v.emplace_back(id, p);
  };
  put(byUid_, op->uid(), pos);
  put(byUnbound_, op->unbound_uid(), pos);
  const uint16_t t = op->traits();
  if (!(t & op_trait::WithEvent)) return;
  auto latest = [this](int event) -> Latest * {
    for (auto &r : recordOf_)
      if (r.event == event) return &r;
    return nullptr;
  };
  auto mark = [this](const Event &e) {
    if (e.id() < 0) return;
    if (size_t(e.id()) >= used_.size()) used_.resize(size_t(e.id()) + 1, 0);
    used_[size_t(e.id())] = 1;
  };
  if (t & op_trait::EventRecord) {
    auto cer = std::static_pointer_cast<CudaEventRecord>(op);
    mark(cer->event());
    auto &v = records_mut(cer->stream().id());
    Record r;
    r.pos = pos;
    r.event = cer->event();
    r.op = cer;
    if (Latest *l = latest(cer->event().id())) {
      l->stream = cer->stream().id();
      l->idx = v.size();
    } else {
      recordOf_.push_back(Latest{cer->event().id(), cer->stream().id(), v.size()});
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

void PathIndex::extend(const Sequence<BoundOp> &path) {
  if (path.size() <= n_) return;
  // a few new ops are inserted in place; a long stretch is appended and sorted once
  const bool bulk = path.size() - n_ > 8;
  if (bulk) {
    byUid_.reserve(path.size());
    byUnbound_.reserve(path.size());
  }
  for (; n_ < path.size(); ++n_) append(path[n_], int(n_), !bulk);
  if (bulk) {
    std::sort(byUid_.begin(), byUid_.end());
    std::sort(byUnbound_.begin(), byUnbound_.end());
  }
  while (size_t(nextEvent_) < used_.size() && used_[size_t(nextEvent_)]) ++nextEvent_;
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
const std::shared_ptr<CudaEventRecord> *first_record(const PathIndex &idx, int from, const Stream &stream) {
  for (const auto &r : idx.records(stream))
    if (r.pos >= from) return &r.op;
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

// Runs once per (frontier op, predecessor) pair of every decision: cached type bits, no casts,
// no reference counting.
Need classify(const BoundOp &pred, const BoundOp &bo) {
  const uint16_t pt = pred.traits(), bt = bo.traits();
  const bool bIsGpu = (bt & op_trait::BoundGpu) != 0;
  if (!(pt & op_trait::BoundGpu)) {
    if (pt & op_trait::IsStart) return Need::None;
    if ((pt & op_trait::Cpu) && (bIsGpu || (bt & op_trait::Cpu))) return Need::None; // the host issues in order
    THROW_RUNTIME("unexpected op combination: " << pred.desc() << " -> " << bo.desc());
  }
  if (bIsGpu) {
    // a flag-driven consumer only needs its GPU predecessors issued earlier, which the
    // sequence order already guarantees
    if (static_cast<const BoundGpuOp &>(bo).unbound()->issue_order_only()) return Need::None;
    return Need::GpuThenGpu;
  }
  if (bt & op_trait::Cpu) return Need::GpuThenCpu;
  THROW_RUNTIME("unexpected op combination: " << pred.desc() << " -> " << bo.desc());
}

// fn(pred) for every graph predecessor of bo, as the bound op found in the path; stops when fn
// returns false.  Walks vertex numbers: no op lists are copied.
template <typename Fn>
void each_pred_in_path(const BoundOpPtr &bo, const Graph<OpBase> &g, const Path &path,
                       const PathIndex &idx, Fn &&fn) {
  const int v = g.find_vertex_unbound(*bo);
  if (v < 0) THROW_RUNTIME("couldn't find " << bo->desc() << " in graph");
  for (int p : g.pred_ids(v)) {
    const OpPtr &gp = g.op_at(p);
    const int i = idx.pos_of_unbound(*gp);
    if (i < 0)
      THROW_RUNTIME("couldn't find " << gp->desc() << " in path (pred of " << bo->desc() << ")");
    if (!fn(path[size_t(i)])) return;
  }
}

bool synced_gpu_cpu(const BoundGpuOp &a, const PathIndex &idx) {
  return host_waited(idx, index_of(idx, a), a.stream());
}

bool synced_gpu_gpu(const BoundGpuOp &a, const BoundGpuOp &b, const PathIndex &idx) {
  if (a.stream() == b.stream()) return true;
  return stream_waited(idx, index_of(idx, a), a.stream(), b.stream());
}

// `aPtr` / `bPtr` own a and b; they are only needed to name a new op after its neighbours
BoundOpPtr sync_gpu_cpu(const BoundOpPtr &aPtr, const BoundOpPtr &bPtr, const PathIndex &idx) {
  const auto &a = static_cast<const BoundGpuOp &>(*aPtr);
  const int ai = index_of(idx, a);
  if (host_waited(idx, ai, a.stream())) return nullptr;
  if (const auto *cer = first_record(idx, ai, a.stream())) {
    auto ces = std::make_shared<CudaEventSync>((*cer)->event());
    ces->update_name({}, {bPtr});
    return ces;
  }
  auto cer = std::make_shared<CudaEventRecord>(idx.new_unique_event(), a.stream());
  cer->update_name({aPtr}, {});
  return cer;
}

BoundOpPtr sync_gpu_gpu(const BoundOpPtr &aPtr, const BoundOpPtr &bPtr, const PathIndex &idx) {
  const auto &a = static_cast<const BoundGpuOp &>(*aPtr);
  const auto &b = static_cast<const BoundGpuOp &>(*bPtr);
  if (a.stream() == b.stream()) return nullptr;
  const int ai = index_of(idx, a);
  if (stream_waited(idx, ai, a.stream(), b.stream())) return nullptr;
  if (const auto *cer = first_record(idx, ai, a.stream())) {
    auto w = std::make_shared<CudaStreamWaitEvent>(b.stream(), (*cer)->event());
    w->update_name({aPtr}, {});
    return w;
  }
  auto cer = std::make_shared<CudaEventRecord>(idx.new_unique_event(), a.stream());
  cer->update_name({}, {bPtr});
  return cer;
}

} // namespace

bool EventSynchronizer::is_synced_gpu_then_cpu(const std::shared_ptr<BoundGpuOp> &a,
                                               const Path &path) {
  return synced_gpu_cpu(*a, PathIndex(path));
}

bool EventSynchronizer::is_synced_gpu_then_gpu(const std::shared_ptr<BoundGpuOp> &a,
                                               const std::shared_ptr<BoundGpuOp> &b,
                                               const Path &path) {
  if (a->stream() == b->stream()) return true;
  return synced_gpu_gpu(*a, *b, PathIndex(path));
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
  bool ok = true;
  each_pred_in_path(bo, g, path, idx, [&](const BoundOpPtr &pred) {
    const Need need = classify(*pred, *bo);
    if (need == Need::GpuThenCpu)
      ok = synced_gpu_cpu(static_cast<const BoundGpuOp &>(*pred), idx);
    else if (need == Need::GpuThenGpu)
      ok = synced_gpu_gpu(static_cast<const BoundGpuOp &>(*pred), static_cast<const BoundGpuOp &>(*bo), idx);
    return ok;
  });
  return ok;
}

std::vector<BoundOpPtr> EventSynchronizer::make_syncs(const BoundOpPtr &bo, const Graph<OpBase> &g,
                                                      const Path &path, bool quiet) {
  return make_syncs(bo, g, path, PathIndex(path), quiet);
}

std::vector<BoundOpPtr> EventSynchronizer::make_syncs(const BoundOpPtr &bo, const Graph<OpBase> &g,
                                                      const Path &path, const PathIndex &idx,
                                                      bool quiet) {
  std::vector<BoundOpPtr> syncs;
  each_pred_in_path(bo, g, path, idx, [&](const BoundOpPtr &pred) {
    const Need need = classify(*pred, *bo);
    BoundOpPtr s;
    if (need == Need::GpuThenCpu) s = sync_gpu_cpu(pred, bo, idx);
    else if (need == Need::GpuThenGpu) s = sync_gpu_gpu(pred, bo, idx);
    if (s) {
      if (!quiet) TENZING_INFO("sync " << s->desc() << " for " << pred->desc() << " -> " << bo->desc());
      // several predecessors on one stream ask for the same op: keep the first
      bool have = false;
      for (const BoundOpPtr &o : syncs) have = have || o->uid() == s->uid();
      if (!have) syncs.push_back(std::move(s));
    }
    return true;
  });
  return syncs;
}

} // namespace tenzing
