#include "tenzing/event_synchronizer.hpp"

namespace tenzing {

namespace {

using Path = Sequence<BoundOp>;

size_t index_of(const Path &path, const OpBase &a) {
  for (size_t i = 0; i < path.size(); ++i)
    if (path[i]->uid() == a.uid()) return i;
  THROW_RUNTIME("couldn't find " << a.desc() << " in path");
}

// first CudaEventRecord on `stream` at or after position `from`
std::shared_ptr<CudaEventRecord> first_record(const Path &path, size_t from, const Stream &stream,
                                              size_t *pos = nullptr) {
  for (size_t i = from; i < path.size(); ++i) {
    if (auto cer = std::dynamic_pointer_cast<CudaEventRecord>(path[i])) {
      if (cer->stream() == stream) {
        if (pos) *pos = i;
        return cer;
      }
    }
  }
  return nullptr;
}

// does some record on a's stream after a have a later CudaEventSync?
bool host_waited(const Path &path, size_t ai, const Stream &stream) {
  for (size_t i = ai; i < path.size(); ++i) {
    auto cer = std::dynamic_pointer_cast<CudaEventRecord>(path[i]);
    if (!cer || !(cer->stream() == stream)) continue;
    for (size_t k = i + 1; k < path.size(); ++k)
      if (auto ces = std::dynamic_pointer_cast<CudaEventSync>(path[k]))
        if (ces->event() == cer->event()) return true;
  }
  return false;
}

// does some record on a's stream after a have a later CudaStreamWaitEvent on `waiter`?
bool stream_waited(const Path &path, size_t ai, const Stream &stream, const Stream &waiter) {
  for (size_t i = ai; i < path.size(); ++i) {
    auto cer = std::dynamic_pointer_cast<CudaEventRecord>(path[i]);
    if (!cer || !(cer->stream() == stream)) continue;
    for (size_t k = i + 1; k < path.size(); ++k)
      if (auto w = std::dynamic_pointer_cast<CudaStreamWaitEvent>(path[k]))
        if (w->event() == cer->event() && w->stream() == waiter) return true;
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
  if (std::dynamic_pointer_cast<Start>(pred)) return r;
  const bool bCpu = bool(std::dynamic_pointer_cast<CpuOp>(bo));
  const bool bGpu = bool(std::dynamic_pointer_cast<BoundGpuOp>(bo));
  const bool pCpu = bool(std::dynamic_pointer_cast<CpuOp>(pred));
  r.pGpu = std::dynamic_pointer_cast<BoundGpuOp>(pred);
  if (pCpu && (bCpu || bGpu)) return r;
  if (r.pGpu && bCpu) {
    r.need = Need::GpuThenCpu;
    return r;
  }
  if (r.pGpu && bGpu) {
    // a flag-driven consumer only needs its GPU predecessors issued earlier, which the
    // sequence order already guarantees
    if (std::static_pointer_cast<BoundGpuOp>(bo)->unbound()->issue_order_only()) return r;
    r.need = Need::GpuThenGpu;
    return r;
  }
  THROW_RUNTIME("unexpected op combination: " << pred->desc() << " -> " << bo->desc());
}

// graph predecessors of bo, as the bound ops found in the path
std::vector<BoundOpPtr> preds_in_path(const BoundOpPtr &bo, const Graph<OpBase> &g,
                                      const Path &path) {
  auto found = g.preds_find_or_find_unbound(bo);
  if (!found.first) THROW_RUNTIME("couldn't find " << bo->desc() << " in graph");
  std::vector<BoundOpPtr> r;
  for (const auto &gp : found.second) {
    auto it = path.find_unbound(*gp);
    if (it == path.end())
      THROW_RUNTIME("couldn't find " << gp->desc() << " in path (pred of " << bo->desc() << ")");
    r.push_back(*it);
  }
  return r;
}

} // namespace

bool EventSynchronizer::is_synced_gpu_then_cpu(const std::shared_ptr<BoundGpuOp> &a,
                                               const Path &path) {
  return host_waited(path, index_of(path, *a), a->stream());
}

bool EventSynchronizer::is_synced_gpu_then_gpu(const std::shared_ptr<BoundGpuOp> &a,
                                               const std::shared_ptr<BoundGpuOp> &b,
                                               const Path &path) {
  if (a->stream() == b->stream()) return true;
  return stream_waited(path, index_of(path, *a), a->stream(), b->stream());
}

BoundOpPtr EventSynchronizer::make_sync_gpu_then_cpu(const std::shared_ptr<BoundGpuOp> &a,
                                                     const std::shared_ptr<CpuOp> &b,
                                                     const Path &path) {
  const size_t ai = index_of(path, *a);
  if (host_waited(path, ai, a->stream())) return nullptr;
  if (auto cer = first_record(path, ai, a->stream())) {
    auto ces = std::make_shared<CudaEventSync>(cer->event());
    ces->update_name({}, {b});
    return ces;
  }
  auto cer = std::make_shared<CudaEventRecord>(path.new_unique_event(), a->stream());
  cer->update_name({a}, {});
  return cer;
}

BoundOpPtr EventSynchronizer::make_sync_gpu_then_gpu(const std::shared_ptr<BoundGpuOp> &a,
                                                     const std::shared_ptr<BoundGpuOp> &b,
                                                     const Path &path) {
  if (a->stream() == b->stream()) return nullptr;
  const size_t ai = index_of(path, *a);
  if (stream_waited(path, ai, a->stream(), b->stream())) return nullptr;
  if (auto cer = first_record(path, ai, a->stream())) {
    auto w = std::make_shared<CudaStreamWaitEvent>(b->stream(), cer->event());
    w->update_name({a}, {});
    return w;
  }
  auto cer = std::make_shared<CudaEventRecord>(path.new_unique_event(), a->stream());
  cer->update_name({}, {b});
  return cer;
}

bool EventSynchronizer::is_synced(const BoundOpPtr &bo, const Graph<OpBase> &g, const Path &path) {
  for (const BoundOpPtr &pred : preds_in_path(bo, g, path)) {
    const PredInfo info = classify(pred, bo);
    if (info.need == Need::GpuThenCpu) {
      if (!is_synced_gpu_then_cpu(info.pGpu, path)) return false;
    } else if (info.need == Need::GpuThenGpu) {
      if (!is_synced_gpu_then_gpu(info.pGpu, std::static_pointer_cast<BoundGpuOp>(bo), path))
        return false;
    }
  }
  return true;
}

std::vector<BoundOpPtr> EventSynchronizer::make_syncs(const BoundOpPtr &bo, const Graph<OpBase> &g,
                                                      const Path &path, bool quiet) {
  std::vector<BoundOpPtr> syncs;
  for (const BoundOpPtr &pred : preds_in_path(bo, g, path)) {
    const PredInfo info = classify(pred, bo);
    BoundOpPtr s;
    if (info.need == Need::GpuThenCpu) {
      s = make_sync_gpu_then_cpu(info.pGpu, std::static_pointer_cast<CpuOp>(bo), path);
    } else if (info.need == Need::GpuThenGpu) {
      s = make_sync_gpu_then_gpu(info.pGpu, std::static_pointer_cast<BoundGpuOp>(bo), path);
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
