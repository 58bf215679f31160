#include "tenzing/cuda/ops_cuda.hpp"

#include <mutex>
#include <typeinfo>
#include <unordered_map>

#include "tenzing/host_streams.hpp"

namespace tenzing {

// the one place that asks the type system; every other part of the library reads the bits
uint16_t OpBase::compute_traits() const {
  uint16_t t = op_trait::Known;
  const std::type_info &ti = typeid(*this);
  if (dynamic_cast<const BoundOp *>(this)) t |= op_trait::Bound;
  if (dynamic_cast<const CpuOp *>(this)) t |= op_trait::Cpu;
  if (ti == typeid(BoundGpuOp)) t |= op_trait::BoundGpu;
  if (dynamic_cast<const GpuOp *>(this)) t |= op_trait::Gpu;
  if (dynamic_cast<const CompoundOp *>(this)) t |= op_trait::Compound;
  if (dynamic_cast<const ChoiceOp *>(this)) t |= op_trait::Choice;
  if (dynamic_cast<const SyncOpBase *>(this)) t |= op_trait::Sync;
  if (dynamic_cast<const HasStream *>(this)) t |= op_trait::WithStream;
  if (dynamic_cast<const HasEvent *>(this)) t |= op_trait::WithEvent;
  if (ti == typeid(CudaEventRecord)) t |= op_trait::EventRecord;
  if (ti == typeid(CudaEventSync)) t |= op_trait::EventSync;
  if (ti == typeid(CudaStreamWaitEvent)) t |= op_trait::StreamWaitEvent;
  if (ti == typeid(Start)) t |= op_trait::IsStart;
  return t;
}

Json BoundGpuOp::json() const {
  Json j = op_->json();
  j["stream"] = stream_.id();
  return j;
}

void BoundGpuOp::run(Platform &plat) {
  if (HostStreams *hs = plat.host_streams()) {
    // issue only: the op runs on its stream's thread, under the epoch it was issued with
    GpuOpPtr op = op_;
    Platform *p = &plat;
    const uint64_t epoch = plat.epoch();
    hs->enqueue(stream_.id(), [op, p, epoch] {
      Platform::EpochScope scope(epoch);
      op->run(*p, nullptr);
    });
    return;
  }
  op_->run(plat, plat.cuda_stream(stream_));
}

void SyncOpBase::build_name() const {
  std::string n(pendingPrefix_);
  if (!pendingPreds_.empty()) {
    n += "-after";
    for (const auto &p : pendingPreds_) n += "-" + p->name();
  }
  if (!pendingSuccs_.empty()) {
    n += "-b4";
    for (const auto &s : pendingSuccs_) n += "-" + s->name();
  }
  name_ = n;
  pendingPrefix_ = nullptr;
  pendingPreds_.clear();
  pendingSuccs_.clear();
}

uint32_t SyncOpBase::intern_small(char kindTag, int a, int b, const OpBase &op) {
  // (kind, a, b) -> the uid intern_key(op.key()) gave the first op of that description
  struct Table {
    std::mutex mu;
    std::unordered_map<uint64_t, uint32_t> ids;
  };
  static Table *t = new Table(); // leaked on purpose, like the key table
  if (a < -1 || b < -1 || a >= (1 << 24) || b >= (1 << 24)) return intern_key(op.key());
  const uint64_t k = (uint64_t(uint8_t(kindTag)) << 56) | (uint64_t(uint32_t(a + 1)) << 28) | uint64_t(uint32_t(b + 1));
  {
    std::lock_guard<std::mutex> lk(t->mu);
    auto it = t->ids.find(k);
    if (it != t->ids.end()) return it->second;
  }
  const uint32_t id = intern_key(op.key());
  std::lock_guard<std::mutex> lk(t->mu);
  t->ids.emplace(k, id);
  return id;
}

Json CudaEventRecord::json() const {
  Json j = Json::object();
  j["name"] = name();
  j["kind"] = kind();
  j["event"] = event_.id();
  j["stream"] = stream_.id();
  return j;
}
void CudaEventRecord::run(Platform &plat) {
  if (HostStreams *hs = plat.host_streams()) return hs->record(event_.id(), stream_.id());
  if (plat.is_null()) return;
  CUDA_RUNTIME(cudaEventRecord(plat.cuda_event(event_), plat.cuda_stream(stream_)));
}

Json CudaStreamWaitEvent::json() const {
  Json j = Json::object();
  j["name"] = name();
  j["kind"] = kind();
  j["event"] = event_.id();
  j["stream"] = stream_.id();
  return j;
}
void CudaStreamWaitEvent::run(Platform &plat) {
  if (HostStreams *hs = plat.host_streams()) return hs->stream_wait(stream_.id(), event_.id());
  if (plat.is_null()) return;
  CUDA_RUNTIME(cudaStreamWaitEvent(plat.cuda_stream(stream_), plat.cuda_event(event_), 0));
}

Json CudaEventSync::json() const {
  Json j = Json::object();
  j["name"] = name();
  j["kind"] = kind();
  j["event"] = event_.id();
  return j;
}
void CudaEventSync::run(Platform &plat) {
  if (HostStreams *hs = plat.host_streams()) return hs->event_sync(event_.id());
  if (plat.is_null()) return;
  CUDA_RUNTIME(cudaEventSynchronize(plat.cuda_event(event_)));
}

Json StreamWait::json() const {
  Json j = Json::object();
  j["name"] = name();
  j["kind"] = kind();
  j["waitee"] = waitee_.id();
  j["waiter"] = waiter_.id();
  j["event"] = event_.id();
  return j;
}
void StreamWait::run(Platform &plat) {
  if (HostStreams *hs = plat.host_streams()) {
    hs->record(event_.id(), waitee_.id());
    return hs->stream_wait(waiter_.id(), event_.id());
  }
  if (plat.is_null()) return;
  cudaEvent_t e = plat.cuda_event(event_);
  CUDA_RUNTIME(cudaEventRecord(e, plat.cuda_stream(waitee_)));
  CUDA_RUNTIME(cudaStreamWaitEvent(plat.cuda_stream(waiter_), e, 0));
}

Json StreamSync::json() const {
  Json j = Json::object();
  j["name"] = name();
  j["kind"] = kind();
  j["stream"] = stream_.id();
  return j;
}
void StreamSync::run(Platform &plat) {
  if (HostStreams *hs = plat.host_streams()) return hs->stream_sync(stream_.id());
  if (plat.is_null()) return;
  CUDA_RUNTIME(cudaStreamSynchronize(plat.cuda_stream(stream_)));
}

BoundOpPtr sync_op_from_json(const Json &j) {
  if (!j.contains("kind") || !j.at("kind").is_string()) return nullptr;
  const std::string &kind = j.at("kind").as_string();
  const std::string name = j.contains("name") ? j.at("name").as_string() : std::string();
  // ids index tables (event pools, path indexes): a damaged file must not ask for a billion entries
  auto geti = [&](const char *k) {
    const int64_t v = j.at(k).as_int();
    if (v < 0 || v >= (int64_t(1) << 20))
      THROW_RUNTIME("serialized op " << j.dump() << ": " << k << " = " << v << " is out of range");
    return static_cast<int>(v);
  };
  if (kind == "CudaEventRecord") {
    return std::make_shared<CudaEventRecord>(Event(geti("event")), Stream(geti("stream")),
                                             name.empty() ? "CER-anon" : name);
  } else if (kind == "CudaStreamWaitEvent") {
    return std::make_shared<CudaStreamWaitEvent>(Stream(geti("stream")), Event(geti("event")),
                                                 name.empty() ? "CSWE-anon" : name);
  } else if (kind == "CudaEventSync") {
    return std::make_shared<CudaEventSync>(Event(geti("event")), name.empty() ? "CES-anon" : name);
  } else if (kind == "StreamWait") {
    auto p = std::make_shared<StreamWait>(Stream(geti("waitee")), Stream(geti("waiter")),
                                          Event(j.contains("event") ? geti("event") : 0));
    if (!name.empty()) p->set_name(name);
    return p;
  } else if (kind == "StreamSync") {
    auto p = std::make_shared<StreamSync>(Stream(geti("stream")));
    if (!name.empty()) p->set_name(name);
    return p;
  }
  return nullptr;
}

} // namespace tenzing
