#include "tenzing/platform.hpp"

#include "tenzing/host_streams.hpp"

#include "tenzing/communicator.hpp"
#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/cuda/ops_cuda.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {

bool Equivalence::merge(const Equivalence &o) {
  if (!truthy_ || !o.truthy_) return false;
  for (const auto &kv : o.streams_.forward())
    if (!streams_.check_or_insert(kv.first, kv.second)) return false;
  for (const auto &kv : o.events_.forward())
    if (!events_.check_or_insert(kv.first, kv.second)) return false;
  return true;
}

std::string Equivalence::str() const {
  if (!truthy_) return "{not equivalent}";
  return "{streams: " + streams_.str() + ", events: " + events_.str() + "}";
}

cudaEvent_t ResourceMap::cuda_event(const Event &e) const {
  auto it = events_.find(e);
  if (it == events_.end()) THROW_RUNTIME("no cudaEvent_t provisioned for event " << e.str());
  return it->second;
}

CudaEventPool::~CudaEventPool() {
  for (cudaEvent_t e : events_)
    if (e) cudaEventDestroy(e);
}

cudaEvent_t CudaEventPool::new_event() {
  if (null_) {
    ++next_;
    return nullptr;
  }
  if (next_ == events_.size()) {
    cudaEvent_t e;
    CUDA_RUNTIME(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    events_.push_back(e);
  }
  return events_[next_++];
}

Platform::Platform(std::shared_ptr<Communicator> comm, Device device)
    : device_(device), pool_(new CudaEventPool(device == Device::Null)), comm_(std::move(comm)) {
  if (!comm_) comm_ = std::make_shared<SelfComm>();
}

Platform::Platform(Platform &&o) noexcept
    : device_(o.device_), cStreams_(std::move(o.cStreams_)), owned_(std::move(o.owned_)),
      resourceMap_(std::move(o.resourceMap_)), pool_(std::move(o.pool_)),
      comm_(std::move(o.comm_)), epoch_(o.epoch_), hostExec_(o.hostExec_),
      hostStreams_(std::move(o.hostStreams_)), deviceBarrier_(std::move(o.deviceBarrier_)),
      streams_(std::move(o.streams_)) {
  o.cStreams_.clear();
  o.owned_.clear();
}

Platform::~Platform() {
  for (size_t i = 0; i < cStreams_.size(); ++i)
    if (owned_[i] && cStreams_[i]) cudaStreamDestroy(cStreams_[i]);
}

Platform Platform::make_n_streams(int n, std::shared_ptr<Communicator> comm, Device device) {
  Platform p(std::move(comm), device);
  p.ensure_streams(n);
  return p;
}

Stream Platform::new_stream() {
  cudaStream_t s = nullptr;
  if (!is_null()) CUDA_RUNTIME(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  Stream id(static_cast<int>(cStreams_.size()));
  cStreams_.push_back(s);
  owned_.push_back(!is_null());
  streams_.push_back(id);
  return id;
}

Stream Platform::adopt_stream(cudaStream_t s) {
  Stream id(static_cast<int>(cStreams_.size()));
  cStreams_.push_back(s);
  owned_.push_back(false);
  streams_.push_back(id);
  return id;
}

void Platform::ensure_streams(int n) {
  while (num_streams() < n) new_stream();
}

cudaStream_t Platform::cuda_stream(const Stream &s) const {
  if (s.id() < 0 || size_t(s.id()) >= cStreams_.size())
    THROW_RUNTIME("platform has no stream " << s.str());
  return cStreams_[s.id()];
}

void Platform::provision_events(const Sequence<BoundOp> &seq) {
  pool_->reset();
  resourceMap_.clear();
  for (const auto &op : seq) {
    if (auto he = std::dynamic_pointer_cast<HasEvent>(op)) {
      for (const Event &e : he->get_events())
        if (!resourceMap_.contains(e)) resourceMap_.insert(e, pool_->new_event());
    }
  }
}

Communicator &Platform::comm() const { return *comm_; }

void Platform::enable_host_streams() {
  if (!is_null()) THROW_RUNTIME("host streams belong to a Null platform with host execution");
  hostExec_ = true;
  if (!hostStreams_ || hostStreams_->num_streams() < num_streams())
    hostStreams_ = std::make_shared<HostStreams>(num_streams());
}

thread_local uint64_t Platform::tlsEpoch_ = 0;
thread_local bool Platform::tlsEpochSet_ = false;

void Platform::synchronize() const {
  if (HostStreams *hs = host_streams()) hs->sync_all();
  if (is_null()) return;
  for (cudaStream_t s : cStreams_) CUDA_RUNTIME(cudaStreamSynchronize(s));
}

} // namespace tenzing
