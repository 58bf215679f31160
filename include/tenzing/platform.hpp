// Runtime substrate: abstract Stream/Event handles, the Platform that maps
// them to real CUDA resources, event pool/resource map, and the Equivalence
// object returned by the "same up to renaming of streams and events" checks.
//
// Reference: include/tenzing/platform.hpp:22-270, include/tenzing/bijection.hpp:3-45,
// src/platform.cpp:8-47.  Differences (by design):
//  * a Platform can be created in Null mode (ids only, no CUDA calls) so the
//    whole SDP core and both solvers run on a machine without a GPU;
//  * streams are non-blocking CUDA streams and can also be adopted from the
//    caller (e.g. torch streams);
//  * the control-plane handle is a tenzing::Communicator, not an MPI_Comm;
//  * Bijection is a true bijection (forward and reverse maps), not the
//    single-map involution of the reference.
#pragma once

#include <cstdint>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include <cuda_runtime_api.h>

#include "tenzing/json.hpp"
#include "tenzing/macro_at.hpp"

namespace tenzing {

class Communicator;
class DeviceBarrier;
class HostStreams;

#define TENZING_ID_HANDLE(NAME)                                                                    \
  class NAME {                                                                                     \
  public:                                                                                          \
    using id_t = int;                                                                              \
    NAME() : id_(0) {}                                                                             \
    explicit NAME(id_t id) : id_(id) {}                                                            \
    id_t id() const { return id_; }                                                                \
    explicit operator id_t() const { return id_; }                                                 \
    bool operator==(const NAME &r) const { return id_ == r.id_; }                                  \
    bool operator!=(const NAME &r) const { return id_ != r.id_; }                                  \
    bool operator<(const NAME &r) const { return id_ < r.id_; }                                    \
    bool operator>(const NAME &r) const { return id_ > r.id_; }                                    \
    std::string str() const { return std::to_string(id_); }                                        \
                                                                                                   \
  private:                                                                                         \
    id_t id_;                                                                                      \
  };                                                                                               \
  inline void to_json(Json &j, const NAME &s) { j = Json(s.id()); }                                \
  inline void from_json(const Json &j, NAME &s) { s = NAME(static_cast<int>(j.as_int())); }

TENZING_ID_HANDLE(Stream)
TENZING_ID_HANDLE(Event)
#undef TENZING_ID_HANDLE

// placeholder for "which CPU executes an op" (reference keeps an unused skeleton,
// platform.hpp:91-128); carried so serialized schedules can name it later
struct CPU {
  int id = 0;
};

// Placeholders for "ops own values" (reference keeps the same unused skeletons,
// platform.hpp:91-128): a value an op produces or consumes, as a scalar or an array.
struct IValue {
  virtual ~IValue() = default;
  virtual size_t bytes() const = 0;
};
template <typename T> struct ScalarValue : IValue {
  T value{};
  size_t bytes() const override { return sizeof(T); }
};
template <typename T> struct ArrayValue : IValue {
  T *data = nullptr;
  size_t count = 0;
  size_t bytes() const override { return count * sizeof(T); }
};

template <typename T> class Bijection {
  std::map<T, T> fwd_, rev_;

public:
  // true if a<->b is consistent with what is recorded so far (and records it)
  bool check_or_insert(const T &a, const T &b) {
    auto f = fwd_.find(a);
    auto r = rev_.find(b);
    if (f == fwd_.end() && r == rev_.end()) {
      fwd_.emplace(a, b);
      rev_.emplace(b, a);
      return true;
    }
    return f != fwd_.end() && r != rev_.end() && f->second == b && r->second == a;
  }
  bool empty() const { return fwd_.empty(); }
  const std::map<T, T> &forward() const { return fwd_; }
  bool operator==(const Bijection &o) const { return fwd_ == o.fwd_; }
  std::string str() const {
    std::string s;
    for (const auto &kv : fwd_) {
      if (!s.empty()) s += ",";
      s += kv.first.str() + "->" + kv.second.str();
    }
    return "{" + s + "}";
  }
};

// result of an equivalence query; truthy iff equivalent, carries the renaming
class Equivalence {
  Bijection<Stream> streams_;
  Bijection<Event> events_;
  bool truthy_ = true;

public:
  static Equivalence falsy() {
    Equivalence e;
    e.truthy_ = false;
    return e;
  }
  bool check_or_insert(const Stream &a, const Stream &b) { return streams_.check_or_insert(a, b); }
  bool check_or_insert(const Event &a, const Event &b) { return events_.check_or_insert(a, b); }
  explicit operator bool() const { return truthy_; }
  bool operator==(const Equivalence &o) const {
    return truthy_ == o.truthy_ && streams_ == o.streams_ && events_ == o.events_;
  }
  // merge another mapping into this one; false if they conflict
  bool merge(const Equivalence &o);
  const Bijection<Stream> &streams() const { return streams_; }
  const Bijection<Event> &events() const { return events_; }
  std::string str() const;
};

// abstract Event -> real cudaEvent_t for the sequence about to run
class ResourceMap {
  std::map<Event, cudaEvent_t> events_;

public:
  bool contains(const Event &e) const { return events_.count(e) != 0; }
  bool insert(const Event &e, cudaEvent_t ce) { return events_.emplace(e, ce).second; }
  cudaEvent_t cuda_event(const Event &e) const;
  size_t size() const { return events_.size(); }
  void clear() { events_.clear(); }
};

// recycles timing-disabled cudaEvent_t across sequences
class CudaEventPool {
  std::vector<cudaEvent_t> events_;
  size_t next_ = 0;
  bool null_ = false;

public:
  explicit CudaEventPool(bool nullDevice = false) : null_(nullDevice) {}
  CudaEventPool(const CudaEventPool &) = delete;
  CudaEventPool &operator=(const CudaEventPool &) = delete;
  ~CudaEventPool();
  cudaEvent_t new_event();
  void reset() { next_ = 0; }
  size_t capacity() const { return events_.size(); }
};

template <typename T> class Sequence;
class BoundOp;

class Platform {
public:
  enum class Device { Null, Cuda };

private:
  Device device_;
  std::vector<cudaStream_t> cStreams_; // index = Stream id
  std::vector<bool> owned_;
  ResourceMap resourceMap_;
  std::unique_ptr<CudaEventPool> pool_;
  std::shared_ptr<Communicator> comm_;
  uint64_t epoch_ = 0;
  bool hostExec_ = false;
  std::shared_ptr<HostStreams> hostStreams_;
  std::shared_ptr<DeviceBarrier> deviceBarrier_;

  static thread_local uint64_t tlsEpoch_;
  static thread_local bool tlsEpochSet_;

public:
  std::vector<Stream> streams_; // the streams decisions may bind GPU ops to

  explicit Platform(std::shared_ptr<Communicator> comm = nullptr, Device device = Device::Cuda);
  Platform(const Platform &) = delete;
  Platform &operator=(const Platform &) = delete;
  Platform(Platform &&) noexcept;
  ~Platform();

  static Platform make_n_streams(int n, std::shared_ptr<Communicator> comm = nullptr,
                                 Device device = Device::Cuda);
  // Null platform: abstract ids only (CPU-only search, CSV replay, tests)
  static Platform make_null(int nStreams, std::shared_ptr<Communicator> comm = nullptr) {
    return make_n_streams(nStreams, std::move(comm), Device::Null);
  }

  Device device() const { return device_; }
  bool is_null() const { return device_ == Device::Null; }

  Stream new_stream();                   // create and own a non-blocking stream
  Stream adopt_stream(cudaStream_t s);   // use a caller-owned stream
  void ensure_streams(int n);            // at least n streams exist
  int num_streams() const { return static_cast<int>(streams_.size()); }

  cudaStream_t cuda_stream(const Stream &s) const;
  cudaEvent_t cuda_event(const Event &e) const { return resourceMap_.cuda_event(e); }

  ResourceMap &resource_map() { return resourceMap_; }
  const ResourceMap &resource_map() const { return resourceMap_; }
  CudaEventPool &event_pool() { return *pool_; }

  // give every Event named by the ops of `seq` a real cudaEvent_t (reference:
  // dfs.hpp:145-167 / mcts.hpp:246-270 do this inline before each benchmark)
  void provision_events(const Sequence<BoundOp> &seq);

  Communicator &comm() const;
  std::shared_ptr<Communicator> comm_ptr() const { return comm_; }
  bool has_comm() const { return bool(comm_); }

  // Iteration epoch for device-side flag protocols: every execution of a
  // sequence happens under a new epoch (all ranks advance it in lock step);
  // fused put/wait kernels use it as the flag value and (epoch & 1) as the
  // staging-buffer parity.
  uint64_t epoch() const { return tlsEpochSet_ ? tlsEpoch_ : epoch_; }
  // A task that runs later than it was issued (host streams) sees the epoch of its issue, the
  // way a kernel keeps the epoch it was launched with.
  struct EpochScope {
    explicit EpochScope(uint64_t e) {
      tlsEpoch_ = e;
      tlsEpochSet_ = true;
    }
    ~EpochScope() { tlsEpochSet_ = false; }
  };
  uint64_t next_epoch() { return ++epoch_; }
  void set_epoch(uint64_t e) { epoch_ = e; }

  // Null platform only: execute the ops' host restatement (tenzing/host_exec.hpp) on the
  // calling thread instead of skipping them, for workloads whose memory is a host-mode
  // SymmetricHeap.  Sequences then really run - payload, flags, epochs - without a GPU.
  void set_host_execution(bool on) { hostExec_ = on; }
  bool host_execution() const { return hostExec_ && is_null(); }
  // Host execution with CUDA's concurrency: one worker thread per stream, events as in CUDA
  // (tenzing/host_streams.hpp).  Without it host-executed GPU ops run inline, in sequence order.
  void enable_host_streams();
  HostStreams *host_streams() const { return host_execution() ? hostStreams_.get() : nullptr; }

  // optional GPU-side all-rank barrier the device-timed benchmarkers use to line up their start
  void set_device_barrier(std::shared_ptr<DeviceBarrier> b) { deviceBarrier_ = std::move(b); }
  const std::shared_ptr<DeviceBarrier> &device_barrier() const { return deviceBarrier_; }

  // block the host until every stream of the platform is idle
  void synchronize() const;
};

} // namespace tenzing
