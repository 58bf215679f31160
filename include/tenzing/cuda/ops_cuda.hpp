// GPU operations and the CUDA synchronisation ops the EventSynchronizer inserts.
//
// Reference: include/tenzing/cuda/ops_cuda.hpp:24-243, src/cuda/ops_cuda.cpp:12-203.
// JSON forms and the CER-/CSWE-/CES- naming scheme are kept byte-compatible
//  because result CSVs are keyed on them.
//
// Sync ops are identified by the resources they act on - (event, stream) -
// not by their (context-dependent) name, so two searches paths that insert
// "the same" record compare equal.  An event id is never reused inside one
// sequence.
#pragma once

#include <vector>

#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/operation.hpp"
#include "tenzing/platform.hpp"

namespace tenzing {

class HasEvent {
public:
  virtual ~HasEvent() = default;
  virtual std::vector<Event> get_events() const = 0;
};
class HasStream {
public:
  virtual ~HasStream() = default;
  virtual std::vector<Stream> get_streams() const = 0;
};

// a kernel-like op that can run in any stream; not executable until bound
class GpuOp : public OpBase {
public:
  std::string kind() const override { return "GpuOp"; }
  virtual void run(cudaStream_t stream) = 0;
  // ops that need more than the stream (communicator, epoch) override this
  virtual void run(Platform &plat, cudaStream_t stream) {
    (void)plat;
    run(stream);
  }
  // True for kernels whose real data dependency is a device-side flag published by ANOTHER
  // rank (wait+unpack style ops).  Their graph predecessors on the GPU then only have to be
  // ISSUED before them (sequence order), not completed, so the EventSynchronizer inserts no
  // CudaEventRecord/CudaStreamWaitEvent for those edges and the kernels may overlap.
  virtual bool issue_order_only() const { return false; }
};
using GpuOpPtr = std::shared_ptr<GpuOp>;

// a GpuOp fixed to one stream
class BoundGpuOp final : public BoundOp, public HasStream {
  GpuOpPtr op_;
  Stream stream_;

public:
  BoundGpuOp(const GpuOpPtr &op, Stream stream) : op_(op), stream_(stream) {}
  std::string name() const override { return op_->name(); }
  std::string desc() const override { return op_->desc() + "(s" + stream_.str() + ")"; }
  std::string kind() const override { return "BoundGpuOp"; }
  std::string key() const override { return op_->key() + "@" + stream_.str(); }
  std::string unbound_key() const override { return op_->key(); }
  Json json() const override;
  CLONE_DEF(BoundGpuOp)
  void run(Platform &plat) override;
  const Stream &stream() const { return stream_; }
  std::vector<Stream> get_streams() const override { return {stream_}; }
  GpuOpPtr unbound() const { return op_; }
};

using OpSet = std::vector<std::shared_ptr<const OpBase>>;

class SyncOpBase : public BoundOp {
protected:
  // The search makes a synchronisation op for every candidate placement and executes a few of
  // them; the descriptive name (prefix + the neighbours' names) is only put together when
  // somebody asks for it.  Identity does not depend on it (key() is kind + event/stream).
  mutable std::string name_;
  mutable const char *pendingPrefix_ = nullptr;
  mutable OpSet pendingPreds_, pendingSuccs_;
  void make_name(const char *prefix, const OpSet &preds, const OpSet &succs) {
    pendingPrefix_ = prefix;
    pendingPreds_ = preds;
    pendingSuccs_ = succs;
  }
  void build_name() const;
  // identity of an op that is fully described by a kind and two small integers, without
  // building and hashing its key string on every new object (src/core/operation.cpp)
  static uint32_t intern_small(char kindTag, int a, int b, const OpBase &op);

public:
  std::string name() const override {
    if (pendingPrefix_) build_name();
    return name_;
  }
  void set_name(const std::string &n) {
    name_ = n;
    pendingPrefix_ = nullptr;
    pendingPreds_.clear();
    pendingSuccs_.clear();
  }
};

class CudaEventRecord final : public SyncOpBase, public HasEvent, public HasStream {
  Event event_;
  Stream stream_;

public:
  CudaEventRecord(Event event, Stream stream, const std::string &name = "CER-anon")
      : event_(event), stream_(stream) {
    name_ = name;
  }
  std::string kind() const override { return "CudaEventRecord"; }
  std::string key() const override { return "CER:e" + event_.str() + ":s" + stream_.str(); }
  uint32_t intern_self() const override { return intern_small('R', event_.id(), stream_.id(), *this); }
  uint32_t intern_unbound_self() const override { return uid(); }
  std::string desc() const override {
    return name() + "(e" + event_.str() + ",s" + stream_.str() + ")";
  }
  Json json() const override;
  CLONE_DEF(CudaEventRecord)
  void update_name(const OpSet &preds, const OpSet &succs) { make_name("CER", preds, succs); }
  void run(Platform &plat) override;
  const Event &event() const { return event_; }
  const Stream &stream() const { return stream_; }
  std::vector<Event> get_events() const override { return {event_}; }
  std::vector<Stream> get_streams() const override { return {stream_}; }
};

class CudaStreamWaitEvent final : public SyncOpBase, public HasEvent, public HasStream {
  Stream stream_;
  Event event_;

public:
  CudaStreamWaitEvent(Stream stream, Event event, const std::string &name = "CSWE-anon")
      : stream_(stream), event_(event) {
    name_ = name;
  }
  std::string kind() const override { return "CudaStreamWaitEvent"; }
  std::string key() const override { return "CSWE:s" + stream_.str() + ":e" + event_.str(); }
  uint32_t intern_self() const override { return intern_small('W', stream_.id(), event_.id(), *this); }
  uint32_t intern_unbound_self() const override { return uid(); }
  std::string desc() const override {
    return name() + "(s" + stream_.str() + ",e" + event_.str() + ")";
  }
  Json json() const override;
  CLONE_DEF(CudaStreamWaitEvent)
  void update_name(const OpSet &preds, const OpSet &succs) { make_name("CSWE", preds, succs); }
  void run(Platform &plat) override;
  const Event &event() const { return event_; }
  const Stream &stream() const { return stream_; }
  std::vector<Event> get_events() const override { return {event_}; }
  std::vector<Stream> get_streams() const override { return {stream_}; }
};

class CudaEventSync final : public SyncOpBase, public HasEvent {
  Event event_;

public:
  explicit CudaEventSync(Event event, const std::string &name = "CES-anon") : event_(event) {
    name_ = name;
  }
  std::string kind() const override { return "CudaEventSync"; }
  std::string key() const override { return "CES:e" + event_.str(); }
  uint32_t intern_self() const override { return intern_small('S', event_.id(), 0, *this); }
  uint32_t intern_unbound_self() const override { return uid(); }
  std::string desc() const override { return name() + "(e" + event_.str() + ")"; }
  Json json() const override;
  CLONE_DEF(CudaEventSync)
  void update_name(const OpSet &preds, const OpSet &succs) { make_name("CES", preds, succs); }
  void run(Platform &plat) override;
  const Event &event() const { return event_; }
  std::vector<Event> get_events() const override { return {event_}; }
};

// legacy: record on `waitee`, make `waiter` wait, in one op
class StreamWait final : public SyncOpBase, public HasEvent, public HasStream {
  Event event_;
  Stream waitee_, waiter_;

public:
  StreamWait(Stream waitee, Stream waiter, Event event)
      : event_(event), waitee_(waitee), waiter_(waiter) {
    name_ = "StreamWait-anon";
  }
  std::string kind() const override { return "StreamWait"; }
  std::string key() const override {
    return "StreamWait:" + waitee_.str() + ">" + waiter_.str() + ":e" + event_.str();
  }
  Json json() const override;
  CLONE_DEF(StreamWait)
  void update_name(const OpSet &preds, const OpSet &succs) {
    make_name("StreamWait", preds, succs);
  }
  void run(Platform &plat) override;
  const Stream &waiter() const { return waiter_; }
  const Stream &waitee() const { return waitee_; }
  const Event &event() const { return event_; }
  std::vector<Event> get_events() const override { return {event_}; }
  std::vector<Stream> get_streams() const override { return {waitee_, waiter_}; }
};

// legacy: host blocks until the stream is idle
class StreamSync final : public SyncOpBase, public HasStream {
  Stream stream_;

public:
  explicit StreamSync(Stream stream) : stream_(stream) { name_ = "StreamSync-anon"; }
  std::string kind() const override { return "StreamSync"; }
  std::string key() const override { return "StreamSync:s" + stream_.str(); }
  Json json() const override;
  CLONE_DEF(StreamSync)
  void update_name(const OpSet &preds, const OpSet &succs) {
    make_name("StreamSync", preds, succs);
  }
  void run(Platform &plat) override;
  const Stream &stream() const { return stream_; }
  std::vector<Stream> get_streams() const override { return {stream_}; }
};

// construct a sync op from its JSON form ("kind" = CudaEventRecord |
// CudaStreamWaitEvent | CudaEventSync | StreamWait | StreamSync); null if other
BoundOpPtr sync_op_from_json(const Json &j);

// a GpuOp from a callable (tests, Python-defined kernels)
class LambdaGpuOp : public GpuOp {
  std::string name_;
  std::function<void(cudaStream_t)> fn_;

public:
  LambdaGpuOp(const std::string &name, std::function<void(cudaStream_t)> fn = nullptr)
      : name_(name), fn_(std::move(fn)) {}
  std::string name() const override { return name_; }
  CLONE_DEF(LambdaGpuOp)
  void run(cudaStream_t s) override {
    if (fn_) fn_(s);
  }
};

} // namespace tenzing
