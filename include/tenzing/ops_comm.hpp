// Host-posted point-to-point ops over the peer-memory transport.
//
// Reference: include/tenzing/mpi/ops_mpi.hpp:17-191, src/mpi/ops_mpi.cpp:11-49 (Isend, Irecv,
// Ialltoallv, Wait, OwningWaitall, MultiWait) and
// include/tenzing/halo_exchange/ops_halo_exchange.hpp:68-92 (OwningIsend, OwningIrecv).
// Same names and graph roles (all are CpuOps, so Pack(GPU) -> Isend needs a host sync and
// Wait -> Unpack(GPU) needs nothing), but no MPI underneath: a send is a copy-engine peer copy
// into the destination rank's buffer followed by a flag publish, both enqueued on a private
// transport stream; a receive is complete when the local flag word reaches the epoch.  These
// are the building blocks of the baseline ("reference structure") styles; the product path is
// the fused kernels.
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/operation.hpp"
#include "tenzing_b200.h"

namespace tenzing {

// the two private streams that play the MPI progress engine; share one per program
class P2PTransport {
public:
  static constexpr int kSendStreams = 4; // as a CUDA-aware MPI pipelines its peer copies
private:
  cudaStream_t send_[kSendStreams] = {nullptr, nullptr, nullptr, nullptr};
  cudaStream_t recv_ = nullptr;
  int next_ = 0;
  bool null_;

public:
  explicit P2PTransport(bool nullDevice = false);
  ~P2PTransport();
  P2PTransport(const P2PTransport &) = delete;
  P2PTransport &operator=(const P2PTransport &) = delete;
  // the stream for the next message (copy and flag publish of one message share a stream,
  // consecutive messages rotate over kSendStreams copy streams)
  cudaStream_t next_send_stream() {
    cudaStream_t s = send_[next_];
    next_ = (next_ + 1) % kSendStreams;
    return s;
  }
  cudaStream_t recv_stream() const { return recv_; }
  bool is_null() const { return null_; }
};

// the analogue of MPI_Request
struct Request {
  enum class Kind { None, Send, Recv } kind = Kind::None;
  std::shared_ptr<P2PTransport> transport;
  cudaEvent_t sent = nullptr;     // Send: recorded after copy + flag publish
  const uint64_t *flag = nullptr; // Recv: local flag word to wait for
  uint64_t epoch = 0;
  bool host = false;              // posted under Platform::host_execution(): no CUDA involved
  ~Request();
  void wait(); // host blocks until the request is complete
};
using RequestPtr = std::shared_ptr<Request>;

class Isend : public CpuOp {
public:
  struct Args {
    const void *buf = nullptr;   // local source (device memory)
    size_t bytes = 0;
    int dest = 0;                // informational (the pointers already name the peer)
    int tag = 0;
    void *peerBuf = nullptr;     // destination in the peer's memory
    uint64_t *peerFlag = nullptr; // flag word of the peer to publish
    RequestPtr request;
    bool operator==(const Args &o) const {
      return buf == o.buf && bytes == o.bytes && peerBuf == o.peerBuf && peerFlag == o.peerFlag &&
             request == o.request;
    }
  };
  Isend(const Args &args, std::shared_ptr<P2PTransport> tr, const std::string &name)
      : args_(args), tr_(std::move(tr)), name_(name) {}
  std::string name() const override { return name_; }
  CLONE_DEF(Isend)
  void run(Platform &plat) override;
  const Args &args() const { return args_; }

protected:
  Args args_;
  std::shared_ptr<P2PTransport> tr_;
  std::string name_;
};

class Irecv : public CpuOp {
public:
  struct Args {
    const uint64_t *flag = nullptr; // local flag the sender publishes
    int source = 0, tag = 0;
    RequestPtr request;
    bool operator==(const Args &o) const { return flag == o.flag && request == o.request; }
  };
  Irecv(const Args &args, std::shared_ptr<P2PTransport> tr, const std::string &name)
      : args_(args), tr_(std::move(tr)), name_(name) {}
  std::string name() const override { return name_; }
  CLONE_DEF(Irecv)
  void run(Platform &plat) override;
  const Args &args() const { return args_; }

protected:
  Args args_;
  std::shared_ptr<P2PTransport> tr_;
  std::string name_;
};

// variants that own their request (reference: OwningIsend / OwningIrecv)
class OwningIsend : public Isend {
public:
  OwningIsend(Args args, std::shared_ptr<P2PTransport> tr, const std::string &name)
      : Isend(with_request(args), std::move(tr), name) {}
  CLONE_DEF(OwningIsend)
  const RequestPtr &request() const { return args_.request; }

private:
  static Args with_request(Args a) {
    a.request = std::make_shared<Request>();
    return a;
  }
};
class OwningIrecv : public Irecv {
public:
  OwningIrecv(Args args, std::shared_ptr<P2PTransport> tr, const std::string &name)
      : Irecv(with_request(args), std::move(tr), name) {}
  CLONE_DEF(OwningIrecv)
  const RequestPtr &request() const { return args_.request; }

private:
  static Args with_request(Args a) {
    a.request = std::make_shared<Request>();
    return a;
  }
};

class Wait : public CpuOp {
  RequestPtr request_;
  std::string name_;

public:
  Wait(RequestPtr request, const std::string &name) : request_(std::move(request)), name_(name) {}
  std::string name() const override { return name_; }
  CLONE_DEF(Wait)
  void run(Platform &plat) override;
};

// waits for a set of requests it keeps alive (reference: OwningWaitall)
class OwningWaitall : public CpuOp {
  std::vector<RequestPtr> reqs_;
  std::string name_;

public:
  explicit OwningWaitall(const std::string &name) : name_(name) {}
  std::string name() const override { return name_; }
  CLONE_DEF(OwningWaitall)
  void add_request(const RequestPtr &r) { reqs_.push_back(r); }
  const std::vector<RequestPtr> &requests() const { return reqs_; }
  void run(Platform &plat) override;
};

// waits for requests one after the other (reference: MultiWait)
class MultiWait : public CpuOp {
  std::vector<RequestPtr> reqs_;
  std::string name_;

public:
  explicit MultiWait(const std::string &name) : name_(name) {}
  std::string name() const override { return name_; }
  CLONE_DEF(MultiWait)
  void add_request(const RequestPtr &r) { reqs_.push_back(r); }
  void run(Platform &plat) override;
};

// personalised all-to-all: slice p of the send buffer goes to peer p's buffer at the
// displacement that peer expects; one flag per peer.  (The reference defines MPI_Ialltoallv
// and never uses it, ops_mpi.hpp:82-119.)
class Ialltoallv : public CpuOp {
public:
  struct Args {
    const void *sendbuf = nullptr;
    std::vector<size_t> sendOffsets, sendBytes; // per peer, in bytes
    std::vector<void *> peerBufs;               // where slice p lands on peer p
    std::vector<uint64_t *> peerFlags;          // peer p's flag word for me
    RequestPtr request;                         // completes when all my slices are on their way
  };
  Ialltoallv(const Args &args, std::shared_ptr<P2PTransport> tr, const std::string &name)
      : args_(args), tr_(std::move(tr)), name_(name) {}
  std::string name() const override { return name_; }
  CLONE_DEF(Ialltoallv)
  void run(Platform &plat) override;

private:
  Args args_;
  std::shared_ptr<P2PTransport> tr_;
  std::string name_;
};

} // namespace tenzing
