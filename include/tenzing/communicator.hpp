// Host control plane + peer-memory directory.
//
// The reference uses MPI for three things: control collectives
// (Bcast / Allreduce MAX / Barrier), set-up traffic, and the data plane
// (CUDA-aware Isend/Irecv).  Here the data plane is device-initiated stores
// into peer GPU memory, so the host side only needs a small collective
// interface.  Implementations:
//   SelfComm      world size 1
//   LoopbackComm  N logical ranks = N threads of one process (CPU tests,
//                 single-process multi-GPU drivers)
//   SocketComm    one process per GPU, TCP on 127.0.0.1 (native launcher,
//                 reads RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT like torchrun)
//   PyComm        (bindings) forwards to torch.distributed (gloo / NCCL)
#pragma once

#include <condition_variable>
#include <functional>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

namespace tenzing {

class Communicator {
public:
  virtual ~Communicator() = default;
  virtual int rank() const = 0;
  virtual int size() const = 0;
  virtual void barrier() = 0;
  virtual void bcast(void *buf, size_t bytes, int root) = 0;
  virtual void allreduce_max(double *vals, size_t n) = 0;
  // out has size()*bytesPerRank bytes; rank r's contribution at r*bytesPerRank
  virtual void allgather(const void *in, void *out, size_t bytesPerRank) = 0;
  virtual std::string kind() const = 0;

  // conveniences built on the primitives
  std::string bcast_string(const std::string &s, int root);
  int64_t bcast_int(int64_t v, int root);
  double max_over_ranks(double v) {
    allreduce_max(&v, 1);
    return v;
  }
  std::vector<std::string> allgather_string(const std::string &s);
};

class SelfComm : public Communicator {
public:
  int rank() const override { return 0; }
  int size() const override { return 1; }
  void barrier() override {}
  void bcast(void *, size_t, int) override {}
  void allreduce_max(double *, size_t) override {}
  void allgather(const void *in, void *out, size_t n) override { std::memcpy(out, in, n); }
  std::string kind() const override { return "self"; }
};

// shared state of an in-process world; create one, then one LoopbackComm per thread
// thrown out of a collective when another rank of the (thread) world has failed
struct LoopbackAborted : std::runtime_error {
  LoopbackAborted() : std::runtime_error("LoopbackWorld: another rank failed, collective abandoned") {}
};

class LoopbackWorld {
  friend class LoopbackComm;
  const int size_;
  std::mutex mu_;
  std::condition_variable cv_;
  int arrived_ = 0;
  uint64_t generation_ = 0;
  bool aborted_ = false; // a rank left with an exception: collectives throw instead of waiting
  std::vector<const void *> src_;
  std::vector<void *> dst_;

public:
  explicit LoopbackWorld(int size) : size_(size), src_(size), dst_(size) {}
  int size() const { return size_; }
  // run body(rank, comm) on `size` threads and join; rethrows the first exception.  When a
  // rank's body throws, the ranks waiting in a collective are released with an exception too
  // (a failed rank must not hang the others).
  void abort();
  static void run(int size, const std::function<void(int, std::shared_ptr<Communicator>)> &body);
};

class LoopbackComm : public Communicator {
  std::shared_ptr<LoopbackWorld> world_;
  int rank_;
  // all ranks arrive; the last arriver runs `atRendezvous` while the others wait
  template <typename F> void rendezvous(const void *src, void *dst, F &&atRendezvous);

public:
  LoopbackComm(std::shared_ptr<LoopbackWorld> world, int rank)
      : world_(std::move(world)), rank_(rank) {}
  int rank() const override { return rank_; }
  int size() const override { return world_->size_; }
  void barrier() override;
  void bcast(void *buf, size_t bytes, int root) override;
  void allreduce_max(double *vals, size_t n) override;
  void allgather(const void *in, void *out, size_t bytesPerRank) override;
  std::string kind() const override { return "loopback"; }
};

// One process per rank; rank 0 is a star hub over TCP.  Messages are small
// (schedules as JSON, timings, IPC handles), so a star is plenty.
class SocketComm : public Communicator {
  int rank_, size_;
  int listenFd_ = -1;
  std::vector<int> fds_; // rank 0: fd per peer (index = rank); others: fds_[0] = hub

  void send_all(int fd, const void *buf, size_t n);
  void recv_all(int fd, void *buf, size_t n);

public:
  SocketComm(int rank, int size, const std::string &addr, int port, double timeoutSec = 120.0);
  ~SocketComm() override;
  // RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT (+portOffset so it does not
  // collide with a torch store on MASTER_PORT); nullptr if RANK is unset
  static std::shared_ptr<Communicator> from_env(int portOffset = 17);
  int rank() const override { return rank_; }
  int size() const override { return size_; }
  void barrier() override;
  void bcast(void *buf, size_t bytes, int root) override;
  void allreduce_max(double *vals, size_t n) override;
  void allgather(const void *in, void *out, size_t bytesPerRank) override;
  std::string kind() const override { return "socket"; }
};

} // namespace tenzing
