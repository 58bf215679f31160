// Symmetric device heap: every rank allocates the same number of bytes and can
// address every other rank's allocation (the data plane that replaces
// CUDA-aware MPI).  Ranks in other processes are mapped through
// CUDA IPC handles exchanged over the Communicator; ranks in the same process
// (LoopbackComm threads, one GPU each) use peer access.  A host-memory mode
// backs CPU-only runs (set-up tests and host execution, tenzing/host_exec.hpp): ranks that
// are threads share pointers, ranks in other processes are mapped through POSIX shared memory.
#pragma once

#include <cstddef>
#include <cstdint>
#include <memory>
#include <vector>

#include "tenzing/communicator.hpp"

namespace tenzing {

class SymmetricHeap {
public:
  enum class Memory { Device, Host };

  // collective over comm
  SymmetricHeap(std::shared_ptr<Communicator> comm, size_t bytes, Memory memory = Memory::Device);
  ~SymmetricHeap();
  SymmetricHeap(const SymmetricHeap &) = delete;
  SymmetricHeap &operator=(const SymmetricHeap &) = delete;

  int rank() const { return rank_; }
  int size() const { return size_; }
  size_t bytes() const { return bytes_; }
  Memory memory() const { return memory_; }
  Communicator &comm() const { return *comm_; }
  std::shared_ptr<Communicator> comm_ptr() const { return comm_; }

  void *local() const { return bases_[size_t(rank_)]; }
  void *peer(int rank) const { return bases_.at(size_t(rank)); }
  template <typename T> T *at(int rank, size_t offsetBytes) const {
    return reinterpret_cast<T *>(static_cast<char *>(peer(rank)) + offsetBytes);
  }

  // Bump allocation.  Symmetric as long as every rank makes the same calls in
  // the same order; returns the byte offset (valid on every rank).
  size_t alloc(size_t bytes, size_t align = 256);
  size_t used() const { return used_; }

  // byte offset of a pointer into ANY rank's allocation (the same word on every rank has the
  // same offset); -1 if the pointer is not in the heap
  int64_t offset_of(const void *p) const {
    for (void *b : bases_) {
      const char *lo = static_cast<const char *>(b);
      const char *q = static_cast<const char *>(p);
      if (b && q >= lo && q < lo + bytes_) return int64_t(q - lo);
    }
    return -1;
  }

  // zero the local allocation (or a range of it)
  void zero(size_t offset = 0, size_t bytes = size_t(-1));

private:
  std::shared_ptr<Communicator> comm_;
  int rank_ = 0, size_ = 1; // cached: the destructor must not call into the communicator
  size_t bytes_;
  Memory memory_;
  size_t used_ = 0;
  std::vector<void *> bases_;
  std::vector<bool> ipcOpened_;
  std::vector<bool> shmMapped_; // host mode: peers in other processes mapped through /dev/shm
  bool localShm_ = false;       // host mode: my allocation is a shared-memory segment
  char shmName_[48] = {0};
};

} // namespace tenzing
