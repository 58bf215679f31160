#include "tenzing/cuda/symmetric_heap.hpp"

#include <cstdio>
#include <cstdlib>
#include <cstring>

#include <atomic>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/macro_at.hpp"

namespace tenzing {

namespace {
struct Hello {
  int64_t pid;
  int64_t hostId;
  int32_t device;
  int32_t pad;
  uint64_t ptr;
  cudaIpcMemHandle_t handle;
};
} // namespace

SymmetricHeap::SymmetricHeap(std::shared_ptr<Communicator> comm, size_t bytes, Memory memory)
    : comm_(std::move(comm)), bytes_(bytes), memory_(memory) {
  if (!comm_) comm_ = std::make_shared<SelfComm>();
  const int n = comm_->size(), me = comm_->rank();
  rank_ = me;
  size_ = n;
  bases_.assign(size_t(n), nullptr);
  ipcOpened_.assign(size_t(n), false);
  shmMapped_.assign(size_t(n), false);

  Hello mine;
  std::memset(&mine, 0, sizeof(mine));
  mine.pid = int64_t(getpid());
  static const int64_t hostId = int64_t(gethostid()); // once: ranks may be threads
  mine.hostId = hostId;
  void *base = nullptr;
  if (memory_ == Memory::Host) {
    // Ranks in other processes (torchrun + gloo, tzrun + sockets) reach this rank's host heap
    // through POSIX shared memory, so host execution works with one process per rank as well;
    // ranks that are threads of this process simply use the pointer.
    int64_t pid = mine.pid;
    std::vector<int64_t> pids(static_cast<size_t>(n));
    comm_->allgather(&pid, pids.data(), sizeof(pid));
    bool otherProcess = false;
    for (int64_t p : pids) otherProcess = otherProcess || p != pid;
    if (otherProcess) {
      static std::atomic<uint64_t> counter{0};
      std::snprintf(shmName_, sizeof(shmName_), "/tzb_%lld_%llu", static_cast<long long>(pid),
                    static_cast<unsigned long long>(counter++));
      const size_t len = bytes_ ? bytes_ : 1;
      int fd = shm_open(shmName_, O_CREAT | O_EXCL | O_RDWR, 0600);
      if (fd >= 0 && ftruncate(fd, off_t(len)) == 0) {
        void *p = mmap(nullptr, len, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
        if (p != MAP_FAILED) {
          base = p; // fresh shared memory reads as zero
          localShm_ = true;
        }
      }
      if (fd >= 0) close(fd);
      if (!localShm_) {
        if (fd >= 0) shm_unlink(shmName_);
        shmName_[0] = 0;
      }
    }
    if (!base) base = std::calloc(bytes_ ? bytes_ : 1, 1);
    if (!base) THROW_RUNTIME("SymmetricHeap: host allocation of " << bytes_ << " bytes failed");
    mine.device = -1;
    static_assert(sizeof(shmName_) <= sizeof(mine.handle), "the name travels in the handle field");
    std::memcpy(&mine.handle, shmName_, sizeof(shmName_));
  } else {
    int dev = 0;
    CUDA_RUNTIME(cudaGetDevice(&dev));
    mine.device = dev;
    CUDA_RUNTIME(cudaMalloc(&base, bytes_ ? bytes_ : 256));
    CUDA_RUNTIME(cudaMemset(base, 0, bytes_ ? bytes_ : 256));
    CUDA_RUNTIME(cudaDeviceSynchronize());
    if (n > 1) CUDA_RUNTIME(cudaIpcGetMemHandle(&mine.handle, base));
  }
  mine.ptr = reinterpret_cast<uint64_t>(base);

  std::vector<Hello> all;
  all.resize(size_t(n));
  comm_->allgather(&mine, all.data(), sizeof(Hello));
  for (int r = 0; r < n; ++r) {
    const Hello &h = all[size_t(r)];
    if (r == me) {
      bases_[size_t(r)] = base;
    } else if (h.pid == mine.pid && h.hostId == mine.hostId) {
      // same address space: the pointer is usable as is (after enabling peer access)
      bases_[size_t(r)] = reinterpret_cast<void *>(h.ptr);
      if (memory_ == Memory::Device && h.device != mine.device) {
        int can = 0;
        CUDA_RUNTIME(cudaDeviceCanAccessPeer(&can, mine.device, h.device));
        if (!can) THROW_RUNTIME("device " << mine.device << " cannot access device " << h.device);
        cudaError_t e = cudaDeviceEnablePeerAccess(h.device, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) {
          (void)cudaGetLastError();
        } else {
          CUDA_RUNTIME(e);
        }
      }
    } else {
      if (memory_ == Memory::Host) {
        // another process: map its shared-memory segment (stays null if it has none, e.g.
        // on another host - then only dry runs work, ops never dereference a peer pointer)
        char name[sizeof(shmName_)];
        std::memcpy(name, &h.handle, sizeof(name));
        name[sizeof(name) - 1] = 0;
        if (name[0] && h.hostId == mine.hostId) {
          int fd = shm_open(name, O_RDWR, 0600);
          if (fd >= 0) {
            void *p = mmap(nullptr, bytes_ ? bytes_ : 1, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
            close(fd);
            if (p != MAP_FAILED) {
              bases_[size_t(r)] = p;
              shmMapped_[size_t(r)] = true;
            }
          }
        }
        continue;
      }
      if (h.hostId != mine.hostId) THROW_RUNTIME("SymmetricHeap: rank " << r << " is on another host");
      void *p = nullptr;
      CUDA_RUNTIME(cudaIpcOpenMemHandle(&p, h.handle, cudaIpcMemLazyEnablePeerAccess));
      bases_[size_t(r)] = p;
      ipcOpened_[size_t(r)] = true;
    }
  }
  comm_->barrier();
  // every peer has mapped the segment: the name can go, the memory lives until the last unmap
  if (localShm_) shm_unlink(shmName_);
}

SymmetricHeap::~SymmetricHeap() {
  const int me = rank_;
  for (size_t r = 0; r < bases_.size(); ++r) {
    if (ipcOpened_[r] && bases_[r]) cudaIpcCloseMemHandle(bases_[r]);
    if (shmMapped_[r] && bases_[r]) munmap(bases_[r], bytes_ ? bytes_ : 1);
  }
  if (memory_ == Memory::Host) {
    if (localShm_) munmap(bases_[size_t(me)], bytes_ ? bytes_ : 1);
    else std::free(bases_[size_t(me)]);
  } else if (bases_[size_t(me)]) {
    cudaFree(bases_[size_t(me)]);
  }
}

size_t SymmetricHeap::alloc(size_t bytes, size_t align) {
  const size_t off = (used_ + align - 1) / align * align;
  if (off + bytes > bytes_)
    THROW_RUNTIME("SymmetricHeap: out of space (" << off + bytes << " > " << bytes_ << ")");
  used_ = off + bytes;
  return off;
}

void SymmetricHeap::zero(size_t offset, size_t bytes) {
  if (bytes == size_t(-1)) bytes = bytes_ - offset;
  char *p = static_cast<char *>(local()) + offset;
  if (memory_ == Memory::Host) {
    std::memset(p, 0, bytes);
  } else {
    CUDA_RUNTIME(cudaMemset(p, 0, bytes));
    CUDA_RUNTIME(cudaDeviceSynchronize());
  }
}

} // namespace tenzing
