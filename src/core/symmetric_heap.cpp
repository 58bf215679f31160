#include "tenzing/cuda/symmetric_heap.hpp"

#include <cstdlib>
#include <cstring>

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

  Hello mine;
  std::memset(&mine, 0, sizeof(mine));
  mine.pid = int64_t(getpid());
  static const int64_t hostId = int64_t(gethostid()); // once: ranks may be threads
  mine.hostId = hostId;
  void *base = nullptr;
  if (memory_ == Memory::Host) {
    base = std::calloc(bytes_ ? bytes_ : 1, 1);
    if (!base) THROW_RUNTIME("SymmetricHeap: host allocation of " << bytes_ << " bytes failed");
    mine.device = -1;
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
        // dry-run across processes (no GPU): ops are no-ops in this mode and never
        // dereference a peer pointer, so the peer base stays null
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
}

SymmetricHeap::~SymmetricHeap() {
  const int me = rank_;
  for (size_t r = 0; r < bases_.size(); ++r)
    if (ipcOpened_[r] && bases_[r]) cudaIpcCloseMemHandle(bases_[r]);
  if (memory_ == Memory::Host) {
    std::free(bases_[size_t(me)]);
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
