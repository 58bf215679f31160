// All-rank barrier executed on the GPUs: every rank's stream publishes an epoch to a flag word
// on every other rank and then waits until all of its own words carry that epoch.  Used to
// line up the start of a measurement across ranks to within a few microseconds (a host
// barrier leaves tens of microseconds of skew, which a device-timed exchange would then
// count as waiting time).  SURVEY section 7.4 item 7.
#pragma once

#include <memory>

#include <cuda_runtime_api.h>

#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/cuda/symmetric_heap.hpp"
#include "tenzing_b200.h"

namespace tenzing {

class DeviceBarrier {
  std::unique_ptr<SymmetricHeap> heap_;
  uint64_t epoch_ = 0;

public:
  // collective
  explicit DeviceBarrier(std::shared_ptr<Communicator> comm)
      : heap_(new SymmetricHeap(std::move(comm), 4096)) {}
  // enqueue the barrier on `stream` (all ranks must call it the same number of times)
  void enqueue(cudaStream_t stream) {
    ++epoch_;
    const int n = heap_->size(), me = heap_->rank();
    for (int r0 = 0; r0 < n; r0 += 32) {
      TzbFlagList l;
      l.n = 0;
      for (int r = r0; r < n && r < r0 + 32; ++r) l.flag[l.n++] = heap_->at<uint64_t>(r, 0) + me;
      CUDA_RUNTIME(cudaError_t(tzb_flags_signal(&l, epoch_, stream)));
    }
    CUDA_RUNTIME(cudaError_t(tzb_flags_wait(heap_->at<uint64_t>(me, 0), n, epoch_, stream)));
  }
};

} // namespace tenzing
