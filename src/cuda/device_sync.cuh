// Device-side building blocks of the cross-GPU flag protocol (sm_100a).
//
// The reference synchronises ranks on the host (MPI_Wait + cudaEventSynchronize,
// SURVEY §2.8).  Here a producer kernel publishes an epoch with a system-scope
// release store into the consumer GPU's flag word and the consumer kernel
// spins with system-scope acquire loads - no host round trip.
#pragma once

#include <cstdint>

#include <cuda_runtime.h>

namespace tzb {

// A flag wait that is not satisfied within this many nanoseconds ends the kernel with a trap
// (the launch then reports cudaErrorLaunchFailed and the process fails) instead of spinning
// for ever behind a schedule that never publishes or a peer that died.  One copy per
// translation unit; tzb_set_wait_timeout_ms() sets all of them.  0 = wait for ever.
static __device__ unsigned long long g_wait_timeout_ns = 20ull * 1000ull * 1000ull * 1000ull;

// host side of the above for the including translation unit
static inline cudaError_t set_wait_timeout_ns_this_tu(unsigned long long ns) {
  return cudaMemcpyToSymbol(g_wait_timeout_ns, &ns, sizeof(ns));
}

// How a kernel gets its epoch.  Normally by value.  A sequence captured into a CUDA graph is
// replayed many times with the same arguments, so there the value comes from a device word that
// the first node of the graph increments (tzb_set_epoch_source / tzb_epoch_advance).
struct EpochArg {
  uint64_t value;
  const uint64_t *dev;
  __device__ __forceinline__ uint64_t get() const { return dev ? *dev : value; }
};
// defined in spmv_kernels.cu: the device word the launch wrappers of THIS THREAD hand to kernels, or null
extern thread_local const uint64_t *g_epochSource;
inline EpochArg ep(uint64_t value) { return EpochArg{value, g_epochSource}; }

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ void st_release_sys(uint64_t *p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// spin until *p >= epoch; back off a little so the polling does not flood L2.  The watchdog
// reads the clock only once the back-off has reached its ceiling, so a wait that is satisfied
// within a few microseconds never pays for it.
__device__ __forceinline__ void wait_flag_geq(const uint64_t *p, uint64_t epoch) {
  unsigned ns = 20;
  unsigned long long t0 = 0;
  while (ld_acquire_sys(p) < epoch) {
    __nanosleep(ns);
    if (ns < 320) {
      ns *= 2;
    } else {
      const unsigned long long limit = g_wait_timeout_ns;
      if (limit) {
        const unsigned long long now = globaltimer_ns();
        if (!t0) t0 = now;
        else if (now - t0 > limit) __trap();
      }
    }
  }
}

// Block-level "this block's part is done": the last block to arrive (of
// `expected`) gets true and the counter is reset to zero for the next launch.
// All threads of the block must call it; prior global/peer writes of the
// whole block are made visible system-wide before the count is taken.
__device__ __forceinline__ bool block_arrive_last(unsigned *counter, unsigned expected) {
  __shared__ bool isLast;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    const unsigned prev = atomicAdd(counter, 1u);
    isLast = (prev == expected - 1);
    if (isLast) {
      *counter = 0;
      __threadfence_system();
    }
  }
  __syncthreads();
  return isLast;
}

// number of SMs of the current device (cached per device; 148 on B200)
inline int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!cached[dev]) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

} // namespace tzb
