// Device-side building blocks of the cross-GPU flag protocol (sm_100a).
//
// The reference synchronises ranks on the host (MPI_Wait + cudaEventSynchronize,
// SURVEY §2.8).  Here a producer kernel publishes an epoch with a system-scope
// release store into the consumer GPU's flag word and the consumer kernel
// spins with system-scope acquire loads - no host round trip.
#pragma once

#include <cstdint>

namespace tzb {

__device__ __forceinline__ void st_release_sys(uint64_t *p, uint64_t v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// spin until *p >= epoch; back off a little so the polling does not flood L2
__device__ __forceinline__ void wait_flag_geq(const uint64_t *p, uint64_t epoch) {
  unsigned ns = 20;
  while (ld_acquire_sys(p) < epoch) {
    __nanosleep(ns);
    if (ns < 320) ns *= 2;
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

} // namespace tzb
