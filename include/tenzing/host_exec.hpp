// Host restatement of the device kernels (include/tenzing_b200.h): the same operations on host
// memory, executed by the calling thread.
//
// Purpose: with a host-memory SymmetricHeap, a Null platform and
// Platform::set_host_execution(true), every op of the workloads EXECUTES instead of being a
// no-op, so a schedule can be run and checked against the oracle without a GPU - and with the
// ranks as threads of one process (LoopbackWorld) the whole cross-rank protocol runs for real:
// payload stores into the peer's memory, epoch flags published with release and polled with
// acquire semantics, double-buffered staging by epoch parity.  Under ThreadSanitizer that is a
// machine check of the protocol's ordering argument (scripts/sanitize.sh).
//
// The flag functions are the CPU counterpart of src/cuda/device_sync.cuh; a wait that is not
// satisfied within `wait_timeout_seconds()` throws, so a schedule that would dead-lock the GPUs
// fails a CPU test instead of hanging it.
//
// No counterpart in the reference (its ops only run on CUDA + MPI).
#pragma once

#include <cstddef>
#include <cstdint>

#include "tenzing_b200.h"

namespace tenzing {
namespace host {

// seconds a flag wait may take before it throws (default 20; env TENZING_HOST_WAIT_SECS)
double wait_timeout_seconds();
void set_wait_timeout_seconds(double s);

// flag <- epoch with release semantics (payload written before it is visible to the acquirer)
void flag_signal(uint64_t *flag, uint64_t epoch);
void flags_signal(const TzbFlagList &list, uint64_t epoch);
// spin (yielding) until every flags[i] >= epoch, acquire semantics
void flags_wait(const uint64_t *flags, int n, uint64_t epoch);

// Box copies: per box, wait for its `wait` flag, copy the loop nest, and publish `signal`:
// as soon as the box is complete if `ordered`, after all boxes otherwise (the device
// kernels' two plan kinds).
void boxcopy(const TzbBox *boxes, int nBoxes, int elemSize, uint64_t epoch, bool ordered);

void copy_bytes(void *dst, const void *src, size_t bytes);

template <typename T>
void spmv_csr(int rows, const int *rowPtr, const int *colInd, const T *val, const T *x, T *y,
              bool accumulate);
template <typename T> void gather(int n, const int *idx, const T *src, T *dst);
template <typename T> void vector_add(int n, const T *a, T *y);
// gather into every peer's buffer, then one flag per peer (tzb_gather_put_*)
template <typename T>
void gather_put(const TzbGatherPutArgs &args, const int *idx, const T *src, uint64_t epoch);
// wait for nFlags flags, then the product (tzb_spmv_csr_wait_*)
template <typename T>
void spmv_csr_wait(int rows, const int *rowPtr, const int *colInd, const T *val, const T *x, T *y,
                   bool accumulate, const uint64_t *flags, int nFlags, uint64_t epoch);
// pull form: wait for the ready flags, product with x read through xPtr, then the acks
// (tzb_spmv_csr_pull_*)
template <typename T>
void spmv_csr_pull(int rows, const int *rowPtr, const int *colInd, const T *val,
                   const T *const *xPtr, T *y, bool accumulate, const uint64_t *ready, int nReady,
                   const TzbFlagList &acks, uint64_t epoch);
// the whole distributed product (tzb_spmv_fused_all_*)
template <typename T>
void spmv_fused_all(const TzbGatherPutArgs &put, const int *sendIdx, int rows, const int *lRowPtr,
                    const int *lColInd, const T *lVal, const int *rRowPtr, const int *rColInd,
                    const T *rVal, const T *x, const T *rx, T *y, const uint64_t *flags, int nFlags,
                    uint64_t epoch);

} // namespace host
} // namespace tenzing
