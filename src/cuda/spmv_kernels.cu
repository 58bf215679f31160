// SpMV-side kernels for sm_100a: CSR product (optionally fused with the flag
// wait that replaces MPI_Wait), gather, gather fused with the peer put, and
// y-accumulate.  All are bandwidth-bound: the design goal is coalesced streams
// of (val, colInd), x gathers served from L2, and no host round trips.
//
// Reference counterparts: include/tenzing/spmv/ops_spmv.cuh:25-59 (scalar
// kernels), :91-126 (cuSPARSE call), :217-304 (Post/Wait Send/Recv),
// src/spmv/ops_spmv.cu:26-47.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdint>

#include "device_sync.cuh"
#include "tenzing_b200.h"

namespace tzb {

std::atomic<uint64_t> g_launches{0};
thread_local const uint64_t *g_epochSource = nullptr;

// the flag-wait watchdog of this translation unit's kernels (see device_sync.cuh)
cudaError_t set_wait_timeout_spmv(unsigned long long ns) { return set_wait_timeout_ns_this_tu(ns); }

namespace {

constexpr int kThreads = 256;

// ---------------------------------------------------------------------------------------
// CSR product, "stream" form.  A block owns a tile of kThreads consecutive rows.  The tile's
// non-zeros are one contiguous range of (val, colInd), so all threads stream that range with
// fully coalesced, mutually independent loads (U pairs per thread in flight, then U gathers
// of x), park the products in shared memory, and every thread then sums the segment of its own
// row.  Compared with one row per lane group (the round-1 kernel, kept below for dense rows)
// there is no rowPtr -> colInd -> x dependency chain per row and no idle lanes when rows are
// shorter than the lane group: measured 25 % of HBM bandwidth before, see profiles/.
//
// COHERENT: x is being written by peer GPUs until the flag wait returns (the remote x of the
// fused styles), so it must not be read through the non-coherent path.
template <typename T, bool COHERENT> __device__ __forceinline__ T load_x(const T *x, int c) {
  if (COHERENT) return __ldcg(x + c);
  return __ldg(x + c);
}

// products of tile rows [r0, r0 + kThreads) into `acc` of the thread that owns row r0 + tid
// tileRows (<= kThreads) rows per tile: small problems use small tiles so that every SM has one
template <typename T, int U, bool COHERENT>
__device__ __forceinline__ T tile_product(int r0, int rows, int tileRows, const int *__restrict__ rowPtr,
                                          const int *__restrict__ colInd, const T *__restrict__ val,
                                          const T *x, T *prod /* kThreads * U */, int *range /* 2 */) {
  const int t = threadIdx.x;
  const int row = r0 + t;
  int b = 0, e = 0;
  if (t < tileRows && row < rows) b = __ldg(rowPtr + row), e = __ldg(rowPtr + row + 1);
  __syncthreads(); // the previous tile's readers of `range` and `prod` are done
  if (t == 0) range[0] = b;
  if (t < tileRows && row == min(r0 + tileRows, rows) - 1) range[1] = e;
  __syncthreads();
  const int B = range[0], E = range[1];
  constexpr int CH = kThreads * U;
  T acc = 0;
  for (int c0 = B; c0 < E; c0 += CH) {
    T v[U];
    int c[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int j = c0 + t + kThreads * u;
      v[u] = 0, c[u] = -1;
      if (j < E) v[u] = __ldcs(val + j), c[u] = __ldcs(colInd + j);
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (c[u] >= 0) v[u] *= load_x<T, COHERENT>(x, c[u]);
#pragma unroll
    for (int u = 0; u < U; ++u) prod[t + kThreads * u] = v[u];
    __syncthreads();
    const int lo = max(b, c0) - c0, hi = min(e, c0 + CH) - c0;
    for (int j = lo; j < hi; ++j) acc += prod[j];
    if (c0 + CH < E) __syncthreads(); // before the next chunk overwrites the products
  }
  return acc;
}

template <typename T, int U, bool WAIT>
__global__ void __launch_bounds__(kThreads)
spmv_csr_stream_kernel(int rows, int tileRows, const int *__restrict__ rowPtr, const int *__restrict__ colInd,
                       const T *__restrict__ val, const T *x, T *__restrict__ y, int accumulate,
                       const uint64_t *flags, int nFlags, EpochArg epochArg) {
  const uint64_t epoch = epochArg.get();
  __shared__ T prod[kThreads * U];
  __shared__ int range[2];
  if (WAIT) {
    // remote x entries are only valid once every source rank has published this epoch
    if (int(threadIdx.x) < nFlags) wait_flag_geq(flags + threadIdx.x, epoch);
    __syncthreads();
  }
  const int tiles = (rows + tileRows - 1) / tileRows;
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int r0 = tile * tileRows;
    const T acc = tile_product<T, U, WAIT>(r0, rows, tileRows, rowPtr, colInd, val, x, prod, range);
    const int row = r0 + threadIdx.x;
    if (int(threadIdx.x) < tileRows && row < rows) y[row] = accumulate ? y[row] + acc : acc;
  }
}

// One row per group of TPR lanes: the form for long rows (more than ~48 non-zeros per row),
// where a row fills its lane group for several trips.
template <typename T, int TPR, bool WAIT>
__global__ void __launch_bounds__(kThreads)
spmv_csr_kernel(int rows, const int *__restrict__ rowPtr, const int *__restrict__ colInd,
                const T *__restrict__ val, const T *x, T *__restrict__ y, int accumulate,
                const uint64_t *flags, int nFlags, EpochArg epochArg) {
  const uint64_t epoch = epochArg.get();
  if (WAIT) {
    if (int(threadIdx.x) < nFlags) wait_flag_geq(flags + threadIdx.x, epoch);
    __syncthreads();
  }
  const int lane = threadIdx.x % TPR;
  // the lanes of MY row only: groups of one warp leave the row loop (or skip empty rows)
  // independently, so a full-warp mask would name lanes that never arrive
  const unsigned groupMask =
      TPR == 32 ? 0xffffffffu : (((1u << TPR) - 1u) << ((threadIdx.x & 31) / TPR * TPR));
  const int groupsPerBlock = kThreads / TPR;
  for (int row = blockIdx.x * groupsPerBlock + threadIdx.x / TPR; row < rows;
       row += gridDim.x * groupsPerBlock) {
    const int begin = __ldg(rowPtr + row), end = __ldg(rowPtr + row + 1);
    T acc = 0;
    for (int j = begin + lane; j < end; j += TPR)
      acc += __ldg(val + j) * load_x<T, WAIT>(x, __ldg(colInd + j));
#pragma unroll
    for (int off = TPR / 2; off > 0; off /= 2) acc += __shfl_down_sync(groupMask, acc, off, TPR);
    if (lane == 0) y[row] = accumulate ? y[row] + acc : acc;
  }
}

// resident blocks per SM of a kernel (occupancy API, cached by the caller)
template <typename K> int blocks_per_sm(K kernel) {
  int n = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kernel, kThreads, 0) != cudaSuccess || n < 1) {
    (void)cudaGetLastError();
    n = 1;
  }
  return n;
}

// rows per tile: 256 where that still gives every SM two tiles, else 128 or 64 (a tile is one
// dependent chain rowPtr -> (val, colInd) -> x -> sum, so a small problem wants many small tiles)
inline int stream_tile_rows(int rows) {
  const long want = 2L * sm_count();
  if (rows >= want * kThreads) return kThreads;
  if (rows >= want * (kThreads / 2)) return kThreads / 2;
  return kThreads / 4;
}
// pairs per thread in flight for a given density: the tile's non-zeros should fit one chunk
inline int stream_unroll(int nnzPerRowHint) { return nnzPerRowHint <= 4 ? 4 : (nnzPerRowHint <= 8 ? 8 : 12); }
constexpr int kStreamMaxNnzPerRow = 48;

template <typename T, bool WAIT>
cudaError_t launch_spmv(int rows, const int *rowPtr, const int *colInd, const T *val, const T *x,
                        T *y, int accumulate, const uint64_t *flags, int nFlags, uint64_t epoch,
                        int nnzPerRowHint, cudaStream_t stream) {
  if (rows <= 0) return cudaSuccess;
  if (WAIT && nFlags > kThreads) return cudaErrorInvalidValue;
  ++g_launches;
  if (nnzPerRowHint <= kStreamMaxNnzPerRow) {
    const int tileRows = stream_tile_rows(rows);
    const int tiles = (rows + tileRows - 1) / tileRows;
#define TZB_CASE(N)                                                                                \
  case N: {                                                                                        \
    static const int perSm = blocks_per_sm(spmv_csr_stream_kernel<T, N, WAIT>);                    \
    const int blocks = std::min(tiles, sm_count() * perSm);                                       \
    spmv_csr_stream_kernel<T, N, WAIT><<<blocks, kThreads, 0, stream>>>(                           \
        rows, tileRows, rowPtr, colInd, val, x, y, accumulate, flags, nFlags, ep(epoch));             \
    break;                                                                                         \
  }
    switch (stream_unroll(nnzPerRowHint)) {
      TZB_CASE(4)
      TZB_CASE(8)
      TZB_CASE(12)
    }
#undef TZB_CASE
    return cudaGetLastError();
  }
  // lanes per row: the largest power of two not above nnz per row (2..32), so few lanes idle
  int tpr = 2;
  while (tpr < 32 && tpr * 2 <= nnzPerRowHint) tpr *= 2;
  const int groupsPerBlock = kThreads / tpr;
  int blocks = (rows + groupsPerBlock - 1) / groupsPerBlock;
  const int maxBlocks = sm_count() * 8; // 8 x 256 threads = full occupancy, grid-stride beyond
  if (blocks > maxBlocks) blocks = maxBlocks;
#define TZB_CASE(N)                                                                                \
  case N:                                                                                          \
    spmv_csr_kernel<T, N, WAIT><<<blocks, kThreads, 0, stream>>>(rows, rowPtr, colInd, val, x, y,  \
                                                                 accumulate, flags, nFlags, ep(epoch)); \
    break
  switch (tpr) {
  default:
    TZB_CASE(32);
  }
#undef TZB_CASE
  return cudaGetLastError();
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
gather_kernel(int n, const int *__restrict__ idx, const T *__restrict__ src, T *__restrict__ dst) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    dst[i] = src[__ldg(idx + i)];
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
vector_add_kernel(int n, const T *__restrict__ a, T *__restrict__ y) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    y[i] += a[i];
}

// Gather straight into the peers' receive buffers (stores travel over NVLink),
// then one release per peer.  Consecutive threads write consecutive peer
// addresses, so the remote stores leave the SM as full 32 B sectors.
template <typename T>
__global__ void __launch_bounds__(kThreads)
gather_put_kernel(TzbGatherPutArgs a, const int *__restrict__ idx, const T *__restrict__ src,
                  EpochArg epochArg, unsigned *counter) {
  const uint64_t epoch = epochArg.get();
  const int n = a.displ[a.nPeers];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    int p = 0;
    while (i >= a.displ[p + 1]) ++p; // <= 16 peers, displ lives in constant param space
    static_cast<T *>(a.peerDst[p])[i - a.displ[p]] = src[__ldg(idx + i)];
  }
  if (block_arrive_last(counter, gridDim.x)) {
    for (int p = threadIdx.x; p < a.nPeers; p += blockDim.x)
      if (a.peerFlag[p]) st_release_sys(a.peerFlag[p], epoch);
  }
}

// The whole distributed product in ONE launch: (1) gather my x entries straight into the
// peers' receive buffers and publish the epoch, (2) local product while those stores and the
// peers' stores are in flight, (3) wait for every peer's flag, (4) accumulate the remote
// product.  No block waits for another block of the same GPU; the only waiting is on flags
// written by other GPUs, whose phase (1) does not depend on anything here.  A rank publishes
// only after ALL its blocks have passed phase (1) while resident blocks may already spin in
// phase (3), so the whole grid has to be co-resident: the launcher sizes it from the occupancy
// API and launches cooperatively, which the driver refuses rather than deadlocks on.
// Blocks whose rows have no remote non-zeros never look at the flags.
template <typename T, int U>
__global__ void __launch_bounds__(kThreads)
spmv_fused_all_kernel(TzbGatherPutArgs a, const int *__restrict__ sendIdx, int rows, int tileRows,
                      const int *__restrict__ lRowPtr, const int *__restrict__ lColInd,
                      const T *__restrict__ lVal, const int *__restrict__ rRowPtr,
                      const int *__restrict__ rColInd, const T *__restrict__ rVal,
                      const T *__restrict__ x, const T *rx, T *__restrict__ y,
                      const uint64_t *flags, int nFlags, EpochArg epochArg, unsigned *counter) {
  const uint64_t epoch = epochArg.get();
  __shared__ T prod[kThreads * U];
  __shared__ int range[2];
  // (1) put
  const int nSend = a.displ[a.nPeers];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nSend; i += gridDim.x * blockDim.x) {
    int p = 0;
    while (i >= a.displ[p + 1]) ++p;
    static_cast<T *>(a.peerDst[p])[i - a.displ[p]] = x[__ldg(sendIdx + i)];
  }
  if (block_arrive_last(counter, gridDim.x)) {
    for (int p = threadIdx.x; p < a.nPeers; p += blockDim.x)
      if (a.peerFlag[p]) st_release_sys(a.peerFlag[p], epoch);
  }
  // (2) local product
  const int tiles = (rows + tileRows - 1) / tileRows;
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int r0 = tile * tileRows;
    const T acc = tile_product<T, U, false>(r0, rows, tileRows, lRowPtr, lColInd, lVal, x, prod, range);
    if (int(threadIdx.x) < tileRows && r0 + int(threadIdx.x) < rows) y[r0 + threadIdx.x] = acc;
  }
  // (3) + (4): wait for the peers (only blocks that need remote entries), remote product;
  // the same thread wrote y[row] above, so there is no inter-thread hazard
  bool waited = false;
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int r0 = tile * tileRows, r1 = min(r0 + tileRows, rows);
    if (__ldg(rRowPtr + r0) == __ldg(rRowPtr + r1)) continue; // block-uniform
    if (!waited) {
      if (int(threadIdx.x) < nFlags) wait_flag_geq(flags + threadIdx.x, epoch);
      __syncthreads();
      waited = true;
    }
    const T acc = tile_product<T, U, true>(r0, rows, tileRows, rRowPtr, rColInd, rVal, rx, prod, range);
    if (int(threadIdx.x) < tileRows && r0 + int(threadIdx.x) < rows) y[r0 + threadIdx.x] += acc;
  }
}

// One-to-many put through NVLS: every store goes to a MULTICAST address, NVSwitch replicates it
// into the same offset of every rank's buffer (the reference sends the same x entries to each
// peer in a loop of MPI_Isend, include/tenzing/spmv/ops_spmv.cuh:217-240).  The epoch flag is
// a multicast store too: word `myRank` of every rank's flag array, after a system-scope fence.
__device__ __forceinline__ void multimem_st(double *p, double v) {
  asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
}
__device__ __forceinline__ void multimem_st(float *p, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
template <typename T>
__global__ void __launch_bounds__(kThreads)
multicast_put_kernel(int n, const T *__restrict__ src, T *mcDst, uint64_t *mcFlag, EpochArg epochArg,
                     unsigned *counter) {
  const uint64_t epoch = epochArg.get();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    multimem_st(mcDst + i, __ldg(src + i));
  if (block_arrive_last(counter, gridDim.x)) {
    if (threadIdx.x == 0)
      asm volatile("multimem.st.release.sys.global.u64 [%0], %1;" ::"l"(mcFlag), "l"(epoch) : "memory");
  }
}

__global__ void epoch_advance_kernel(uint64_t *word) { *word += 1; }
__global__ void epoch_set_kernel(uint64_t *word, uint64_t value) { *word = value; }

int grid_for(int n) {
  int blocks = (n + kThreads - 1) / kThreads;
  if (blocks > sm_count() * 8) blocks = sm_count() * 8;
  return blocks < 1 ? 1 : blocks;
}

template <typename T>
cudaError_t launch_gather_put(const TzbGatherPutArgs *args, const int *idx, const T *src,
                              uint64_t epoch, uint32_t *counter, cudaStream_t stream) {
  if (!args || args->nPeers < 0 || args->nPeers > 16) return cudaErrorInvalidValue;
  const int n = args->displ[args->nPeers];
  ++g_launches;
  gather_put_kernel<T><<<grid_for(n), kThreads, 0, stream>>>(*args, idx, src, ep(epoch), counter);
  return cudaGetLastError();
}

// average nnz per row is not known to the C entry points; 8 lanes per row suits
// the ~10 nnz/row matrices of the shipped workloads (the C++ ops pass the real value)
constexpr int kDefaultNnzPerRow = 8;

} // namespace

// C++ entry with an explicit lanes-per-row hint (used by the SpMV ops)
cudaError_t spmv_csr(int rows, const int *rowPtr, const int *colInd, const double *val,
                     const double *x, double *y, bool accumulate, int nnzPerRow,
                     cudaStream_t stream) {
  return launch_spmv<double, false>(rows, rowPtr, colInd, val, x, y, accumulate, nullptr, 0, 0,
                                    nnzPerRow, stream);
}
cudaError_t spmv_csr(int rows, const int *rowPtr, const int *colInd, const float *val,
                     const float *x, float *y, bool accumulate, int nnzPerRow,
                     cudaStream_t stream) {
  return launch_spmv<float, false>(rows, rowPtr, colInd, val, x, y, accumulate, nullptr, 0, 0,
                                   nnzPerRow, stream);
}
cudaError_t spmv_csr_wait(int rows, const int *rowPtr, const int *colInd, const double *val,
                          const double *x, double *y, bool accumulate, int nnzPerRow,
                          const uint64_t *flags, int nFlags, uint64_t epoch, cudaStream_t stream) {
  return launch_spmv<double, true>(rows, rowPtr, colInd, val, x, y, accumulate, flags, nFlags,
                                   epoch, nnzPerRow, stream);
}
cudaError_t spmv_csr_wait(int rows, const int *rowPtr, const int *colInd, const float *val,
                          const float *x, float *y, bool accumulate, int nnzPerRow,
                          const uint64_t *flags, int nFlags, uint64_t epoch, cudaStream_t stream) {
  return launch_spmv<float, true>(rows, rowPtr, colInd, val, x, y, accumulate, flags, nFlags,
                                  epoch, nnzPerRow, stream);
}

template <typename T>
cudaError_t launch_spmv_fused_all(const TzbGatherPutArgs *a, const int *sendIdx, int rows,
                                  const int *lRowPtr, const int *lColInd, const T *lVal,
                                  const int *rRowPtr, const int *rColInd, const T *rVal, const T *x,
                                  const T *rx, T *y, const uint64_t *flags, int nFlags,
                                  uint64_t epoch, uint32_t *counter, int nnzPerRow,
                                  cudaStream_t stream) {
  if (!a || a->nPeers < 0 || a->nPeers > 16 || nFlags > kThreads) return cudaErrorInvalidValue;
  int tileRows = stream_tile_rows(rows);
  const int tiles = std::max(1, (rows + tileRows - 1) / tileRows);
  const int nSend = a->displ[a->nPeers];
  TzbGatherPutArgs args = *a;
  EpochArg epochArg = ep(epoch);
  void *params[] = {&args,    &sendIdx, &rows, &tileRows, &lRowPtr, &lColInd, &lVal,  &rRowPtr, &rColInd,
                    &rVal,    &x,       &rx,   &y,       &flags,   &nFlags, &epochArg, &counter};
  ++g_launches;
#define TZB_CASE(N)                                                                                \
  case N: {                                                                                        \
    static const int perSm = blocks_per_sm(spmv_fused_all_kernel<T, N>);                           \
    /* enough blocks for the tiles and for the gather, never more than fit at once */              \
    const int want = std::max(tiles, (nSend + kThreads - 1) / kThreads);                           \
    const int blocks = std::max(1, std::min(want, sm_count() * perSm));                            \
    return cudaLaunchCooperativeKernel(reinterpret_cast<void *>(spmv_fused_all_kernel<T, N>),      \
                                       dim3(blocks), dim3(kThreads), params, 0, stream);           \
  }
  switch (stream_unroll(nnzPerRow)) {
    TZB_CASE(4)
    TZB_CASE(8)
    TZB_CASE(12)
  }
#undef TZB_CASE
  return cudaErrorInvalidValue;
}

cudaError_t spmv_fused_all(const TzbGatherPutArgs *a, const int *sendIdx, int rows,
                           const int *lRowPtr, const int *lColInd, const double *lVal,
                           const int *rRowPtr, const int *rColInd, const double *rVal,
                           const double *x, const double *rx, double *y, const uint64_t *flags,
                           int nFlags, uint64_t epoch, uint32_t *counter, int nnzPerRow,
                           cudaStream_t stream) {
  return launch_spmv_fused_all<double>(a, sendIdx, rows, lRowPtr, lColInd, lVal, rRowPtr, rColInd,
                                       rVal, x, rx, y, flags, nFlags, epoch, counter, nnzPerRow, stream);
}
cudaError_t spmv_fused_all(const TzbGatherPutArgs *a, const int *sendIdx, int rows,
                           const int *lRowPtr, const int *lColInd, const float *lVal,
                           const int *rRowPtr, const int *rColInd, const float *rVal,
                           const float *x, const float *rx, float *y, const uint64_t *flags,
                           int nFlags, uint64_t epoch, uint32_t *counter, int nnzPerRow,
                           cudaStream_t stream) {
  return launch_spmv_fused_all<float>(a, sendIdx, rows, lRowPtr, lColInd, lVal, rRowPtr, rColInd,
                                      rVal, x, rx, y, flags, nFlags, epoch, counter, nnzPerRow, stream);
}

// Pull form of the remote product ("gather fused with the off-diagonal SpMV"): no staging
// copy of x at all.  xPtr[c] points at the owner's copy of remote column c (a peer pointer
// into that rank's x), so the gather happens inside the product, over NVLink.  Every block
// first waits until all ranks have declared their x ready for this epoch; the last block to
// finish tells every rank that this rank is done reading (ackFlags), after which the owners
// may overwrite their x.
template <typename T, int TPR>
__global__ void __launch_bounds__(kThreads)
spmv_csr_pull_kernel(int rows, const int *__restrict__ rowPtr, const int *__restrict__ colInd,
                     const T *__restrict__ val, const T *const *__restrict__ xPtr,
                     T *__restrict__ y, int accumulate, const uint64_t *ready, int nReady,
                     TzbFlagList acks, EpochArg epochArg, unsigned *counter) {
  const uint64_t epoch = epochArg.get();
  if (int(threadIdx.x) < nReady) wait_flag_geq(ready + threadIdx.x, epoch);
  __syncthreads();
  const int lane = threadIdx.x % TPR;
  const unsigned groupMask =
      TPR == 32 ? 0xffffffffu : (((1u << TPR) - 1u) << ((threadIdx.x & 31) / TPR * TPR));
  const int groupsPerBlock = kThreads / TPR;
  for (int row = blockIdx.x * groupsPerBlock + threadIdx.x / TPR; row < rows;
       row += gridDim.x * groupsPerBlock) {
    const int begin = __ldg(rowPtr + row), end = __ldg(rowPtr + row + 1);
    T acc = 0;
    for (int j = begin + lane; j < end; j += TPR) {
      const T *xp = reinterpret_cast<const T *>(
          __ldg(reinterpret_cast<const unsigned long long *>(xPtr) + __ldg(colInd + j)));
      acc += __ldg(val + j) * *xp; // peer load; x does not change while the product runs
    }
#pragma unroll
    for (int off = TPR / 2; off > 0; off /= 2) acc += __shfl_down_sync(groupMask, acc, off, TPR);
    if (lane == 0) y[row] = accumulate ? y[row] + acc : acc;
  }
  if (block_arrive_last(counter, gridDim.x)) {
    if (int(threadIdx.x) < acks.n && acks.flag[threadIdx.x])
      st_release_sys(acks.flag[threadIdx.x], epoch);
  }
}

template <typename T>
cudaError_t launch_spmv_pull(int rows, const int *rowPtr, const int *colInd, const T *val,
                             const T *const *xPtr, T *y, int accumulate, const uint64_t *ready,
                             int nReady, const TzbFlagList *acks, uint64_t epoch, uint32_t *counter,
                             int nnzPerRowHint, cudaStream_t stream) {
  if (!acks || nReady > kThreads || acks->n > 32) return cudaErrorInvalidValue;
  int tpr = 2;
  while (tpr < 32 && tpr * 2 <= nnzPerRowHint) tpr *= 2;
  const int groupsPerBlock = kThreads / tpr;
  // at least one block even without rows: the handshake must still happen
  int blocks = (std::max(rows, 1) + groupsPerBlock - 1) / groupsPerBlock;
  if (blocks > sm_count() * 8) blocks = sm_count() * 8;
  ++g_launches;
#define TZB_CASE(N)                                                                                \
  case N:                                                                                          \
    spmv_csr_pull_kernel<T, N><<<blocks, kThreads, 0, stream>>>(                                   \
        rows, rowPtr, colInd, val, xPtr, y, accumulate, ready, nReady, *acks, ep(epoch), counter);    \
    break
  switch (tpr) {
    TZB_CASE(2);
    TZB_CASE(4);
    TZB_CASE(8);
    TZB_CASE(16);
    TZB_CASE(32);
  }
#undef TZB_CASE
  return cudaGetLastError();
}

cudaError_t spmv_csr_pull(int rows, const int *rowPtr, const int *colInd, const double *val,
                          const double *const *xPtr, double *y, bool accumulate,
                          const uint64_t *ready, int nReady, const TzbFlagList *acks, uint64_t epoch,
                          uint32_t *counter, int nnzPerRow, cudaStream_t stream) {
  return launch_spmv_pull<double>(rows, rowPtr, colInd, val, xPtr, y, accumulate, ready, nReady,
                                  acks, epoch, counter, nnzPerRow, stream);
}
cudaError_t spmv_csr_pull(int rows, const int *rowPtr, const int *colInd, const float *val,
                          const float *const *xPtr, float *y, bool accumulate,
                          const uint64_t *ready, int nReady, const TzbFlagList *acks, uint64_t epoch,
                          uint32_t *counter, int nnzPerRow, cudaStream_t stream) {
  return launch_spmv_pull<float>(rows, rowPtr, colInd, val, xPtr, y, accumulate, ready, nReady,
                                 acks, epoch, counter, nnzPerRow, stream);
}

} // namespace tzb

using namespace tzb;

extern "C" {

int tzb_spmv_csr_f64(int rows, const int *rowPtr, const int *colInd, const double *val,
                     const double *x, double *y, int accumulate, void *stream) {
  return launch_spmv<double, false>(rows, rowPtr, colInd, val, x, y, accumulate, nullptr, 0, 0,
                                    kDefaultNnzPerRow, static_cast<cudaStream_t>(stream));
}
int tzb_spmv_csr_f32(int rows, const int *rowPtr, const int *colInd, const float *val,
                     const float *x, float *y, int accumulate, void *stream) {
  return launch_spmv<float, false>(rows, rowPtr, colInd, val, x, y, accumulate, nullptr, 0, 0,
                                   kDefaultNnzPerRow, static_cast<cudaStream_t>(stream));
}
int tzb_spmv_csr_wait_f64(int rows, const int *rowPtr, const int *colInd, const double *val,
                          const double *x, double *y, int accumulate, const uint64_t *flags,
                          int nFlags, uint64_t epoch, void *stream) {
  return launch_spmv<double, true>(rows, rowPtr, colInd, val, x, y, accumulate, flags, nFlags,
                                   epoch, kDefaultNnzPerRow, static_cast<cudaStream_t>(stream));
}
int tzb_spmv_csr_wait_f32(int rows, const int *rowPtr, const int *colInd, const float *val,
                          const float *x, float *y, int accumulate, const uint64_t *flags,
                          int nFlags, uint64_t epoch, void *stream) {
  return launch_spmv<float, true>(rows, rowPtr, colInd, val, x, y, accumulate, flags, nFlags,
                                  epoch, kDefaultNnzPerRow, static_cast<cudaStream_t>(stream));
}
int tzb_spmv_csr_pull_f64(int rows, const int *rowPtr, const int *colInd, const double *val,
                          const double *const *xPtr, double *y, int accumulate,
                          const uint64_t *ready, int nReady, const TzbFlagList *acks, uint64_t epoch,
                          uint32_t *counter, void *stream) {
  return launch_spmv_pull<double>(rows, rowPtr, colInd, val, xPtr, y, accumulate, ready, nReady,
                                  acks, epoch, counter, kDefaultNnzPerRow,
                                  static_cast<cudaStream_t>(stream));
}
int tzb_spmv_csr_pull_f32(int rows, const int *rowPtr, const int *colInd, const float *val,
                          const float *const *xPtr, float *y, int accumulate,
                          const uint64_t *ready, int nReady, const TzbFlagList *acks, uint64_t epoch,
                          uint32_t *counter, void *stream) {
  return launch_spmv_pull<float>(rows, rowPtr, colInd, val, xPtr, y, accumulate, ready, nReady,
                                 acks, epoch, counter, kDefaultNnzPerRow,
                                 static_cast<cudaStream_t>(stream));
}

int tzb_gather_f64(int n, const int *idx, const double *src, double *dst, void *stream) {
  if (n <= 0) return 0;
  ++g_launches;
  gather_kernel<double><<<grid_for(n), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(n, idx, src, dst);
  return cudaGetLastError();
}
int tzb_gather_f32(int n, const int *idx, const float *src, float *dst, void *stream) {
  if (n <= 0) return 0;
  ++g_launches;
  gather_kernel<float><<<grid_for(n), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(n, idx, src, dst);
  return cudaGetLastError();
}
int tzb_vector_add_f64(int n, const double *a, double *y, void *stream) {
  if (n <= 0) return 0;
  ++g_launches;
  vector_add_kernel<double><<<grid_for(n), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(n, a, y);
  return cudaGetLastError();
}
int tzb_vector_add_f32(int n, const float *a, float *y, void *stream) {
  if (n <= 0) return 0;
  ++g_launches;
  vector_add_kernel<float><<<grid_for(n), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(n, a, y);
  return cudaGetLastError();
}
int tzb_gather_put_f64(const TzbGatherPutArgs *args, const int *idx, const double *src,
                       uint64_t epoch, uint32_t *counter, void *stream) {
  return launch_gather_put<double>(args, idx, src, epoch, counter, static_cast<cudaStream_t>(stream));
}
int tzb_gather_put_f32(const TzbGatherPutArgs *args, const int *idx, const float *src,
                       uint64_t epoch, uint32_t *counter, void *stream) {
  return launch_gather_put<float>(args, idx, src, epoch, counter, static_cast<cudaStream_t>(stream));
}

int tzb_spmv_fused_all_f64(const TzbGatherPutArgs *put, const int *sendIdx, int rows,
                           const int *lRowPtr, const int *lColInd, const double *lVal,
                           const int *rRowPtr, const int *rColInd, const double *rVal,
                           const double *x, const double *rx, double *y, const uint64_t *flags,
                           int nFlags, uint64_t epoch, uint32_t *counter, void *stream) {
  return launch_spmv_fused_all<double>(put, sendIdx, rows, lRowPtr, lColInd, lVal, rRowPtr, rColInd,
                                       rVal, x, rx, y, flags, nFlags, epoch, counter,
                                       kDefaultNnzPerRow, static_cast<cudaStream_t>(stream));
}
int tzb_spmv_fused_all_f32(const TzbGatherPutArgs *put, const int *sendIdx, int rows,
                           const int *lRowPtr, const int *lColInd, const float *lVal,
                           const int *rRowPtr, const int *rColInd, const float *rVal,
                           const float *x, const float *rx, float *y, const uint64_t *flags,
                           int nFlags, uint64_t epoch, uint32_t *counter, void *stream) {
  return launch_spmv_fused_all<float>(put, sendIdx, rows, lRowPtr, lColInd, lVal, rRowPtr, rColInd,
                                      rVal, x, rx, y, flags, nFlags, epoch, counter,
                                      kDefaultNnzPerRow, static_cast<cudaStream_t>(stream));
}

int tzb_multicast_put_f64(int n, const double *src, double *mcDst, uint64_t *mcFlag, uint64_t epoch,
                          uint32_t *counter, void *stream) {
  if (n < 0 || !mcFlag) return cudaErrorInvalidValue;
  ++g_launches;
  multicast_put_kernel<double><<<grid_for(std::max(n, 1)), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      n, src, mcDst, mcFlag, ep(epoch), counter);
  return cudaGetLastError();
}
int tzb_multicast_put_f32(int n, const float *src, float *mcDst, uint64_t *mcFlag, uint64_t epoch,
                          uint32_t *counter, void *stream) {
  if (n < 0 || !mcFlag) return cudaErrorInvalidValue;
  ++g_launches;
  multicast_put_kernel<float><<<grid_for(std::max(n, 1)), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      n, src, mcDst, mcFlag, ep(epoch), counter);
  return cudaGetLastError();
}

void tzb_set_epoch_source(const uint64_t *devWord) { g_epochSource = devWord; }

int tzb_epoch_advance(uint64_t *devWord, void *stream) {
  if (!devWord) return cudaErrorInvalidValue;
  epoch_advance_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(devWord);
  return cudaGetLastError();
}

int tzb_epoch_set(uint64_t *devWord, uint64_t value, void *stream) {
  if (!devWord) return cudaErrorInvalidValue;
  epoch_set_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(devWord, value);
  return cudaGetLastError();
}

uint64_t tzb_launch_count(void) { return g_launches.load(); }

} // extern "C"
