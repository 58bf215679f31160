// Batched strided box copies for sm_100a: one kernel moves any set of halo
// regions (pack, unpack, pack fused with the put into a peer GPU, direct put
// into the peer's ghost cells), waits on / publishes the cross-GPU flags of
// each region, and is cut into equal work items so all 148 SMs stay busy
// whatever the mix of thick faces, thin faces, edges and corners.
//
// Reference counterparts: pack/unpack kernels
// src/halo_exchange/ops_halo_exchange.cu:519-664 (which ignore their thread
// indices) and the MPI_Isend/Irecv/Wait ops around them.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "device_sync.cuh"
#include "tenzing_b200.h"

namespace tzb {

extern std::atomic<uint64_t> g_launches;
cudaError_t set_wait_timeout_spmv(unsigned long long ns);

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kBlocksPerSM = 4;
// a work item moves about this many bytes: small enough that the largest
// workload yields several items per resident block, large enough to amortise
// the per-item bookkeeping
constexpr int64_t kItemBytes = 32 * 1024;
constexpr int kThin = 8; // elements per thread in flight, thin rows (also the item quantum)

struct alignas(16) BoxDesc {
  const char *src;
  char *dst;
  int64_t s1, s2, s3, d1, d2, d3; // element strides
  int32_t n1, n2, n3, run;
  uint64_t *signal;
  const uint64_t *wait;
  int32_t nItems; // work items this box was cut into
  // thin rows (run * 8 bytes <= 32): how each side is accessed, see thin_item().
  // bits 0-1 source mode, 2-3 destination mode (kElem / kDense / kWindow), 4-5 element offset of
  // the source piece inside its aligned 32-byte window, 6-7 the same for the destination
  int32_t thin;
  uint64_t *ready; // published at kernel start (permission for the partner to overwrite), or null
};

enum ThinMode : int { kElem = 0, kDense = 1, kWindow = 2 };
constexpr int kThinEnabled = 256; // BoxDesc::thin: this box takes the sector path

// rows [i1, i1+c1) x [i2, i2+c2) of plane i3 of box `box`
struct alignas(8) WorkItem {
  int32_t box, i3, i2, c2, i1, c1;
};

struct PlanHeader {
  int32_t nBoxes, nItems, elemSize, pad;
  // followed by BoxDesc[nBoxes], WorkItem[nItems]
};

// Memory-level parallelism is what this kernel lives on: the regions are far
// larger than L2 reuse could help and every byte is touched once, so each lane
// keeps kThick/kThin independent 8-byte loads in flight before the first store
// (thick rows: one warp per row; thin rows: lanes run over (row, element)
// pairs so the dense side of a pack/unpack stays fully coalesced).

// ---------------------------------------------------------------------------------------
// Thin rows, one row per thread ("sector path", 8-byte elements, run <= 4).
//
// An x-face row piece of the XYZQ grid is 3 doubles inside a 4 KB row.  Measured on B200
// (scripts/probes/thin_probe.cu, profiles/): DRAM delivers a whole 128-byte line per piece
// whatever the load width or engine (8-byte LDG, 256-bit LDG, TMA tile with L2 promotion off
// all read 100.7 MB for 18.9 MB of payload and take the same 22 us), so the lever is not the
// bytes but the instruction count and, for stores, avoiding read-modify-write: a 24-byte
// store into a 32-byte sector makes L2 fetch the sector first (2x the time of the loads),
// a full-sector store does not.  Therefore:
//   * a strided piece that lies inside one aligned 32-byte window is LOADED with one 256-bit
//     access (kWindow) instead of `run` scalar ones;
//   * where the caller says the rest of the destination window is padding it may clobber
//     (TzbBox::dstSlackLo/Hi), the piece is STORED as one full 256-bit sector;
//   * the dense side of a pack / unpack goes through shared memory so that global (or peer)
//     memory sees contiguous 16-byte accesses.
struct U256 {
  unsigned long long a, b, c, d;
};
__device__ __forceinline__ U256 ld256(const void *p) {
  U256 v;
  asm volatile("ld.global.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(v.a), "=l"(v.b), "=l"(v.c), "=l"(v.d) : "l"(p));
  return v;
}
__device__ __forceinline__ void st256(void *p, const U256 &v) {
  asm volatile("st.global.v4.u64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(v.a), "l"(v.b), "l"(v.c), "l"(v.d) : "memory");
}

constexpr int kThinRows = kThreads;               // rows per chunk: one row per thread in flight
constexpr int kThinSmemElems = kThinRows * 4;     // dense image of a chunk (run <= 4)
// One row per thread is enough: 1024 resident threads per SM keep 128 KB of DRAM lines in
// flight, several times what the measured 24 G rows/s ceiling needs.

// elements [off, off + 4) of a window, off uniform: no dynamic register indexing
__device__ __forceinline__ void window_extract(const U256 &w, int off, unsigned long long (&e)[4]) {
  e[0] = off == 0 ? w.a : (off == 1 ? w.b : (off == 2 ? w.c : w.d));
  e[1] = off == 0 ? w.b : (off == 1 ? w.c : w.d);
  e[2] = off == 0 ? w.c : w.d;
  e[3] = w.d;
}
__device__ __forceinline__ U256 window_build(const unsigned long long (&e)[4], int off) {
  U256 w;
  w.a = e[0];
  w.b = off == 0 ? e[1] : e[0];
  w.c = off == 0 ? e[2] : (off == 1 ? e[1] : e[0]);
  w.d = off == 0 ? e[3] : (off == 1 ? e[2] : (off == 2 ? e[1] : e[0]));
  return w;
}

// One work item of a thin box of 8-byte elements; `sm` holds kThinSmemElems elements.
// srcMode / dstMode: kWindow = strided piece inside one aligned 32-byte window (a single 256-bit
// access; for a destination the rest of the window is padding), kDense = the item's rows are
// contiguous there (moved through shared memory with 16-byte accesses).
template <int srcMode, int dstMode>
__device__ __forceinline__ void thin_item_t(const BoxDesc &b, const WorkItem &w, unsigned long long *sm) {
  const int run = b.run;
  const int srcOff = (b.thin >> 4) & 3, dstOff = (b.thin >> 6) & 3;
  const int rows = w.c1 * w.c2, c1 = w.c1;
  const unsigned long long *src = reinterpret_cast<const unsigned long long *>(b.src) + int64_t(w.i3) * b.s3 +
                                  int64_t(w.i2) * b.s2 + int64_t(w.i1) * b.s1;
  unsigned long long *dst = reinterpret_cast<unsigned long long *>(b.dst) + int64_t(w.i3) * b.d3 +
                            int64_t(w.i2) * b.d2 + int64_t(w.i1) * b.d1;
  const int t = threadIdx.x;
#pragma unroll 1
  for (int r0 = 0; r0 < rows; r0 += kThinRows) {
    const int nr = min(kThinRows, rows - r0);
    const int n = nr * run; // elements of this chunk on a dense side
    if (srcMode == kDense) {
      const unsigned long long *sp = src + int64_t(r0) * run;
      if (((reinterpret_cast<uintptr_t>(sp) & 15) == 0) && (n & 1) == 0) {
#pragma unroll 2
        for (int i = t; i < n / 2; i += kThreads)
          reinterpret_cast<ulonglong2 *>(sm)[i] = __ldcs(reinterpret_cast<const ulonglong2 *>(sp) + i);
      } else {
#pragma unroll 2
        for (int i = t; i < n; i += kThreads) sm[i] = __ldcs(sp + i);
      }
      __syncthreads();
    }
    const int r = r0 + t;
    const bool live = t < nr;
    int j2 = 0, j1 = r;
    if (w.c2 > 1) j2 = r / c1, j1 = r - j2 * c1;
    unsigned long long e[4] = {0, 0, 0, 0};
    if (live) {
      if (srcMode == kWindow) {
        const U256 win = ld256(src + int64_t(j2) * b.s2 + int64_t(j1) * b.s1 - srcOff);
        window_extract(win, srcOff, e);
      } else {
        const unsigned long long *p = sm + t * run;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (i < run) e[i] = p[i];
      }
    }
    if (srcMode == kDense && dstMode == kDense) __syncthreads();
    if (dstMode == kDense) {
      if (live) {
        unsigned long long *p = sm + t * run;
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (i < run) p[i] = e[i];
      }
      __syncthreads();
      unsigned long long *dp = dst + int64_t(r0) * run;
      if (((reinterpret_cast<uintptr_t>(dp) & 15) == 0) && (n & 1) == 0) {
#pragma unroll 1
        for (int i = t; i < n / 2; i += kThreads)
          reinterpret_cast<ulonglong2 *>(dp)[i] = reinterpret_cast<const ulonglong2 *>(sm)[i];
      } else {
#pragma unroll 1
        for (int i = t; i < n; i += kThreads) dp[i] = sm[i];
      }
    } else if (live) {
      st256(dst + int64_t(j2) * b.d2 + int64_t(j1) * b.d1 - dstOff, window_build(e, dstOff));
    }
    __syncthreads(); // shared memory is free for the next chunk
  }
}

__device__ __forceinline__ void thin_item(const BoxDesc &b, const WorkItem &w, unsigned long long *sm) {
  switch (b.thin & 15) { // block-uniform
  case kWindow | (kDense << 2): thin_item_t<kWindow, kDense>(b, w, sm); break;   // pack
  case kDense | (kWindow << 2): thin_item_t<kDense, kWindow>(b, w, sm); break;   // unpack
  default: thin_item_t<kWindow, kWindow>(b, w, sm); break;                       // direct
  }
}

// 256-bit accesses on thick rows (TZB_WIDE256=1): off by default until measured over NVLink
__device__ int kWide256 = 0;

// one work item: rows [i1, i1+c1) x [i2, i2+c2) of plane i3
template <typename T, bool VEC, int kThick, int kThin>
__device__ __forceinline__ void copy_item(const BoxDesc &b, const WorkItem &w, int lane, int warp) {
  const T *src = reinterpret_cast<const T *>(b.src) + int64_t(w.i3) * b.s3;
  T *dst = reinterpret_cast<T *>(b.dst) + int64_t(w.i3) * b.d3;
  const int run = b.run;
  const int rows = w.c1 * w.c2;
  if (run >= 64) {
    for (int r = warp; r < rows; r += kWarps) {
      const int j2 = r / w.c1, j1 = r - j2 * w.c1;
      const T *s = src + int64_t(w.i2 + j2) * b.s2 + int64_t(w.i1 + j1) * b.s1;
      T *d = dst + int64_t(w.i2 + j2) * b.d2 + int64_t(w.i1 + j1) * b.d1;
      if (VEC && sizeof(T) == 8 && kWide256 && (run & 3) == 0 &&
          ((reinterpret_cast<uintptr_t>(s) | reinterpret_cast<uintptr_t>(d)) & 31) == 0) {
        // both sides 32-byte aligned (grid rows from the first interior cell are): 256-bit
        // loads and stores, a warp moves 1 KB per instruction
        const int quads = run >> 2;
        int p = lane;
        for (; p + 32 < quads; p += 64) {
          const U256 a = ld256(s + 4 * p), b = ld256(s + 4 * (p + 32));
          st256(d + 4 * p, a);
          st256(d + 4 * (p + 32), b);
        }
        for (; p < quads; p += 32) st256(d + 4 * p, ld256(s + 4 * p));
        continue;
      }
      if (VEC && sizeof(T) == 8) {
        // 16-byte accesses on whichever side is 16-byte aligned (staging buffers always are;
        // grid rows start at x = nGhost and usually are not): over NVLink a warp then stores
        // 512 contiguous, aligned bytes instead of 256 that straddle sectors
        const bool sVec = (reinterpret_cast<uintptr_t>(s) & 15) == 0;
        const bool dVec = (reinterpret_cast<uintptr_t>(d) & 15) == 0;
        if ((run & 1) == 0 && (sVec || dVec)) {
          const int pairs = run >> 1;
          constexpr int kPairs = kThick / 2;
          int p = lane;
          for (; p + 32 * (kPairs - 1) < pairs; p += 32 * kPairs) {
            double2 v[kPairs];
#pragma unroll
            for (int u = 0; u < kPairs; ++u) {
              const double *sp = reinterpret_cast<const double *>(s) + 2 * (p + 32 * u);
              if (sVec) v[u] = __ldcs(reinterpret_cast<const double2 *>(sp));
              else v[u] = make_double2(__ldcs(sp), __ldcs(sp + 1));
            }
#pragma unroll
            for (int u = 0; u < kPairs; ++u) {
              double *dp = reinterpret_cast<double *>(d) + 2 * (p + 32 * u);
              if (dVec) *reinterpret_cast<double2 *>(dp) = v[u];
              else dp[0] = v[u].x, dp[1] = v[u].y;
            }
          }
          for (; p < pairs; p += 32) {
            d[2 * p] = s[2 * p];
            d[2 * p + 1] = s[2 * p + 1];
          }
          continue;
        }
      }
      int e = lane;
      for (; e + 32 * (kThick - 1) < run; e += 32 * kThick) {
        T v[kThick];
#pragma unroll
        for (int u = 0; u < kThick; ++u) v[u] = __ldcs(s + e + 32 * u);
#pragma unroll
        for (int u = 0; u < kThick; ++u) d[e + 32 * u] = v[u];
      }
      for (; e < run; e += 32) d[e] = s[e];
    }
  } else {
    // the plan sizes thin items to a multiple of kThreads * kThin elements where it can, so
    // the scalar tail below is the exception
    const int total = rows * run;
    const int c1 = w.c1;
    auto offsets = [&](int f, int64_t &so, int64_t &dof) {
      const int r = f / run, e = f - r * run;
      const int j2 = r / c1, j1 = r - j2 * c1;
      so = int64_t(w.i2 + j2) * b.s2 + int64_t(w.i1 + j1) * b.s1 + e;
      dof = int64_t(w.i2 + j2) * b.d2 + int64_t(w.i1 + j1) * b.d1 + e;
    };
    int f = threadIdx.x;
    for (; f + kThreads * (kThin - 1) < total; f += kThreads * kThin) {
      T v[kThin];
      int64_t dof[kThin];
#pragma unroll
      for (int u = 0; u < kThin; ++u) {
        int64_t so;
        offsets(f + kThreads * u, so, dof[u]);
        v[u] = src[so];
      }
#pragma unroll
      for (int u = 0; u < kThin; ++u) dst[dof[u]] = v[u];
    }
    for (; f < total; f += kThreads) {
      int64_t so, dof;
      offsets(f, so, dof);
      dst[dof] = src[so];
    }
  }
}

// Memory-level parallelism is what these kernels live on: the regions are far larger than L2
// reuse could help and every byte is touched once, so each lane keeps kThick/kThin
// independent 8-byte loads in flight before the first store (thick rows: one warp per row;
// thin rows: lanes run over (row, element) pairs so the dense side of a pack/unpack stays
// fully coalesced).
//
// Full-occupancy variant: no flag waits, flags (if any) published once by the last block.
template <typename T, bool VEC>
__global__ void __launch_bounds__(kThreads, kBlocksPerSM)
box_copy_kernel(const BoxDesc *__restrict__ boxes, const WorkItem *__restrict__ items, int nItems,
                int nBoxes, int anySignal, EpochArg epochArg, unsigned *counter) {
  const uint64_t epoch = epochArg.get();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ alignas(16) unsigned long long thinSm[kThinSmemElems];
  for (int it = blockIdx.x; it < nItems; it += gridDim.x) {
    const WorkItem w = items[it];
    const BoxDesc &b = boxes[w.box];
    if (sizeof(T) == 8 && (b.thin & kThinEnabled)) thin_item(b, w, thinSm);
    else copy_item<T, VEC, 8, kThin>(b, w, lane, warp);
  }
  if (anySignal) {
    // one system-scope fence per block; the last block to finish publishes every flag
    if (block_arrive_last(counter, gridDim.x)) {
      for (int bx = threadIdx.x; bx < nBoxes; bx += kThreads)
        if (boxes[bx].signal) st_release_sys(boxes[bx].signal, epoch);
    }
  }
}

// Synchronising variant: waits for region flags, and for ordered plans publishes each region's
// flag as soon as that region is complete - a block counts the items it finished of the current
// region and hands the count in (one system-scope fence) when it moves on to another region.
// Ordered plans (MIN_BLOCKS = 2) are launched with a grid of 2 blocks per SM.  Registers are
// bounded as for the full-occupancy variant (4 blocks of 256 threads fit an SM), so half of
// every SM's register file and block slots stay free for the partner kernel they are meant to
// overlap with.
template <typename T, int MIN_BLOCKS>
__global__ void __launch_bounds__(kThreads, kBlocksPerSM)
box_copy_sync_kernel(const BoxDesc *__restrict__ boxes, const WorkItem *__restrict__ items,
                     int nItems, int nBoxes, int anySignal, EpochArg epochArg, unsigned *counter) {
  const uint64_t epoch = epochArg.get();
  constexpr bool kPerBox = MIN_BLOCKS < kBlocksPerSM; // the ordered (half-occupancy) variant
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ unsigned readyMask; // boxes whose wait flag this block has already seen
  __shared__ alignas(16) unsigned long long thinSm[kThinSmemElems];
  if (threadIdx.x == 0) readyMask = 0;
  // permissions first (a few of the earliest blocks repeat the same stores: whichever of them
  // the hardware starts first gets them out, nothing here may wait before that)
  if (blockIdx.x < 4)
    for (int bx = threadIdx.x; bx < nBoxes; bx += kThreads)
      if (boxes[bx].ready) st_release_sys(boxes[bx].ready, epoch);
  __syncthreads();
  int curBox = -1;
  unsigned curCount = 0;
  auto hand_in = [&](int box, unsigned n) {
    __syncthreads(); // all threads' stores of these items are issued
    if (threadIdx.x == 0) {
      __threadfence_system();
      const unsigned prev = atomicAdd(counter + 1 + box, n);
      if (prev + n == unsigned(boxes[box].nItems)) {
        counter[1 + box] = 0;
        if (boxes[box].signal) st_release_sys(boxes[box].signal, epoch);
      }
    }
  };
  for (int it = blockIdx.x; it < nItems; it += gridDim.x) {
    const WorkItem w = items[it];
    const BoxDesc &b = boxes[w.box];
    if (kPerBox) {
      if (w.box != curBox) {
        if (curBox >= 0) hand_in(curBox, curCount);
        curBox = w.box;
        curCount = 0;
      }
      ++curCount;
    }
    if (b.wait) {
      // the data of this box is only there once the producer published this epoch
      // the mask remembers the first 32 boxes only; later ones are polled every time (a
      // satisfied flag costs one load)
      if (threadIdx.x == 0 && !(w.box < 32 && ((readyMask >> w.box) & 1u))) {
        wait_flag_geq(b.wait, epoch);
        if (w.box < 32) readyMask |= 1u << w.box;
      }
      __syncthreads();
    }
    if (sizeof(T) == 8 && (b.thin & kThinEnabled)) thin_item(b, w, thinSm);
    else copy_item<T, true, 8, kThin>(b, w, lane, warp);
  }
  if (kPerBox) {
    if (curBox >= 0) hand_in(curBox, curCount);
  } else if (anySignal) {
    if (block_arrive_last(counter, gridDim.x)) {
      for (int bx = threadIdx.x; bx < nBoxes; bx += kThreads)
        if (boxes[bx].signal) st_release_sys(boxes[bx].signal, epoch);
    }
  }
}

// ---------------------------------------------------------------------------------------
// Opt-in (TZB_BULK=1): the copy engine of the SM does the moving.
// For plans whose every row is 16-byte aligned on both sides and at least 512 bytes long (the
// "thick" regions: 4 KB rows of the y/z faces and edges, i.e. the data that crosses NVLink
// un-staged), one thread per block drives a ring of shared-memory stages with
//   cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes   (global -> smem)
//   cp.async.bulk.global.shared::cta.bulk_group                         (smem -> global/peer)
// so no registers or LSU slots are spent on the payload and the NVLink side sees one bulk
// store per row instead of thousands of 16-byte ones.  Blocks are one warp; up to seven fit
// on an SM next to each other (32 KB of stages each).
// Status: validated on B200s against the halo oracle (SASS: UBLKCP, SYNCS.ARRIVE.TRANS64).  It
// moves an 18.9 MB face in the same 8.9 us as the LSU path locally and is no faster over NVLink
// (81.8 vs 79.9 us for the thick put at 2 GPUs), so it stays off unless TZB_BULK=1.
constexpr int kBulkStages = 4;
constexpr int kBulkStageBytes = 8192;

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__global__ void __launch_bounds__(32)
box_bulk_kernel(const BoxDesc *__restrict__ boxes, const WorkItem *__restrict__ items, int nItems,
                int nBoxes, int elemSize, int anySignal, EpochArg epochArg, unsigned *counter) {
  const uint64_t epoch = epochArg.get();
  __shared__ alignas(128) unsigned char stage[kBulkStages][kBulkStageBytes];
  __shared__ alignas(8) uint64_t bar[kBulkStages];
  if (threadIdx.x == 0) {
    for (int s = 0; s < kBulkStages; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[s])) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncwarp();
  if (threadIdx.x == 0) {
    uint32_t g = 0; // segments issued so far by this block: stage = g % kBulkStages, use = g / kBulkStages
    for (int it = blockIdx.x; it < nItems; it += gridDim.x) {
      const WorkItem w = items[it];
      const BoxDesc &b = boxes[w.box];
      const int64_t rowBytes = int64_t(b.run) * elemSize;
      const int cpr = int((rowBytes + kBulkStageBytes - 1) / kBulkStageBytes); // chunks per row
      const int nseg = w.c1 * w.c2 * cpr;
      const char *src = b.src + int64_t(w.i3) * b.s3 * elemSize;
      char *dst = b.dst + int64_t(w.i3) * b.d3 * elemSize;
      auto locate = [&](int k, const char *&sp, char *&dp, uint32_t &bytes) {
        const int r = k / cpr, c = k - r * cpr;
        const int j2 = r / w.c1, j1 = r - j2 * w.c1;
        const int64_t off = int64_t(c) * kBulkStageBytes;
        sp = src + (int64_t(w.i2 + j2) * b.s2 + int64_t(w.i1 + j1) * b.s1) * elemSize + off;
        dp = dst + (int64_t(w.i2 + j2) * b.d2 + int64_t(w.i1 + j1) * b.d1) * elemSize + off;
        const int64_t left = rowBytes - off;
        bytes = uint32_t(left < kBulkStageBytes ? left : kBulkStageBytes);
      };
      auto issue_load = [&](int k) {
        const char *sp;
        char *dp;
        uint32_t bytes;
        locate(k, sp, dp, bytes);
        const uint32_t s = (g + uint32_t(k)) % kBulkStages;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%1], %0;" ::"r"(bytes),
                     "r"(smem_u32(&bar[s]))
                     : "memory");
        asm volatile(
            "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                smem_u32(stage[s])),
            "l"(sp), "r"(bytes), "r"(smem_u32(&bar[s]))
            : "memory");
      };
      const int pre = nseg < kBulkStages ? nseg : kBulkStages;
      for (int k = 0; k < pre; ++k) issue_load(k);
      for (int k = 0; k < nseg; ++k) {
        const uint32_t n = g + uint32_t(k);
        const uint32_t s = n % kBulkStages, parity = (n / kBulkStages) & 1u;
        asm volatile(
            "{\n\t"
            ".reg .pred P1;\n\t"
            "WAIT_%=:\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
            "@P1 bra DONE_%=;\n\t"
            "bra WAIT_%=;\n\t"
            "DONE_%=:\n\t"
            "}" ::"r"(smem_u32(&bar[s])),
            "r"(parity)
            : "memory");
        const char *sp;
        char *dp;
        uint32_t bytes;
        locate(k, sp, dp, bytes);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dp),
                     "r"(smem_u32(stage[s])), "r"(bytes)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        // refill the stage of the PREVIOUS segment: its store has finished reading shared
        // memory once at most one bulk group (the one just committed) is still pending
        if (k >= 1 && k - 1 + kBulkStages < nseg) {
          asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          issue_load(k - 1 + kBulkStages);
        }
      }
      // all stages free again before the next item's prologue reuses them
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      g += uint32_t(nseg);
    }
    // stores complete (not merely read) before this block reports in
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    asm volatile("fence.proxy.async;" ::: "memory");
  }
  __syncwarp();
  if (anySignal) {
    if (block_arrive_last(counter, gridDim.x)) {
      for (int bx = threadIdx.x; bx < nBoxes; bx += 32)
        if (boxes[bx].signal) st_release_sys(boxes[bx].signal, epoch);
    }
  }
}

__global__ void flags_signal_kernel(TzbFlagList l, EpochArg epochArg) {
  const uint64_t epoch = epochArg.get();
  if (int(threadIdx.x) < l.n) {
    __threadfence_system();
    st_release_sys(l.flag[threadIdx.x], epoch);
  }
}

__global__ void flags_wait_kernel(const uint64_t *flags, int n, EpochArg epochArg) {
  const uint64_t epoch = epochArg.get();
  for (int i = threadIdx.x; i < n; i += blockDim.x) wait_flag_geq(flags + i, epoch);
}

// L2 flush for the benchmarks: write the buffer (evicts everything else), then read it back so
// that what stays in L2 is CLEAN - otherwise the next kernel pays for writing back up to 126 MB
// of the flush buffer's dirty lines inside its own timed region.
__global__ void __launch_bounds__(kThreads) flush_write_kernel(uint4 *p, size_t n16, unsigned v) {
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n16; i += stride) p[i] = make_uint4(v, v, v, v);
}
__global__ void __launch_bounds__(kThreads) flush_read_kernel(const uint4 *p, size_t n16, unsigned *sink) {
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  unsigned acc = 0;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n16; i += stride) {
    const uint4 v = __ldcg(p + i);
    acc ^= v.x ^ v.y ^ v.z ^ v.w;
  }
  if (acc == 0x9e3779b9u) *sink = acc; // never true for a buffer of equal words; keeps the loads alive
}

// keeps a stream busy for `ns` nanoseconds: lets the host queue what follows before it starts
__global__ void delay_kernel(unsigned long long ns) {
  const unsigned long long t0 = globaltimer_ns();
  while (globaltimer_ns() - t0 < ns) __nanosleep(200);
}

__global__ void __launch_bounds__(kThreads)
copy_bytes_kernel(uint4 *__restrict__ dst, const uint4 *__restrict__ src, size_t n16) {
  size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (; i + 3 * stride < n16; i += 4 * stride) {
    const uint4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
    dst[i] = a;
    dst[i + stride] = b;
    dst[i + 2 * stride] = c;
    dst[i + 3 * stride] = d;
  }
  for (; i < n16; i += stride) dst[i] = src[i];
}

} // namespace
} // namespace tzb

using namespace tzb;

extern "C" {

// TZB_THIN_SECTOR=0 plans thin boxes for the element-wise path of round 1 (A/B comparisons)
static int thin_sector_level() {
  static const int level = [] {
    const char *e = std::getenv("TZB_THIN_SECTOR");
    return e ? std::atoi(e) : 1;
  }();
  return level;
}
static bool thin_sector_enabled() { return thin_sector_level() > 0; }

static int boxcopy_plan_impl(const TzbBox *boxes, int nBoxes, int elemSize, TzbBoxPlanInfo *info,
                             void *hostImage, bool ordered) {
  if (!info || nBoxes < 0 || (elemSize != 4 && elemSize != 8)) return cudaErrorInvalidValue;
  std::vector<BoxDesc> descs(nBoxes);
  std::vector<WorkItem> items;
  size_t payload = 0;
  // item size: enough items to give every resident block several, within [4 KB, kItemBytes]
  int64_t totalBytes = 0;
  for (int b = 0; b < nBoxes; ++b)
    totalBytes += int64_t(boxes[b].run) * elemSize * boxes[b].n1 * boxes[b].n2 * boxes[b].n3;
  const int64_t itemBytes =
      std::min<int64_t>(kItemBytes, std::max<int64_t>(4096, totalBytes / (int64_t(sm_count()) * kBlocksPerSM * 4)));
  for (int b = 0; b < nBoxes; ++b) {
    const TzbBox &in = boxes[b];
    if (in.n1 < 0 || in.n2 < 0 || in.n3 < 0 || in.run < 0) return cudaErrorInvalidValue;
    BoxDesc &d = descs[b];
    d.src = static_cast<const char *>(in.src);
    d.dst = static_cast<char *>(in.dst);
    d.s1 = in.s1, d.s2 = in.s2, d.s3 = in.s3;
    d.d1 = in.d1, d.d2 = in.d2, d.d3 = in.d3;
    d.n1 = in.n1, d.n2 = in.n2, d.n3 = in.n3, d.run = in.run;
    d.signal = in.signal;
    d.wait = in.wait;
    d.ready = in.ready;
    d.thin = 0;
    if (elemSize == 8 && in.run >= 1 && in.run <= 4 && thin_sector_enabled()) {
      // how can each side of this thin box be accessed? (see thin_item)
      auto dense = [&](int64_t t1, int64_t t2) { return t1 == in.run && (in.n2 <= 1 || t2 == t1 * in.n1); };
      auto window = [&](const void *p, int64_t t1, int64_t t2, int64_t t3, int &off) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        off = int((a & 31) / 8);
        return (a & 7) == 0 && (t1 * 8) % 32 == 0 && (t2 * 8) % 32 == 0 && (t3 * 8) % 32 == 0 &&
               off + in.run <= 4;
      };
      int so = 0, dof = 0;
      const bool sDense = dense(in.s1, in.s2), dDense = dense(in.d1, in.d2);
      const bool sWin = !sDense && window(in.src, in.s1, in.s2, in.s3, so);
      // a full-sector store is only allowed where the rest of the window is the caller's padding
      const bool dWin = !dDense && window(in.dst, in.d1, in.d2, in.d3, dof) && in.dstSlackLo >= dof &&
                        in.dstSlackHi >= 4 - dof - in.run;
      // pack (window -> dense), unpack (dense -> window) and direct (window -> window); every
      // other combination stays on the element-wise path of copy_item
      // Measured (profiles/README.md): only the unpack gains - its full-sector stores spare L2
      // the fetch of a partially written sector; pack and direct copies are bound by DRAM line
      // fetches either way and the element-wise path keeps more loads in flight per thread.
      // TZB_THIN_SECTOR=2 routes all three through the sector path (A/B comparisons).
      if ((sDense && dWin) || (thin_sector_level() >= 2 && ((sWin && dDense) || (sWin && dWin)))) {
        const int sm = sDense ? kDense : kWindow, dm = dDense ? kDense : kWindow;
        d.thin = kThinEnabled | sm | (dm << 2) | ((sWin ? so : 0) << 4) | ((dWin ? dof : 0) << 6);
      }
    }
    const int64_t rowBytes = int64_t(in.run) * elemSize;
    const int64_t volume = rowBytes * in.n1 * in.n2 * in.n3;
    payload += size_t(volume);
    int count = 0;
    if (volume == 0) {
      // an empty box still has to wait/publish its flags
      if (in.signal || in.wait || in.ready) {
        items.push_back(WorkItem{b, 0, 0, 0, 0, 0});
        count = 1;
      }
    } else {
      // rows per item, then cut (i2, i1) into c2 x c1 tiles of about that many rows
      // thick rows are copied one warp per row: never hand a block fewer rows than it has warps
      const int64_t minRows = in.run >= 64 ? kWarps : 1;
      int64_t rowsPerItem =
          std::max<int64_t>(minRows, itemBytes / std::max<int64_t>(rowBytes, 1));
      if (in.run < 64) {
        // thin rows: make rows*run a multiple of the unrolled batch (kThreads * kThin elements)
        const int64_t batch = int64_t(kThreads) * kThin;
        int64_t g = batch, r = in.run;
        while (r) { const int64_t t = g % r; g = r; r = t; } // gcd(batch, run)
        const int64_t quantum = batch / g;
        if (int64_t(in.n1) * in.n2 >= quantum) rowsPerItem = (rowsPerItem + quantum - 1) / quantum * quantum;
      }
      int c1 = int(std::min<int64_t>(in.n1, rowsPerItem));
      int c2 = int(std::min<int64_t>(in.n2, std::max<int64_t>(1, rowsPerItem / c1)));
      for (int i3 = 0; i3 < in.n3; ++i3)
        for (int i2 = 0; i2 < in.n2; i2 += c2)
          for (int i1 = 0; i1 < in.n1; i1 += c1) {
            items.push_back(WorkItem{b, i3, i2, std::min(c2, in.n2 - i2), i1, std::min(c1, in.n1 - i1)});
            ++count;
          }
    }
    d.nItems = count;
  }
  // interleave the boxes' items so that concurrently running blocks touch
  // different regions (and different peers) instead of draining one box at a time
  if (nBoxes > 1 && !ordered) {
    std::vector<std::vector<WorkItem>> perBox(nBoxes);
    for (const WorkItem &w : items) perBox[w.box].push_back(w);
    std::vector<WorkItem> mixed;
    mixed.reserve(items.size());
    std::vector<size_t> pos(nBoxes, 0);
    // proportional round-robin: each sweep takes from box b a share matching its size
    size_t maxItems = 1;
    for (auto &v : perBox) maxItems = std::max(maxItems, v.size());
    for (size_t step = 1; step <= maxItems; ++step) {
      for (int b = 0; b < nBoxes; ++b) {
        const size_t want = perBox[b].size() * step / maxItems;
        while (pos[b] < want) mixed.push_back(perBox[b][pos[b]++]);
      }
    }
    items.swap(mixed);
  }
  info->nBoxes = nBoxes;
  info->nItems = int(items.size());
  info->elemSize = elemSize;
  info->reserved = 0; // bit 0: some box publishes a flag; bit 1: some box waits for one
  for (int b = 0; b < nBoxes; ++b) {
    if (boxes[b].signal) info->reserved |= 1;
    if (boxes[b].wait || boxes[b].ready) info->reserved |= 2; // the synchronising kernel
  }
  if (ordered) info->reserved |= 4; // publish per box, share the SMs with a partner kernel
  {
    // bit 6: most of the payload goes to another GPU - the link is the bottleneck
    int64_t remoteBytes = 0;
    for (int b = 0; b < nBoxes; ++b)
      if (boxes[b].dstRemote)
        remoteBytes += int64_t(boxes[b].run) * elemSize * boxes[b].n1 * boxes[b].n2 * boxes[b].n3;
    bool thinSource = false; // strided thin sources need the full grid's loads in flight
    for (int b = 0; b < nBoxes; ++b) thinSource = thinSource || (boxes[b].run > 0 && boxes[b].run < 64);
    if (remoteBytes > 0 && remoteBytes * 2 >= totalBytes && !thinSource) info->reserved |= 64;
  }
  for (int b = 0; b < nBoxes; ++b) { // some thick box has a 16-byte aligned side: use the vector variant
    const TzbBox &in = boxes[b];
    if (elemSize != 8 || in.run < 64 || (in.run & 1)) continue;
    auto aligned = [&](const void *p, int64_t s1, int64_t s2, int64_t s3) {
      return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (s1 & 1) == 0 && (s2 & 1) == 0 && (s3 & 1) == 0;
    };
    if (aligned(in.src, in.s1, in.s2, in.s3) || aligned(in.dst, in.d1, in.d2, in.d3)) info->reserved |= 8;
  }
  // bit 4: every row of every box can go through the bulk-copy engine (16-byte aligned on both
  // sides, a multiple of 16 bytes and at least 512 bytes long); see box_bulk_kernel
  bool bulk = nBoxes > 0;
  for (int b = 0; b < nBoxes && bulk; ++b) {
    const TzbBox &in = boxes[b];
    const int64_t rowBytes = int64_t(in.run) * elemSize;
    auto ok = [&](const void *p, int64_t s1, int64_t s2, int64_t s3) {
      return (reinterpret_cast<uintptr_t>(p) & 15) == 0 && (s1 * elemSize) % 16 == 0 &&
             (s2 * elemSize) % 16 == 0 && (s3 * elemSize) % 16 == 0;
    };
    bulk = rowBytes >= 512 && rowBytes % 16 == 0 && ok(in.src, in.s1, in.s2, in.s3) &&
           ok(in.dst, in.d1, in.d2, in.d3);
  }
  if (bulk) info->reserved |= 16;
  info->planBytes = sizeof(PlanHeader) + sizeof(BoxDesc) * size_t(nBoxes) + sizeof(WorkItem) * items.size();
  info->scratchBytes = (sizeof(unsigned) * size_t(nBoxes + 1) + 255) / 256 * 256; // arrival counters
  info->payloadBytes = payload;
  if (hostImage) {
    char *p = static_cast<char *>(hostImage);
    PlanHeader h{nBoxes, int(items.size()), elemSize, 0};
    std::memcpy(p, &h, sizeof(h));
    p += sizeof(h);
    if (nBoxes) std::memcpy(p, descs.data(), sizeof(BoxDesc) * size_t(nBoxes));
    p += sizeof(BoxDesc) * size_t(nBoxes);
    if (!items.empty()) std::memcpy(p, items.data(), sizeof(WorkItem) * items.size());
  }
  return cudaSuccess;
}

int tzb_boxcopy_plan(const TzbBox *boxes, int nBoxes, int elemSize, TzbBoxPlanInfo *info,
                     void *hostImage) {
  return boxcopy_plan_impl(boxes, nBoxes, elemSize, info, hostImage, false);
}

int tzb_boxcopy_plan_ordered(const TzbBox *boxes, int nBoxes, int elemSize, TzbBoxPlanInfo *info,
                             void *hostImage) {
  return boxcopy_plan_impl(boxes, nBoxes, elemSize, info, hostImage, true);
}

static bool bulk_enabled() {
  static const bool on = [] {
    const char *e = std::getenv("TZB_BULK");
    return e && std::atoi(e) != 0;
  }();
  return on;
}

int tzb_boxcopy_launch(const void *devPlan, const TzbBoxPlanInfo *info, uint64_t epoch,
                       void *scratch, void *stream) {
  if (!devPlan || !info) return cudaErrorInvalidValue;
  static const bool wideSet = [] {
    const char *e = std::getenv("TZB_WIDE256");
    const int v = (e && std::atoi(e) != 0) ? 1 : 0;
    if (v) (void)cudaMemcpyToSymbol(kWide256, &v, sizeof(v));
    return true;
  }();
  (void)wideSet;
  if (info->nItems == 0) return cudaSuccess;
  const char *p = static_cast<const char *>(devPlan) + sizeof(PlanHeader);
  const BoxDesc *boxes = reinterpret_cast<const BoxDesc *>(p);
  const WorkItem *items =
      reinterpret_cast<const WorkItem *>(p + sizeof(BoxDesc) * size_t(info->nBoxes));
  // ordered plans are meant to run beside a partner kernel (a put beside the unpack that follows
  // it region by region): their grid is at most two blocks per SM and their registers are
  // bounded so that four blocks fit (checked below with the occupancy API), which leaves room
  // for the partner launched from another stream - room, not a guarantee: a consumer that spins
  // on flags of THIS GPU's producer must still be ordered behind it by an event (the workloads
  // do that for self-neighbours)
  const bool waits = (info->reserved & 2) != 0, ordered = (info->reserved & 4) != 0;
  const bool vec = (info->reserved & 8) != 0;
  // Grid size.  Every block works through kItemsPerBlock (4, TZB_ITEMS_PER_BLOCK overrides;
  // 0 = a resident grid that strides over all items) work items and retires: with short-lived
  // blocks a kernel launched from another stream gets SM slots as blocks retire, which is what
  // lets e.g. the unpack of early regions run beside the put of later ones (measured: 288 ->
  // 244 us for the 8-GPU exchange, profiles/README.md).  Ordered plans always use a resident
  // grid of half the SM slots: their consumer may spin before the local producer has finished,
  // so the producer must be guaranteed a place next to it.
  static const int itemsPerBlock = [] {
    const char *e = std::getenv("TZB_ITEMS_PER_BLOCK");
    return e ? std::max(0, std::atoi(e)) : 4;
  }();
  // Link-bound plans (thick rows into another GPU): NVLink is saturated by two blocks per SM
  // (TZB_LINK_BLOCKS_PER_SM overrides), and every further resident block would only hold an SM
  // slot for the whole transfer that a kernel on another stream - the unpack of what has
  // already arrived - could use.
  static const int linkBlocksPerSM = [] {
    const char *e = std::getenv("TZB_LINK_BLOCKS_PER_SM");
    return e ? std::max(1, std::atoi(e)) : 2; // measured: 195 / 166 / 155 us with 4 / 1 / 2 (profiles/README.md)
  }();
  const bool linkBound = (info->reserved & 64) != 0;
  auto grid = [&](int cap) {
    if (linkBound && !ordered) return std::min(info->nItems, sm_count() * linkBlocksPerSM);
    if (itemsPerBlock > 0 && !ordered)
      return std::max(1, (info->nItems + itemsPerBlock - 1) / itemsPerBlock);
    return std::min(info->nItems, cap);
  };
  const int anySignal = info->reserved & 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  unsigned *ctr = static_cast<unsigned *>(scratch);
  ++g_launches;
  if (ordered) {
    static const bool roomChecked = [] {
      int n = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, box_copy_sync_kernel<double, kBlocksPerSM / 2>,
                                                        kThreads, 0) != cudaSuccess)
        (void)cudaGetLastError();
      else if (n < kBlocksPerSM)
        std::fprintf(stderr, "tzb: ordered box copy fits only %d blocks per SM (expected %d): kernels meant "
                             "to overlap with it may have to wait\n", n, kBlocksPerSM);
      return true;
    }();
    (void)roomChecked;
    const int blocks = grid(sm_count() * (kBlocksPerSM / 2));
    if (info->elemSize == 8)
      box_copy_sync_kernel<double, kBlocksPerSM / 2><<<blocks, kThreads, 0, s>>>(
          boxes, items, info->nItems, info->nBoxes, anySignal, ep(epoch), ctr);
    else
      box_copy_sync_kernel<float, kBlocksPerSM / 2><<<blocks, kThreads, 0, s>>>(
          boxes, items, info->nItems, info->nBoxes, anySignal, ep(epoch), ctr);
  } else if (waits) {
    // strictly ordered after the local producers (graph edge): full occupancy is safe
    const int blocks = grid(sm_count() * kBlocksPerSM);
    if (info->elemSize == 8)
      box_copy_sync_kernel<double, kBlocksPerSM><<<blocks, kThreads, 0, s>>>(
          boxes, items, info->nItems, info->nBoxes, anySignal, ep(epoch), ctr);
    else
      box_copy_sync_kernel<float, kBlocksPerSM><<<blocks, kThreads, 0, s>>>(
          boxes, items, info->nItems, info->nBoxes, anySignal, ep(epoch), ctr);
  } else if ((info->reserved & 16) && bulk_enabled()) {
    // opt-in bulk-copy path (see box_bulk_kernel); one warp per block, 7 blocks per SM
    const int blocks = grid(sm_count() * 7);
    box_bulk_kernel<<<blocks, 32, 0, s>>>(boxes, items, info->nItems, info->nBoxes, info->elemSize,
                                          anySignal, ep(epoch), ctr);
  } else {
    const int blocks = grid(sm_count() * kBlocksPerSM);
    if (info->elemSize == 8 && vec)
      box_copy_kernel<double, true><<<blocks, kThreads, 0, s>>>(boxes, items, info->nItems,
                                                                info->nBoxes, anySignal, ep(epoch), ctr);
    else if (info->elemSize == 8)
      box_copy_kernel<double, false><<<blocks, kThreads, 0, s>>>(boxes, items, info->nItems,
                                                                 info->nBoxes, anySignal, ep(epoch), ctr);
    else
      box_copy_kernel<float, false><<<blocks, kThreads, 0, s>>>(boxes, items, info->nItems,
                                                                info->nBoxes, anySignal, ep(epoch), ctr);
  }
  return cudaGetLastError();
}

// Host interpreter of a plan image: performs the copies item by item with memcpy.  A checking
// tool for the planner (every element of every box must be covered by exactly one work item);
// flags are neither waited for nor published.  Returns the number of items executed, -1 on a
// malformed image.
long long tzb_boxcopy_host_execute(const void *hostImage, size_t imageBytes) {
  if (!hostImage || imageBytes < sizeof(PlanHeader)) return -1;
  PlanHeader h;
  std::memcpy(&h, hostImage, sizeof(h));
  if (h.nBoxes < 0 || h.nItems < 0 || (h.elemSize != 4 && h.elemSize != 8)) return -1;
  const size_t need = sizeof(PlanHeader) + sizeof(BoxDesc) * size_t(h.nBoxes) + sizeof(WorkItem) * size_t(h.nItems);
  if (imageBytes < need) return -1;
  const char *p = static_cast<const char *>(hostImage) + sizeof(PlanHeader);
  const BoxDesc *boxes = reinterpret_cast<const BoxDesc *>(p);
  const WorkItem *items = reinterpret_cast<const WorkItem *>(p + sizeof(BoxDesc) * size_t(h.nBoxes));
  const int64_t es = h.elemSize;
  for (int it = 0; it < h.nItems; ++it) {
    const WorkItem &w = items[it];
    if (w.box < 0 || w.box >= h.nBoxes) return -1;
    const BoxDesc &b = boxes[w.box];
    if (w.i3 < 0 || w.i3 >= std::max(b.n3, 1) || w.i2 < 0 || w.i2 + w.c2 > b.n2 || w.i1 < 0 ||
        w.i1 + w.c1 > b.n1)
      return -1;
    for (int j2 = 0; j2 < w.c2; ++j2)
      for (int j1 = 0; j1 < w.c1; ++j1) {
        const int64_t so = int64_t(w.i3) * b.s3 + int64_t(w.i2 + j2) * b.s2 + int64_t(w.i1 + j1) * b.s1;
        const int64_t dof = int64_t(w.i3) * b.d3 + int64_t(w.i2 + j2) * b.d2 + int64_t(w.i1 + j1) * b.d1;
        std::memcpy(b.dst + dof * es, b.src + so * es, size_t(b.run) * size_t(es));
      }
  }
  return h.nItems;
}

int tzb_flags_signal(const TzbFlagList *list, uint64_t epoch, void *stream) {
  if (!list || list->n < 0 || list->n > 32) return cudaErrorInvalidValue;
  if (list->n == 0) return cudaSuccess;
  ++g_launches;
  flags_signal_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(*list, ep(epoch));
  return cudaGetLastError();
}

int tzb_flags_wait(const uint64_t *flags, int n, uint64_t epoch, void *stream) {
  if (n <= 0) return cudaSuccess;
  ++g_launches;
  flags_wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(flags, n, ep(epoch));
  return cudaGetLastError();
}

int tzb_set_wait_timeout_ms(unsigned ms) {
  const unsigned long long ns = (unsigned long long)ms * 1000000ull;
  cudaError_t e = set_wait_timeout_ns_this_tu(ns);
  if (e == cudaSuccess) e = set_wait_timeout_spmv(ns);
  return e;
}

int tzb_l2_flush(void *buf, size_t bytes, unsigned value, void *stream) {
  if (!buf || bytes < 32) return cudaErrorInvalidValue;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const size_t n16 = (bytes - 16) / 16; // the last 16 bytes are the read kernel's sink
  const int blocks = sm_count() * 8;
  flush_write_kernel<<<blocks, kThreads, 0, s>>>(static_cast<uint4 *>(buf), n16, value);
  flush_read_kernel<<<blocks, kThreads, 0, s>>>(static_cast<const uint4 *>(buf), n16,
                                                reinterpret_cast<unsigned *>(static_cast<char *>(buf) + n16 * 16));
  return cudaGetLastError();
}

int tzb_delay_us(unsigned us, void *stream) {
  delay_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>((unsigned long long)us * 1000ull);
  return cudaGetLastError();
}

int tzb_copy_bytes(void *dst, const void *src, size_t bytes, void *stream) {
  if (bytes % 16) return cudaErrorInvalidValue;
  if (!bytes) return cudaSuccess;
  ++g_launches;
  copy_bytes_kernel<<<sm_count() * kBlocksPerSM, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<uint4 *>(dst), static_cast<const uint4 *>(src), bytes / 16);
  return cudaGetLastError();
}

} // extern "C"

// ---------------------------------------------------------------------------
// Allocation-free box copy.
//
// tzb_boxcopy_launch() expects the caller to have put the plan image and a
// zeroed counter into device memory it allocated itself.  Some callers cannot
// do that: code running inside another framework's memory pool (where a stray
// cudaMalloc fragments or deadlocks the pool), code on a latency-critical path
// where the implicit device synchronisation of cudaMalloc/cudaFree is not
// acceptable, and plug-ins that are handed buffers and a stream and nothing
// else.  For them the plan image and the counter live in one fixed-size array
// that is part of this library's device image.  Every GPU that loads the
// library carries these TZB_STATIC_SCRATCH_BYTES until the process exits.
// ---------------------------------------------------------------------------
namespace tzb {
namespace {

// layout: [0, 4096) arrival counters (must be zero at kernel start), [4096, ...) plan image
static __device__ __align__(256) unsigned char g_static_scratch[TZB_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
constexpr size_t kStaticCounterBytes = 4096; // arrival counters: up to 1023 boxes

struct StaticScratch {
  std::atomic<unsigned char *> ptr[kMaxDevices];
  StaticScratch() {
    for (auto &p : ptr) p.store(nullptr);
  }
};
StaticScratch g_static;

thread_local char g_lastError[256] = "";

void set_error(const char *fmt, long long a = 0, long long b = 0) {
  std::snprintf(g_lastError, sizeof(g_lastError), fmt, a, b);
}

// Device address of the array on the CURRENT device.  In host code the name
// g_static_scratch is only a stand-in for the symbol, not an address a kernel
// could use, so the address is asked for once per device and remembered.
cudaError_t static_scratch_address(unsigned char **out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  unsigned char *p = g_static.ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_static_scratch);
    if (e != cudaSuccess) return e;
    g_static.ptr[dev].store(p, std::memory_order_release); // same answer from every thread
  }
  *out = p;
  return cudaSuccess;
}

} // namespace
} // namespace tzb

extern "C" {

const char *tzb_last_error(void) { return tzb::g_lastError; }

int tzb_boxcopy_static(const TzbBox *boxes, int nBoxes, int elemSize, uint64_t epoch, void *stream) {
  using namespace tzb;
  g_lastError[0] = 0;
  if (!boxes || nBoxes <= 0) {
    set_error("tzb_boxcopy_static: need at least one box");
    return cudaErrorInvalidValue;
  }
  TzbBoxPlanInfo info;
  int err = tzb_boxcopy_plan(boxes, nBoxes, elemSize, &info, nullptr);
  if (err != cudaSuccess) {
    set_error("tzb_boxcopy_static: invalid box description");
    return err;
  }
  if (info.scratchBytes > kStaticCounterBytes) {
    set_error("tzb_boxcopy_static: %lld boxes need more counters than the built-in array keeps (%lld)",
              (long long)nBoxes, (long long)(kStaticCounterBytes / sizeof(unsigned) - 1));
    return cudaErrorInvalidValue;
  }
  if (info.planBytes + kStaticCounterBytes > size_t(TZB_STATIC_SCRATCH_BYTES)) {
    set_error("tzb_boxcopy_static: this copy needs %lld bytes of plan scratch but the built-in "
              "array holds %lld; split the boxes over several calls or use tzb_boxcopy_launch",
              (long long)info.planBytes, (long long)(TZB_STATIC_SCRATCH_BYTES - kStaticCounterBytes));
    return cudaErrorInvalidValue;
  }
  if (info.nItems == 0) return cudaSuccess;
  unsigned char *base = nullptr;
  cudaError_t ce = static_scratch_address(&base);
  if (ce != cudaSuccess) {
    set_error("tzb_boxcopy_static: cudaGetSymbolAddress failed (%lld)", (long long)ce);
    return ce;
  }
  // host image: host memory only; the device side is the built-in array
  std::vector<char> image(info.planBytes);
  err = tzb_boxcopy_plan(boxes, nBoxes, elemSize, &info, image.data());
  if (err != cudaSuccess) return err;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // the array still holds whatever the previous call left: the counter has to start at zero
  ce = cudaMemsetAsync(base, 0, kStaticCounterBytes, s);
  if (ce != cudaSuccess) return ce;
  // pageable source: the runtime stages it before returning, so `image` may go out of scope
  ce = cudaMemcpyAsync(base + kStaticCounterBytes, image.data(), info.planBytes,
                       cudaMemcpyHostToDevice, s);
  if (ce != cudaSuccess) return ce;
  return tzb_boxcopy_launch(base + kStaticCounterBytes, &info, epoch, base, stream);
}

} // extern "C"
