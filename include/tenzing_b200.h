/* C ABI of the tenzing-b200 kernel library (libtenzing_b200_kernels.so).
 *
 * Every hot op the schedule search can place is one of these sm_100a kernels.
 * All functions return a cudaError_t value as int (0 = success), launch on the
 * given stream and do not synchronise.  Pointers are device pointers unless
 * stated otherwise; "peer" pointers are device pointers into another GPU's
 * memory (CUDA IPC / peer access over NVLink) and may be used wherever a
 * destination or flag is expected.
 *
 * Cross-GPU protocol: a producer kernel writes its payload, then publishes a
 * 64-bit epoch to a flag word with release semantics at system scope; a
 * consumer kernel spins on its own flag word with acquire semantics until the
 * value is >= the epoch it was launched with.  Epochs only grow.
 */
#ifndef TENZING_B200_H
#define TENZING_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------------------------------------------------------- SpMV ---- */
/* Replaces the reference's cuSPARSE call (include/tenzing/spmv/ops_spmv.cuh:91-126)
 * and its unused scalar kernel (ops_spmv.cuh:25-39).
 * y = A x (accumulate == 0) or y += A x (accumulate != 0); CSR, 32-bit indices. */
int tzb_spmv_csr_f64(int rows, const int *rowPtr, const int *colInd, const double *val,
                     const double *x, double *y, int accumulate, void *stream);
int tzb_spmv_csr_f32(int rows, const int *rowPtr, const int *colInd, const float *val,
                     const float *x, float *y, int accumulate, void *stream);

/* dst[i] = src[idx[i]] - the reference's `scatter` kernel, really a gather
 * (ops_spmv.cuh:41-59). */
int tzb_gather_f64(int n, const int *idx, const double *src, double *dst, void *stream);
int tzb_gather_f32(int n, const int *idx, const float *src, float *dst, void *stream);

/* y[i] += a[i] - the reference's VectorAdd is an empty function
 * (src/spmv/ops_spmv.cu:45-47). */
int tzb_vector_add_f64(int n, const double *a, double *y, void *stream);
int tzb_vector_add_f32(int n, const float *a, float *y, void *stream);

/* Gather fused with the send (replaces Scatter + PostSend,
 * ops_spmv.cuh:167-215,217-240): entry i of the send list goes to
 * peerDst[p][i - displ[p]] where p is the peer whose [displ[p], displ[p+1])
 * range contains i (ranges may be empty).  After the last block finishes,
 * flag[p] <- epoch is published (release, system scope) for every peer whose
 * flag pointer is not NULL.
 * `counter` is one zeroed uint32 of scratch in device memory (left zero). */
typedef struct {
  int nPeers;
  int displ[17];        /* nPeers + 1 entries, displ[0] = 0 */
  void *peerDst[16];    /* base of this rank's slot in each peer's rx buffer */
  uint64_t *peerFlag[16];
} TzbGatherPutArgs;
int tzb_gather_put_f64(const TzbGatherPutArgs *args, const int *idx, const double *src,
                       uint64_t epoch, uint32_t *counter, void *stream);
int tzb_gather_put_f32(const TzbGatherPutArgs *args, const int *idx, const float *src,
                       uint64_t epoch, uint32_t *counter, void *stream);

/* Receive fused with the off-diagonal product (replaces WaitRecv + the remote
 * cuSPARSE call + VectorAdd, ops_spmv.cuh:276-304): every block first waits
 * until all nFlags local flag words are >= epoch, then y (+)= A_remote rx. */
int tzb_spmv_csr_wait_f64(int rows, const int *rowPtr, const int *colInd, const double *val,
                          const double *x, double *y, int accumulate, const uint64_t *flags,
                          int nFlags, uint64_t epoch, void *stream);
int tzb_spmv_csr_wait_f32(int rows, const int *rowPtr, const int *colInd, const float *val,
                          const float *x, float *y, int accumulate, const uint64_t *flags,
                          int nFlags, uint64_t epoch, void *stream);

/* Pull form: the gather of the remote x entries
 * happens inside the product.  xPtr[c] is the address of remote column c in its owner's x (a
 * peer pointer); every block waits until the nReady local `ready` flags reached the epoch (all
 * owners declared their x final), then y (+)= A_remote x; the last block publishes the epoch to
 * the flags in `acks` ("I have finished reading your x").  `counter`: one zeroed uint32. */
struct TzbFlagList_;
int tzb_spmv_csr_pull_f64(int rows, const int *rowPtr, const int *colInd, const double *val,
                          const double *const *xPtr, double *y, int accumulate,
                          const uint64_t *ready, int nReady, const struct TzbFlagList_ *acks,
                          uint64_t epoch, uint32_t *counter, void *stream);
int tzb_spmv_csr_pull_f32(int rows, const int *rowPtr, const int *colInd, const float *val,
                          const float *const *xPtr, float *y, int accumulate,
                          const uint64_t *ready, int nReady, const struct TzbFlagList_ *acks,
                          uint64_t epoch, uint32_t *counter, void *stream);

/* One-to-many put through NVLS multicast (no counterpart in the reference, whose PostSend
 * loops over MPI_Isend): src[0..n) is stored to the MULTICAST address mcDst, so NVSwitch writes
 * it to the same offset of every rank's buffer bound to that multicast object; then the epoch is
 * multicast-stored to mcFlag (the multicast address of this rank's word in every rank's flag
 * array).  `counter`: one zeroed uint32.  The multicast mapping comes from the caller
 * (torch.distributed._symmetric_memory in the Python front end, tenzing_b200/ops/nvls.py). */
int tzb_multicast_put_f64(int n, const double *src, double *mcDst, uint64_t *mcFlag, uint64_t epoch,
                          uint32_t *counter, void *stream);
int tzb_multicast_put_f32(int n, const float *src, float *mcDst, uint64_t *mcFlag, uint64_t epoch,
                          uint32_t *counter, void *stream);

/* The whole distributed product in one launch (no counterpart in the reference, where it is
 * seven ops): gather-put of my x entries to every peer + flag publish, y = A_local x, wait for
 * the nFlags local flags, y += A_remote rx.  `counter`: one zeroed uint32 of scratch. */
int tzb_spmv_fused_all_f64(const TzbGatherPutArgs *put, const int *sendIdx, int rows,
                           const int *lRowPtr, const int *lColInd, const double *lVal,
                           const int *rRowPtr, const int *rColInd, const double *rVal,
                           const double *x, const double *rx, double *y, const uint64_t *flags,
                           int nFlags, uint64_t epoch, uint32_t *counter, void *stream);
int tzb_spmv_fused_all_f32(const TzbGatherPutArgs *put, const int *sendIdx, int rows,
                           const int *lRowPtr, const int *lColInd, const float *lVal,
                           const int *rRowPtr, const int *rColInd, const float *rVal,
                           const float *x, const float *rx, float *y, const uint64_t *flags,
                           int nFlags, uint64_t epoch, uint32_t *counter, void *stream);

/* ----------------------------------------------------------- box copies ---- */
/* One description serves pack, unpack, pack+put and direct put of a halo
 * region in either storage order (replaces pack/unpack kernels
 * src/halo_exchange/ops_halo_exchange.cu:519-664 and the Isend/Irecv/Wait
 * around them).  A box is a 4-level loop nest; the innermost `run` elements
 * are contiguous in both source and destination. Strides are in elements. */
typedef struct {
  const void *src;
  void *dst;                 /* may be a peer pointer */
  int64_t s1, s2, s3;        /* source strides of loop levels 1 (fastest) .. 3 */
  int64_t d1, d2, d3;        /* destination strides */
  int32_t n1, n2, n3, run;   /* extents */
  uint64_t *signal;          /* flag to publish after the box is complete (may be peer), or NULL */
  const uint64_t *wait;      /* local flag to wait on before touching the box, or NULL */
  /* Elements directly before / after every destination run that are padding the copy may
   * overwrite with anything (0 = none).  With them a thin run (an x-face ghost piece is 24 bytes)
   * can be stored as one full 32-byte sector, which spares L2 the read-modify-write of a
   * partial-sector store. */
  int32_t dstSlackLo, dstSlackHi;
  /* non-zero: dst is memory of ANOTHER GPU.  A plan whose payload mostly crosses NVLink is
   * bound by the link, not by the SMs: it is launched with a small resident grid (one block
   * per SM keeps the link full) so that kernels on other streams find free SM slots beside it. */
  int32_t dstRemote, reserved0;
  /* "You may overwrite what I gave you last time": a (peer) flag published at the START of the
   * kernel, before anything is copied, or NULL.  A put that stores straight into memory its
   * partner also reads (ghost cells) pairs it with `wait`: every rank grants its partners
   * permission for this epoch first, then waits for theirs before its first store. */
  uint64_t *ready;
} TzbBox;

/* A plan is the device image of the boxes plus a table that cuts them into
 * block-sized work items.  Build it on the host, copy `planBytes` to the
 * device once, launch many times. */
typedef struct {
  int32_t nBoxes;
  int32_t nItems;
  int32_t elemSize;          /* 4 or 8 */
  int32_t reserved;          /* internal flags set by the plan functions */
  size_t planBytes;          /* size of the host/device image */
  size_t scratchBytes;       /* zeroed device scratch needed at launch (per-box counters) */
  size_t payloadBytes;       /* bytes moved per launch */
} TzbBoxPlanInfo;

/* Pass hostImage == NULL to query sizes only. */
int tzb_boxcopy_plan(const TzbBox *boxes, int nBoxes, int elemSize, TzbBoxPlanInfo *info,
                     void *hostImage);
/* Same, but the boxes are worked through one after the other in the order given (instead of
 * interleaved) and every box's flag is published as soon as that box is complete, so a consumer
 * kernel running at the same time can start on early boxes while later ones are still in flight. */
int tzb_boxcopy_plan_ordered(const TzbBox *boxes, int nBoxes, int elemSize, TzbBoxPlanInfo *info,
                             void *hostImage);
/* devPlan: device copy of the image. scratch: info.scratchBytes of zeroed device
 * memory (the kernel leaves it zeroed again). */
int tzb_boxcopy_launch(const void *devPlan, const TzbBoxPlanInfo *info, uint64_t epoch,
                       void *scratch, void *stream);

/* Box copy for callers that cannot allocate device memory.  (This entry point has
 * no counterpart in the reference.)
 *
 * tzb_boxcopy_launch needs a device copy of the plan and a zeroed counter, which
 * the caller normally gets from cudaMalloc.  That is not possible everywhere:
 * inside another framework's caching allocator a raw cudaMalloc can fragment or
 * stall the pool, on a latency-critical path the implicit synchronisation of
 * cudaMalloc/cudaFree is unwelcome, and a plug-in may simply be given buffers
 * and a stream and no right to allocate.  This variant keeps the plan and the
 * counter in an array of TZB_STATIC_SCRATCH_BYTES that is built into the
 * library's device image (so every GPU that loads the library holds that much
 * memory until the process exits) and calls no cudaMalloc/cudaFree or any other
 * allocating routine.  Source and destination buffers still come from the caller.
 *
 *  - A copy whose plan does not fit is refused with cudaErrorInvalidValue and a
 *    message in tzb_last_error(); the plan needs about 96 bytes per box plus 24
 *    bytes per ~32 KB of payload, so the array covers roughly 300 MB per call.
 *  - The first call on each device looks up the array's device address; make
 *    that call (with that device current) before any stream capture starts.
 *  - `stream` must be a stream of the CURRENT device: the array that is used is
 *    the current device's copy.
 *  - Every call stages its plan (a few KB to 250 KB) from ordinary host memory to
 *    the array with cudaMemcpyAsync; for pageable memory the runtime copies the
 *    bytes before the call returns, so the call costs that much host time.
 *  - All calls share the one array, which keeps the previous call's contents;
 *    the counter is cleared on the stream before each launch.  Two calls whose
 *    kernels may overlap on the GPU (different streams, no ordering between
 *    them) race on it: order them, or use tzb_boxcopy_launch for one of them.
 */
#define TZB_STATIC_SCRATCH_BYTES (256 * 1024)
int tzb_boxcopy_static(const TzbBox *boxes, int nBoxes, int elemSize, uint64_t epoch,
                       void *stream);
/* message of the last failed tzb_boxcopy_static call on this thread ("" if none) */
const char *tzb_last_error(void);

/* Host interpreter of a plan image (the bytes tzb_boxcopy_plan wrote): executes the work items
 * one by one with memcpy on HOST pointers; flags are ignored.  A checking tool for the planner -
 * the result must equal the plain loop nest over the boxes.  Returns the number of items
 * executed, -1 for a malformed image. */
long long tzb_boxcopy_host_execute(const void *hostImage, size_t imageBytes);

/* ----------------------------------------------------------------- flags ---- */
/* publish epoch to up to 32 (possibly peer) flags; wait for n contiguous local flags. */
typedef struct TzbFlagList_ {
  int n;
  uint64_t *flag[32];
} TzbFlagList;
int tzb_flags_signal(const TzbFlagList *list, uint64_t epoch, void *stream);
int tzb_flags_wait(const uint64_t *flags, int n, uint64_t epoch, void *stream);
/* Watchdog of every device-side flag wait of this library (the reference's MPI_Wait has none
 * and hangs with a dead rank): a wait that stays unsatisfied for `ms` milliseconds
 * ends its kernel with a trap, so the launch fails with cudaErrorLaunchFailed instead of
 * spinning until the job is killed.  Default 20000 ms; 0 = wait for ever.  Applies to the
 * current device; call it before launching (it is a synchronous copy to a device symbol). */
int tzb_set_wait_timeout_ms(unsigned ms);

/* ---------------------------------------------------- CUDA-graph replay ---- */
/* Every kernel above takes its epoch by value.  A launch sequence captured into a CUDA graph is
 * replayed with the same arguments, so for capture the epoch can come from a device word
 * instead: while an epoch source is set (for the calling thread; set it around the capture only), every
 * launch of this library hands the word's address to its kernel, which reads the epoch from it
 * and ignores the by-value argument.  tzb_epoch_advance enqueues `*devWord += 1` - the first node
 * of such a graph.  (The reference has no counterpart: its benchmark loop re-issues every op
 * from the host, src/benchmarker.cpp:93-97.) */
void tzb_set_epoch_source(const uint64_t *devWord); /* per calling thread; NULL = by value again */
int tzb_epoch_advance(uint64_t *devWord, void *stream);
int tzb_epoch_set(uint64_t *devWord, uint64_t value, void *stream);

/* ------------------------------------------------------------ diagnostics -- */
/* L2 flush between timed steps: writes `bytes` (> L2 size) of `buf` with `value`, then reads
 * them back, so L2 ends up full of CLEAN lines of the flush buffer and the timed kernel does
 * not pay for writing the flush's dirty lines back.  The last 16 bytes of buf are scratch. */
int tzb_l2_flush(void *buf, size_t bytes, unsigned value, void *stream);
/* occupies the stream for `us` microseconds (one idle thread): a measurement harness enqueues it
 * first so that everything behind it is already queued when the timed region starts */
int tzb_delay_us(unsigned us, void *stream);
/* plain device copy used as the local bandwidth yardstick in the benchmarks */
int tzb_copy_bytes(void *dst, const void *src, size_t bytes, void *stream);
/* number of kernels launched through this library since load (for bench.py's launch count) */
uint64_t tzb_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* TENZING_B200_H */
