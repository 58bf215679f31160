#include "tenzing/host_exec.hpp"

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>

#include "tenzing/macro_at.hpp"

namespace tenzing {
namespace host {

namespace {
double &timeout_ref() {
  static double t = [] {
    const char *e = std::getenv("TENZING_HOST_WAIT_SECS");
    return e ? std::atof(e) : 20.0;
  }();
  return t;
}
} // namespace

double wait_timeout_seconds() { return timeout_ref(); }
void set_wait_timeout_seconds(double s) { timeout_ref() = s; }

namespace {
// Fault injection for the protocol checks (scripts/sanitize.sh): TENZING_FAULT=relaxed_flags
// publishes flags WITHOUT release semantics, which leaves the payload unordered with respect to
// the flag - ThreadSanitizer must then report the host-executed exchanges as racy.
bool relaxed_flags() {
  static const bool on = [] {
    const char *e = std::getenv("TENZING_FAULT");
    return e && std::string(e) == "relaxed_flags";
  }();
  return on;
}
} // namespace

void flag_signal(uint64_t *flag, uint64_t epoch) {
  if (relaxed_flags()) return __atomic_store_n(flag, epoch, __ATOMIC_RELAXED);
  __atomic_store_n(flag, epoch, __ATOMIC_RELEASE);
}

void flags_signal(const TzbFlagList &list, uint64_t epoch) {
  for (int i = 0; i < list.n; ++i)
    if (list.flag[i]) flag_signal(list.flag[i], epoch);
}

void flags_wait(const uint64_t *flags, int n, uint64_t epoch) {
  using clk = std::chrono::steady_clock;
  const auto t0 = clk::now();
  for (int i = 0; i < n; ++i) {
    unsigned spins = 0;
    while (__atomic_load_n(flags + i, __ATOMIC_ACQUIRE) < epoch) {
      if (++spins < 64) continue;
      std::this_thread::yield();
      if ((spins & 1023) == 0 &&
          std::chrono::duration<double>(clk::now() - t0).count() > timeout_ref())
        THROW_RUNTIME("flag " << i << " of " << n << " did not reach epoch " << epoch << " within "
                              << timeout_ref() << " s (value "
                              << __atomic_load_n(flags + i, __ATOMIC_ACQUIRE)
                              << "): the schedule dead-locks or a rank is gone");
    }
  }
}

void copy_bytes(void *dst, const void *src, size_t bytes) { std::memcpy(dst, src, bytes); }

void boxcopy(const TzbBox *boxes, int nBoxes, int elemSize, uint64_t epoch, bool ordered) {
  const size_t es = size_t(elemSize);
  for (int b = 0; b < nBoxes; ++b) // permissions first, as the kernel does
    if (boxes[b].ready) flag_signal(boxes[b].ready, epoch);
  for (int b = 0; b < nBoxes; ++b) {
    const TzbBox &bx = boxes[b];
    if (bx.wait) flags_wait(bx.wait, 1, epoch);
    const char *src = static_cast<const char *>(bx.src);
    char *dst = static_cast<char *>(bx.dst);
    const size_t runBytes = size_t(bx.run) * es;
    for (int64_t i3 = 0; i3 < bx.n3; ++i3)
      for (int64_t i2 = 0; i2 < bx.n2; ++i2)
        for (int64_t i1 = 0; i1 < bx.n1; ++i1)
          std::memcpy(dst + (i1 * bx.d1 + i2 * bx.d2 + i3 * bx.d3) * int64_t(es),
                      src + (i1 * bx.s1 + i2 * bx.s2 + i3 * bx.s3) * int64_t(es), runBytes);
    if (ordered && bx.signal) flag_signal(bx.signal, epoch);
  }
  if (!ordered)
    for (int b = 0; b < nBoxes; ++b)
      if (boxes[b].signal) flag_signal(boxes[b].signal, epoch);
}

template <typename T>
void spmv_csr(int rows, const int *rowPtr, const int *colInd, const T *val, const T *x, T *y,
              bool accumulate) {
  for (int r = 0; r < rows; ++r) {
    T acc = 0;
    for (int k = rowPtr[r]; k < rowPtr[r + 1]; ++k) acc += val[k] * x[colInd[k]];
    y[r] = accumulate ? y[r] + acc : acc;
  }
}

template <typename T> void gather(int n, const int *idx, const T *src, T *dst) {
  for (int i = 0; i < n; ++i) dst[i] = src[idx[i]];
}

template <typename T> void vector_add(int n, const T *a, T *y) {
  for (int i = 0; i < n; ++i) y[i] += a[i];
}

template <typename T>
void gather_put(const TzbGatherPutArgs &a, const int *idx, const T *src, uint64_t epoch) {
  for (int p = 0; p < a.nPeers; ++p) {
    T *dst = static_cast<T *>(a.peerDst[p]);
    for (int i = a.displ[p]; i < a.displ[p + 1]; ++i) dst[i - a.displ[p]] = src[idx[i]];
  }
  for (int p = 0; p < a.nPeers; ++p)
    if (a.peerFlag[p]) flag_signal(a.peerFlag[p], epoch);
}

template <typename T>
void spmv_csr_wait(int rows, const int *rowPtr, const int *colInd, const T *val, const T *x, T *y,
                   bool accumulate, const uint64_t *flags, int nFlags, uint64_t epoch) {
  flags_wait(flags, nFlags, epoch);
  spmv_csr(rows, rowPtr, colInd, val, x, y, accumulate);
}

template <typename T>
void spmv_csr_pull(int rows, const int *rowPtr, const int *colInd, const T *val,
                   const T *const *xPtr, T *y, bool accumulate, const uint64_t *ready, int nReady,
                   const TzbFlagList &acks, uint64_t epoch) {
  flags_wait(ready, nReady, epoch);
  for (int r = 0; r < rows; ++r) {
    T acc = 0;
    for (int k = rowPtr[r]; k < rowPtr[r + 1]; ++k) acc += val[k] * *xPtr[colInd[k]];
    y[r] = accumulate ? y[r] + acc : acc;
  }
  flags_signal(acks, epoch);
}

template <typename T>
void spmv_fused_all(const TzbGatherPutArgs &put, const int *sendIdx, int rows, const int *lRowPtr,
                    const int *lColInd, const T *lVal, const int *rRowPtr, const int *rColInd,
                    const T *rVal, const T *x, const T *rx, T *y, const uint64_t *flags, int nFlags,
                    uint64_t epoch) {
  gather_put(put, sendIdx, x, epoch);
  spmv_csr(rows, lRowPtr, lColInd, lVal, x, y, false);
  flags_wait(flags, nFlags, epoch);
  spmv_csr(rows, rRowPtr, rColInd, rVal, rx, y, true);
}

#define TZ_INSTANTIATE(T)                                                                          \
  template void spmv_csr<T>(int, const int *, const int *, const T *, const T *, T *, bool);       \
  template void gather<T>(int, const int *, const T *, T *);                                       \
  template void vector_add<T>(int, const T *, T *);                                                \
  template void gather_put<T>(const TzbGatherPutArgs &, const int *, const T *, uint64_t);         \
  template void spmv_csr_wait<T>(int, const int *, const int *, const T *, const T *, T *, bool,   \
                                 const uint64_t *, int, uint64_t);                                 \
  template void spmv_csr_pull<T>(int, const int *, const int *, const T *, const T *const *, T *,  \
                                 bool, const uint64_t *, int, const TzbFlagList &, uint64_t);      \
  template void spmv_fused_all<T>(const TzbGatherPutArgs &, const int *, int, const int *,         \
                                  const int *, const T *, const int *, const int *, const T *,     \
                                  const T *, const T *, T *, const uint64_t *, int, uint64_t);
TZ_INSTANTIATE(double)
TZ_INSTANTIATE(float)
#undef TZ_INSTANTIATE

} // namespace host
} // namespace tenzing
