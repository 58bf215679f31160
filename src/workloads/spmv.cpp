#include "tenzing/spmv/row_part_spmv.hpp"
#include "tenzing/flag_synchronizer.hpp"
#include "tenzing/host_exec.hpp"

#include <cstring>
#include <set>

#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/halo/halo_exchange.hpp" // CudaFreeDeleter
#include "tenzing_b200.h"

namespace tzb {
// C++ kernel entries with an explicit lanes-per-row hint (src/cuda/spmv_kernels.cu)
cudaError_t spmv_csr(int, const int *, const int *, const double *, const double *, double *, bool,
                     int, cudaStream_t);
cudaError_t spmv_csr(int, const int *, const int *, const float *, const float *, float *, bool,
                     int, cudaStream_t);
cudaError_t spmv_csr_wait(int, const int *, const int *, const double *, const double *, double *,
                          bool, int, const uint64_t *, int, uint64_t, cudaStream_t);
cudaError_t spmv_csr_wait(int, const int *, const int *, const float *, const float *, float *,
                          bool, int, const uint64_t *, int, uint64_t, cudaStream_t);
cudaError_t spmv_csr_pull(int, const int *, const int *, const double *, const double *const *,
                          double *, bool, const uint64_t *, int, const TzbFlagList *, uint64_t,
                          uint32_t *, int, cudaStream_t);
cudaError_t spmv_csr_pull(int, const int *, const int *, const float *, const float *const *, float *,
                          bool, const uint64_t *, int, const TzbFlagList *, uint64_t, uint32_t *, int,
                          cudaStream_t);
cudaError_t spmv_fused_all(const TzbGatherPutArgs *, const int *, int, const int *, const int *,
                           const double *, const int *, const int *, const double *, const double *,
                           const double *, double *, const uint64_t *, int, uint64_t, uint32_t *, int,
                           cudaStream_t);
cudaError_t spmv_fused_all(const TzbGatherPutArgs *, const int *, int, const int *, const int *,
                           const float *, const int *, const int *, const float *, const float *,
                           const float *, float *, const uint64_t *, int, uint64_t, uint32_t *, int,
                           cudaStream_t);
} // namespace tzb

namespace tenzing {

namespace {

std::shared_ptr<void> to_device(const void *host, size_t bytes, bool null) {
  if (null) {
    void *p = std::malloc(bytes ? bytes : 1);
    if (bytes) std::memcpy(p, host, bytes);
    return std::shared_ptr<void>(p, [](void *q) { std::free(q); });
  }
  void *p = nullptr;
  CUDA_RUNTIME(cudaMalloc(&p, bytes ? bytes : 16));
  if (bytes) CUDA_RUNTIME(cudaMemcpy(p, host, bytes, cudaMemcpyHostToDevice));
  return std::shared_ptr<void>(p, CudaFreeDeleter());
}

template <typename Scalar>
DeviceCsr<Scalar> upload(const CsrMat<int, Scalar> &m, bool null) {
  DeviceCsr<Scalar> d;
  d.rows = int(m.num_rows());
  d.cols = int(m.num_cols());
  d.nnz = m.nnz();
  d.rowPtr = to_device(m.rowPtr_.data(), m.rowPtr_.size() * sizeof(int), null);
  d.colInd = to_device(m.colInd_.data(), m.colInd_.size() * sizeof(int), null);
  d.val = to_device(m.val_.data(), m.val_.size() * sizeof(Scalar), null);
  return d;
}

size_t up256(size_t v) { return (v + 255) / 256 * 256; }

} // namespace

template <typename Scalar>
RowPartSpmv<Scalar>::RowPartSpmv(std::shared_ptr<Communicator> comm,
                                 const CsrMat<int, Scalar> &myRows, int64_t n,
                                 SymmetricHeap::Memory memory)
    : n_(n) {
  if (!comm) comm = std::make_shared<SelfComm>();
  const int me = comm->rank(), np = comm->size();
  const bool null = memory == SymmetricHeap::Memory::Host;
  range_ = get_partition(n, me, np);
  if (myRows.num_cols() != n)
    THROW_RUNTIME("row-partitioned SpMV needs a square matrix: " << n << " rows, "
                                                                  << myRows.num_cols() << " columns");
  if (myRows.num_rows() != range_.extent())
    THROW_RUNTIME("rank " << me << " holds " << myRows.num_rows() << " rows, expected "
                          << range_.extent());
  split_ = split_local_remote(myRows, range_.lb, range_.ub);
  lA_ = upload(split_.local, null);
  rA_ = upload(split_.remote, null);

  // what I need, grouped by owner (globals are ascending, so groups are contiguous)
  const std::vector<int> &need = split_.globals;
  {
    size_t i = 0;
    while (i < need.size()) {
      const int owner = get_owner(n, need[i], np);
      const Range orange = get_partition(n, owner, np);
      size_t j = i;
      while (j < need.size() && need[j] < orange.ub) ++j;
      Param p;
      p.peer = owner, p.displ = int(i), p.count = int(j - i);
      recvParams_.push_back(p);
      i = j;
    }
  }

  // everybody learns everybody's need list (set-up only; the reference does an
  // all-pairs Isend/Probe/Recv here, row_part_spmv.cuh:321-403)
  std::string mine(reinterpret_cast<const char *>(need.data()), need.size() * sizeof(int));
  const std::vector<std::string> all = comm->allgather_string(mine);
  std::vector<int> sendIdx;
  for (int r = 0; r < np; ++r) {
    if (r == me) continue;
    const int *theirs = reinterpret_cast<const int *>(all[size_t(r)].data());
    const size_t cnt = all[size_t(r)].size() / sizeof(int);
    const int *lo = std::lower_bound(theirs, theirs + cnt, int(range_.lb));
    const int *hi = std::lower_bound(theirs, theirs + cnt, int(range_.ub));
    if (lo == hi) continue;
    Param p;
    p.peer = r, p.displ = int(sendIdx.size()), p.count = int(hi - lo), p.dstDispl = int(lo - theirs);
    for (const int *q = lo; q != hi; ++q) sendIdx.push_back(*q - int(range_.lb));
    sendParams_.push_back(p);
  }
  numSend_ = int(sendIdx.size());
  xSendIdx_ = to_device(sendIdx.data(), sendIdx.size() * sizeof(int), null);
  const uint32_t zero = 0;
  counter_ = to_device(&zero, sizeof(zero), null);

  std::set<int> ps;
  for (const auto &p : sendParams_) ps.insert(p.peer);
  for (const auto &p : recvParams_) ps.insert(p.peer);
  peers_.assign(ps.begin(), ps.end());
  if (peers_.size() > 16) THROW_RUNTIME("more than 16 SpMV peers are not supported");

  // symmetric allocation sized by the largest rank
  double sizes[3] = {double(range_.extent()), double(need.size()), double(numSend_)};
  comm->allreduce_max(sizes, 3);
  const size_t maxRows = size_t(sizes[0]), maxRemote = size_t(sizes[1]), maxSend = size_t(sizes[2]);
  const size_t es = sizeof(Scalar);
  const size_t flagBytes = sizeof(uint64_t) * size_t(kFlagSets) * size_t(np); // one set per style + acks
  const size_t total = 3 * up256(maxRows * es) + 2 * up256(maxRemote * es) + up256(maxSend * es) +
                       up256(flagBytes) + 4096;
  heap_.reset(new SymmetricHeap(comm, total, memory));
  lxOff_ = heap_->alloc(maxRows * es);
  lyOff_ = heap_->alloc(maxRows * es);
  ryOff_ = heap_->alloc(maxRows * es);
  rxOff_[0] = heap_->alloc(maxRemote * es);
  rxOff_[1] = heap_->alloc(maxRemote * es);
  sendOff_ = heap_->alloc(maxSend * es);
  flagOff_ = heap_->alloc(flagBytes);

  // pull style: where every remote column lives in its owner's x (peer pointers; a rank in
  // another process whose host heap could not be mapped leaves null entries - dry runs only)
  std::vector<const Scalar *> ptrs(need.size(), nullptr);
  for (size_t c = 0; c < need.size(); ++c) {
    const int owner = get_owner(n, need[c], np);
    if (!heap_->peer(owner)) continue;
    ptrs[c] = lx(owner) + (need[c] - get_partition(n, owner, np).lb);
  }
  xPtrs_ = to_device(ptrs.data(), ptrs.size() * sizeof(const Scalar *), null);
}

template <typename Scalar> size_t RowPartSpmv<Scalar>::hbm_bytes() const {
  const size_t es = sizeof(Scalar);
  const size_t nnz = size_t(lA_.nnz + rA_.nnz), rows = size_t(lA_.rows);
  // val + colInd per nnz, two rowPtr streams, x read once (local + remote), y written/updated
  return nnz * (es + 4) + 2 * (rows + 1) * 4 + (rows + size_t(num_remote_cols())) * es + 2 * rows * es;
}

template <typename Scalar>
CsrMat<int, Scalar> distribute_rows(Communicator &comm, const CsrMat<int, Scalar> *whole, int root) {
  const int me = comm.rank(), np = comm.size();
  int64_t dims[2] = {0, 0};
  if (me == root) {
    if (!whole) THROW_RUNTIME("distribute_rows: the root must pass the matrix");
    dims[0] = whole->num_rows(), dims[1] = whole->num_cols();
  }
  comm.bcast(dims, sizeof(dims), root);
  CsrMat<int, Scalar> mine;
  for (int r = 0; r < np; ++r) {
    const Range rg = get_partition(dims[0], r, np);
    // root serialises block r: [nnz][rowPtr...][colInd...][val...]
    std::string buf;
    if (me == root) {
      const CsrMat<int, Scalar> blk = whole->retain_rows(rg.lb, rg.ub);
      const int64_t nnz = blk.nnz();
      buf.append(reinterpret_cast<const char *>(&nnz), sizeof(nnz));
      buf.append(reinterpret_cast<const char *>(blk.rowPtr_.data()), blk.rowPtr_.size() * sizeof(int));
      buf.append(reinterpret_cast<const char *>(blk.colInd_.data()), blk.colInd_.size() * sizeof(int));
      buf.append(reinterpret_cast<const char *>(blk.val_.data()), blk.val_.size() * sizeof(Scalar));
    }
    buf = comm.bcast_string(buf, root);
    if (r != me) continue;
    const char *p = buf.data();
    int64_t nnz = 0;
    std::memcpy(&nnz, p, sizeof(nnz));
    p += sizeof(nnz);
    mine = CsrMat<int, Scalar>(rg.extent(), dims[1]);
    std::memcpy(mine.rowPtr_.data(), p, mine.rowPtr_.size() * sizeof(int));
    p += mine.rowPtr_.size() * sizeof(int);
    mine.colInd_.resize(size_t(nnz));
    mine.val_.resize(size_t(nnz));
    std::memcpy(mine.colInd_.data(), p, size_t(nnz) * sizeof(int));
    p += size_t(nnz) * sizeof(int);
    std::memcpy(mine.val_.data(), p, size_t(nnz) * sizeof(Scalar));
  }
  return mine;
}
template CsrMat<int, double> distribute_rows<double>(Communicator &, const CsrMat<int, double> *, int);
template CsrMat<int, float> distribute_rows<float>(Communicator &, const CsrMat<int, float> *, int);

template <typename Scalar>
std::shared_ptr<RowPartSpmv<Scalar>> make_band_spmv(std::shared_ptr<Communicator> comm, int64_t n,
                                                    int64_t bw, int nnzPerRow, uint64_t seed,
                                                    SymmetricHeap::Memory memory) {
  if (!comm) comm = std::make_shared<SelfComm>();
  const Range r = get_partition(n, comm->rank(), comm->size());
  auto rows = random_band_rows<int, Scalar>(n, bw, nnzPerRow, r.lb, r.ub, seed);
  return std::make_shared<RowPartSpmv<Scalar>>(comm, rows, n, memory);
}
template std::shared_ptr<RowPartSpmv<double>> make_band_spmv<double>(std::shared_ptr<Communicator>,
                                                                     int64_t, int64_t, int, uint64_t,
                                                                     SymmetricHeap::Memory);
template std::shared_ptr<RowPartSpmv<float>> make_band_spmv<float>(std::shared_ptr<Communicator>,
                                                                   int64_t, int64_t, int, uint64_t,
                                                                   SymmetricHeap::Memory);

// ---------------------------------------------------------------- ops
namespace {

template <typename Scalar> using SpmvPtr = std::shared_ptr<RowPartSpmv<Scalar>>;
template <typename Scalar> using St = typename RowPartSpmv<Scalar>::Style;

// Flag channels of one flag set, for the FlagSynchronizer: the heap offsets of the words of ALL
// source ranks (rank s publishes word s on every rank; the same op on the other ranks is what
// fills the remaining words here).
template <typename Scalar> std::vector<uint64_t> set_channels(const RowPartSpmv<Scalar> &s, int set) {
  std::vector<uint64_t> out;
  const uint64_t *base = set == RowPartSpmv<Scalar>::kFlagSets - 1
                             ? s.ack_flag(s.rank(), 0)
                             : s.flag(s.rank(), St<Scalar>(set), 0);
  for (int r = 0; r < s.size(); ++r) out.push_back(uint64_t(s.heap().offset_of(base + r)));
  return out;
}
template <typename Scalar> constexpr int ack_set() { return RowPartSpmv<Scalar>::kFlagSets - 1; }

cudaError_t k_gather(int n, const int *idx, const double *s, double *d, cudaStream_t st) {
  return cudaError_t(tzb_gather_f64(n, idx, s, d, st));
}
cudaError_t k_gather(int n, const int *idx, const float *s, float *d, cudaStream_t st) {
  return cudaError_t(tzb_gather_f32(n, idx, s, d, st));
}
cudaError_t k_add(int n, const double *a, double *y, cudaStream_t st) {
  return cudaError_t(tzb_vector_add_f64(n, a, y, st));
}
cudaError_t k_add(int n, const float *a, float *y, cudaStream_t st) {
  return cudaError_t(tzb_vector_add_f32(n, a, y, st));
}
cudaError_t k_gather_put(const TzbGatherPutArgs *a, const int *idx, const double *s, uint64_t e,
                         uint32_t *c, cudaStream_t st) {
  return cudaError_t(tzb_gather_put_f64(a, idx, s, e, c, st));
}
cudaError_t k_gather_put(const TzbGatherPutArgs *a, const int *idx, const float *s, uint64_t e,
                         uint32_t *c, cudaStream_t st) {
  return cudaError_t(tzb_gather_put_f32(a, idx, s, e, c, st));
}

// ly = A_local * lx
template <typename Scalar> class LocalSpMV : public GpuOp {
  SpmvPtr<Scalar> s_;

public:
  explicit LocalSpMV(SpmvPtr<Scalar> s) : s_(std::move(s)) {}
  std::string name() const override { return "yl"; }
  CLONE_DEF(LocalSpMV)
  void run(Platform &plat, cudaStream_t st) override {
    if (s_->null_device() && plat.host_execution()) {
      const auto &A = s_->lA();
      host::spmv_csr(A.rows, A.row_ptr(), A.col_ind(), A.values(), s_->lx(), s_->ly(), false);
      return;
    }
    run(st);
  }
  void run(cudaStream_t st) override {
    if (s_->null_device()) return;
    const auto &A = s_->lA();
    CUDA_RUNTIME(tzb::spmv_csr(A.rows, A.row_ptr(), A.col_ind(), A.values(), s_->lx(), s_->ly(),
                               false, A.nnz_per_row(), st));
  }
};

// remote product A_remote * rx.  intoLy: accumulate into ly (yl must precede),
// else write ry.  wait: fuse the wait for every rank's flag of this epoch.
template <typename Scalar> class RemoteSpMV : public GpuOp, public HasFlags {
  SpmvPtr<Scalar> s_;
  bool intoLy_, wait_;
  St<Scalar> style_;

public:
  std::vector<uint64_t> flags_published() const override { return {}; }
  std::vector<uint64_t> flags_awaited() const override {
    return wait_ ? set_channels(*s_, int(style_)) : std::vector<uint64_t>{};
  }
  RemoteSpMV(SpmvPtr<Scalar> s, bool intoLy, bool wait, St<Scalar> style)
      : s_(std::move(s)), intoLy_(intoLy), wait_(wait), style_(style) {}
  std::string name() const override { return "yr"; }
  CLONE_DEF(RemoteSpMV)
  void run(cudaStream_t) override { THROW_RUNTIME("yr needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t st) override {
    const auto &A = s_->rA();
    const Scalar *x = s_->rx(s_->rank(), int(plat.epoch() & 1));
    Scalar *y = intoLy_ ? s_->ly() : s_->ry();
    if (s_->null_device() && plat.host_execution()) {
      if (wait_)
        host::spmv_csr_wait(A.rows, A.row_ptr(), A.col_ind(), A.values(), x, y, intoLy_,
                            s_->flag(s_->rank(), style_, 0), s_->size(), plat.epoch());
      else
        host::spmv_csr(A.rows, A.row_ptr(), A.col_ind(), A.values(), x, y, intoLy_);
      return;
    }
    if (s_->null_device() || plat.is_null()) return;
    if (wait_) {
      CUDA_RUNTIME(tzb::spmv_csr_wait(A.rows, A.row_ptr(), A.col_ind(), A.values(), x, y, intoLy_,
                                      A.nnz_per_row(), s_->flag(s_->rank(), style_, 0), s_->size(),
                                      plat.epoch(), st));
    } else {
      CUDA_RUNTIME(tzb::spmv_csr(A.rows, A.row_ptr(), A.col_ind(), A.values(), x, y, intoLy_,
                                 A.nnz_per_row(), st));
    }
  }
};

// the reference's Scatter, named "Pack": xSendBuf[i] = lx[xSendIdx[i]]
template <typename Scalar> class Scatter : public GpuOp {
  SpmvPtr<Scalar> s_;

public:
  explicit Scatter(SpmvPtr<Scalar> s) : s_(std::move(s)) {}
  std::string name() const override { return "Pack"; }
  CLONE_DEF(Scatter)
  void run(Platform &plat, cudaStream_t st) override {
    if (s_->null_device() && plat.host_execution()) {
      host::gather(s_->num_send(), s_->x_send_idx(), s_->lx(), s_->x_send_buf());
      return;
    }
    run(st);
  }
  void run(cudaStream_t st) override {
    if (s_->null_device()) return;
    CUDA_RUNTIME(k_gather(s_->num_send(), s_->x_send_idx(), s_->lx(), s_->x_send_buf(), st));
  }
};

// ly += ry
template <typename Scalar> class VectorAdd : public GpuOp {
  SpmvPtr<Scalar> s_;

public:
  explicit VectorAdd(SpmvPtr<Scalar> s) : s_(std::move(s)) {}
  std::string name() const override { return "y"; }
  CLONE_DEF(VectorAdd)
  void run(Platform &plat, cudaStream_t st) override {
    if (s_->null_device() && plat.host_execution()) {
      host::vector_add(int(s_->row_range().extent()), s_->ry(), s_->ly());
      return;
    }
    run(st);
  }
  void run(cudaStream_t st) override {
    if (s_->null_device()) return;
    CUDA_RUNTIME(k_add(int(s_->row_range().extent()), s_->ry(), s_->ly(), st));
  }
};

template <typename Scalar>
TzbGatherPutArgs make_put_args(const RowPartSpmv<Scalar> &s, St<Scalar> style, int parity);

// gather fused with the put into every peer's rx, then one flag per rank
template <typename Scalar> class PackPut : public GpuOp, public HasFlags {
  SpmvPtr<Scalar> s_;
  St<Scalar> style_;

public:
  std::vector<uint64_t> flags_published() const override { return set_channels(*s_, int(style_)); }
  std::vector<uint64_t> flags_awaited() const override { return {}; }
  PackPut(SpmvPtr<Scalar> s, St<Scalar> style) : s_(std::move(s)), style_(style) {}
  std::string name() const override { return "PackPut"; }
  CLONE_DEF(PackPut)
  void run(cudaStream_t) override { THROW_RUNTIME("PackPut needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t st) override {
    if (s_->null_device() && plat.host_execution()) {
      const TzbGatherPutArgs a = make_put_args(*s_, style_, int(plat.epoch() & 1));
      host::gather_put(a, s_->x_send_idx(), s_->lx(), plat.epoch());
      return;
    }
    if (s_->null_device() || plat.is_null()) return;
    const TzbGatherPutArgs a = make_put_args(*s_, style_, int(plat.epoch() & 1));
    CUDA_RUNTIME(k_gather_put(&a, s_->x_send_idx(), s_->lx(), plat.epoch(), s_->counter(), st));
  }
};

// one slot per rank, in rank order; ranks that get no data still get the flag so the handshake
// is all-to-all and nobody can run more than one epoch ahead of anybody
template <typename Scalar>
TzbGatherPutArgs make_put_args(const RowPartSpmv<Scalar> &s, St<Scalar> style, int parity) {
  const int np = s.size(), me = s.rank();
  TzbGatherPutArgs a;
  std::memset(&a, 0, sizeof(a));
  a.nPeers = np;
  size_t k = 0;
  const auto &sp = s.send_params();
  int displ = 0;
  for (int r = 0; r < np; ++r) {
    a.displ[r] = displ;
    a.peerFlag[r] = s.flag(r, style, me);
    if (k < sp.size() && sp[k].peer == r) {
      a.peerDst[r] = s.rx(r, parity) + sp[k].dstDispl;
      displ += sp[k].count;
      ++k;
    }
  }
  a.displ[np] = displ;
  return a;
}

// the whole product as one kernel launch (put, local product, flag wait, remote product)
template <typename Scalar> class FusedAllSpMV : public GpuOp, public HasFlags {
  SpmvPtr<Scalar> s_;

public:
  std::vector<uint64_t> flags_published() const override {
    return set_channels(*s_, int(St<Scalar>::FusedAll));
  }
  std::vector<uint64_t> flags_awaited() const override { return flags_published(); }
  bool publishes_before_waiting() const override { return true; }
  explicit FusedAllSpMV(SpmvPtr<Scalar> s) : s_(std::move(s)) {}
  std::string name() const override { return "SpMVAll"; }
  CLONE_DEF(FusedAllSpMV)
  void run(cudaStream_t) override { THROW_RUNTIME("SpMVAll needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t st) override {
    const bool hostExec = s_->null_device() && plat.host_execution();
    if (!hostExec && (s_->null_device() || plat.is_null())) return;
    const int parity = int(plat.epoch() & 1);
    const auto style = St<Scalar>::FusedAll;
    const TzbGatherPutArgs a = make_put_args(*s_, style, parity);
    const auto &L = s_->lA();
    const auto &R = s_->rA();
    if (hostExec) {
      host::spmv_fused_all(a, s_->x_send_idx(), L.rows, L.row_ptr(), L.col_ind(), L.values(),
                           R.row_ptr(), R.col_ind(), R.values(), s_->lx(),
                           static_cast<const Scalar *>(s_->rx(s_->rank(), parity)), s_->ly(),
                           s_->flag(s_->rank(), style, 0), s_->size(), plat.epoch());
      return;
    }
    CUDA_RUNTIME(tzb::spmv_fused_all(&a, s_->x_send_idx(), L.rows, L.row_ptr(), L.col_ind(),
                                     L.values(), R.row_ptr(), R.col_ind(), R.values(), s_->lx(),
                                     s_->rx(s_->rank(), parity), s_->ly(),
                                     s_->flag(s_->rank(), style, 0), s_->size(), plat.epoch(),
                                     s_->counter(), std::max(L.nnz_per_row(), 2), st));
  }
};

// ---- pull style (built on request, not a default alternative) ----
cudaError_t k_pull(int rows, const int *rp, const int *ci, const double *v, const double *const *xp,
                   double *y, bool acc, const uint64_t *ready, int nReady, const TzbFlagList *acks,
                   uint64_t e, uint32_t *c, int nnz, cudaStream_t st) {
  return tzb::spmv_csr_pull(rows, rp, ci, v, xp, y, acc, ready, nReady, acks, e, c, nnz, st);
}
cudaError_t k_pull(int rows, const int *rp, const int *ci, const float *v, const float *const *xp,
                   float *y, bool acc, const uint64_t *ready, int nReady, const TzbFlagList *acks,
                   uint64_t e, uint32_t *c, int nnz, cudaStream_t st) {
  return tzb::spmv_csr_pull(rows, rp, ci, v, xp, y, acc, ready, nReady, acks, e, c, nnz, st);
}

// "my x is final for this epoch": one flag on every rank
template <typename Scalar> class XReady : public GpuOp, public HasFlags {
  SpmvPtr<Scalar> s_;

public:
  std::vector<uint64_t> flags_published() const override {
    return set_channels(*s_, int(St<Scalar>::FusedPull));
  }
  std::vector<uint64_t> flags_awaited() const override { return {}; }
  explicit XReady(SpmvPtr<Scalar> s) : s_(std::move(s)) {}
  std::string name() const override { return "XReady"; }
  CLONE_DEF(XReady)
  void run(cudaStream_t) override { THROW_RUNTIME("XReady needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t st) override {
    TzbFlagList l;
    l.n = 0;
    for (int r = 0; r < s_->size(); ++r) l.flag[l.n++] = s_->flag(r, St<Scalar>::FusedPull, s_->rank());
    if (s_->null_device() && plat.host_execution()) return host::flags_signal(l, plat.epoch());
    if (s_->null_device() || plat.is_null()) return;
    CUDA_RUNTIME(cudaError_t(tzb_flags_signal(&l, plat.epoch(), st)));
  }
};

// ly += A_remote * x read from the owners, then "done reading" to every rank
template <typename Scalar> class PullSpMV : public GpuOp, public HasFlags {
  SpmvPtr<Scalar> s_;

public:
  std::vector<uint64_t> flags_published() const override { return set_channels(*s_, ack_set<Scalar>()); }
  std::vector<uint64_t> flags_awaited() const override {
    return set_channels(*s_, int(St<Scalar>::FusedPull));
  }
  explicit PullSpMV(SpmvPtr<Scalar> s) : s_(std::move(s)) {}
  std::string name() const override { return "yrPull"; }
  CLONE_DEF(PullSpMV)
  void run(cudaStream_t) override { THROW_RUNTIME("yrPull needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t st) override {
    const bool hostExec = s_->null_device() && plat.host_execution();
    if (!hostExec && (s_->null_device() || plat.is_null())) return;
    const auto &A = s_->rA();
    TzbFlagList acks;
    acks.n = 0;
    for (int r = 0; r < s_->size(); ++r) acks.flag[acks.n++] = s_->ack_flag(r, s_->rank());
    const uint64_t *ready = s_->flag(s_->rank(), St<Scalar>::FusedPull, 0);
    if (hostExec)
      return host::spmv_csr_pull(A.rows, A.row_ptr(), A.col_ind(), A.values(), s_->x_ptrs(), s_->ly(),
                                 true, ready, s_->size(), acks, plat.epoch());
    CUDA_RUNTIME(k_pull(A.rows, A.row_ptr(), A.col_ind(), A.values(), s_->x_ptrs(), s_->ly(), true,
                        ready, s_->size(), &acks, plat.epoch(), s_->counter(), A.nnz_per_row(), st));
  }
};

// every rank has finished reading my x: it may be overwritten from here on
template <typename Scalar> class XFree : public GpuOp, public HasFlags {
  SpmvPtr<Scalar> s_;

public:
  std::vector<uint64_t> flags_published() const override { return {}; }
  std::vector<uint64_t> flags_awaited() const override { return set_channels(*s_, ack_set<Scalar>()); }
  explicit XFree(SpmvPtr<Scalar> s) : s_(std::move(s)) {}
  std::string name() const override { return "XFree"; }
  CLONE_DEF(XFree)
  void run(cudaStream_t) override { THROW_RUNTIME("XFree needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t st) override {
    const uint64_t *acks = s_->ack_flag(s_->rank(), 0);
    if (s_->null_device() && plat.host_execution()) return host::flags_wait(acks, s_->size(), plat.epoch());
    if (s_->null_device() || plat.is_null()) return;
    CUDA_RUNTIME(cudaError_t(tzb_flags_wait(acks, s_->size(), plat.epoch(), st)));
  }
};

// host-driven transport of the reference structure (copy engine + flags)
struct SpmvTransport {
  cudaStream_t send = nullptr, recv = nullptr;
  bool null;
  explicit SpmvTransport(bool n) : null(n) {
    if (null) return;
    CUDA_RUNTIME(cudaStreamCreateWithFlags(&send, cudaStreamNonBlocking));
    CUDA_RUNTIME(cudaStreamCreateWithFlags(&recv, cudaStreamNonBlocking));
  }
  ~SpmvTransport() {
    if (send) cudaStreamDestroy(send);
    if (recv) cudaStreamDestroy(recv);
  }
};

template <typename Scalar> class PostSend : public CpuOp, public HasFlags {
  SpmvPtr<Scalar> s_;
  std::shared_ptr<SpmvTransport> tr_;

public:
  std::vector<uint64_t> flags_published() const override {
    return set_channels(*s_, int(St<Scalar>::Reference));
  }
  std::vector<uint64_t> flags_awaited() const override { return {}; }
  PostSend(SpmvPtr<Scalar> s, std::shared_ptr<SpmvTransport> tr) : s_(std::move(s)), tr_(std::move(tr)) {}
  std::string name() const override { return "PostSend"; }
  CLONE_DEF(PostSend)
  void run(Platform &plat) override {
    const int parity = int(plat.epoch() & 1), me = s_->rank();
    if (tr_->null && plat.host_execution()) {
      for (const auto &p : s_->send_params())
        host::copy_bytes(s_->rx(p.peer, parity) + p.dstDispl, s_->x_send_buf() + p.displ,
                         size_t(p.count) * sizeof(Scalar));
      for (int r = 0; r < s_->size(); ++r)
        host::flag_signal(s_->flag(r, St<Scalar>::Reference, me), plat.epoch());
      return;
    }
    if (tr_->null || plat.is_null()) return;
    for (const auto &p : s_->send_params())
      CUDA_RUNTIME(cudaMemcpyAsync(s_->rx(p.peer, parity) + p.dstDispl, s_->x_send_buf() + p.displ,
                                   size_t(p.count) * sizeof(Scalar), cudaMemcpyDeviceToDevice,
                                   tr_->send));
    TzbFlagList l;
    l.n = 0;
    for (int r = 0; r < s_->size(); ++r) l.flag[l.n++] = s_->flag(r, St<Scalar>::Reference, me);
    CUDA_RUNTIME(cudaError_t(tzb_flags_signal(&l, plat.epoch(), tr_->send)));
  }
};

template <typename Scalar> class WaitSend : public CpuOp {
  std::shared_ptr<SpmvTransport> tr_;

public:
  explicit WaitSend(std::shared_ptr<SpmvTransport> tr) : tr_(std::move(tr)) {}
  std::string name() const override { return "WaitSend"; }
  CLONE_DEF(WaitSend)
  void run(Platform &plat) override {
    if (tr_->null || plat.is_null()) return;
    CUDA_RUNTIME(cudaStreamSynchronize(tr_->send));
  }
};

template <typename Scalar> class WaitRecv : public CpuOp, public HasFlags {
  SpmvPtr<Scalar> s_;
  std::shared_ptr<SpmvTransport> tr_;

public:
  std::vector<uint64_t> flags_published() const override { return {}; }
  std::vector<uint64_t> flags_awaited() const override {
    return set_channels(*s_, int(St<Scalar>::Reference));
  }
  WaitRecv(SpmvPtr<Scalar> s, std::shared_ptr<SpmvTransport> tr) : s_(std::move(s)), tr_(std::move(tr)) {}
  std::string name() const override { return "WaitRecv"; }
  CLONE_DEF(WaitRecv)
  void run(Platform &plat) override {
    if (tr_->null && plat.host_execution()) {
      host::flags_wait(s_->flag(s_->rank(), St<Scalar>::Reference, 0), s_->size(), plat.epoch());
      return;
    }
    if (tr_->null || plat.is_null()) return;
    CUDA_RUNTIME(cudaError_t(tzb_flags_wait(s_->flag(s_->rank(), St<Scalar>::Reference, 0), s_->size(),
                                            plat.epoch(), tr_->recv)));
    CUDA_RUNTIME(cudaStreamSynchronize(tr_->recv));
  }
};

} // namespace

template <typename Scalar>
SpMV<Scalar>::SpMV(const std::string &name, const std::shared_ptr<RowPartSpmv<Scalar>> &s, Style style)
    : name_(name) {
  Graph<OpBase> &g = graph_;
  OpPtr yl = std::make_shared<LocalSpMV<Scalar>>(s);
  if (style == Style::Reference) {
    auto tr = std::make_shared<SpmvTransport>(s->null_device());
    OpPtr pack = std::make_shared<Scatter<Scalar>>(s);
    OpPtr postSend = std::make_shared<PostSend<Scalar>>(s, tr);
    OpPtr postRecv = std::make_shared<NoOp>("PostRecv");
    OpPtr waitSend = std::make_shared<WaitSend<Scalar>>(tr);
    OpPtr waitRecv = std::make_shared<WaitRecv<Scalar>>(s, tr);
    OpPtr yr = std::make_shared<RemoteSpMV<Scalar>>(s, false, false, style);
    OpPtr y = std::make_shared<VectorAdd<Scalar>>(s);
    g.start_then(yl);
    g.start_then(postRecv);
    g.start_then(pack);
    g.then(pack, postSend);
    g.then(postSend, waitSend);
    g.then(postSend, waitRecv);
    g.then(postRecv, waitSend);
    g.then(postRecv, waitRecv);
    g.then(waitRecv, yr);
    g.then(yl, y);
    g.then(yr, y);
    g.then_finish(y);
    g.then_finish(waitSend);
  } else if (style == Style::FusedAll) {
    OpPtr all = std::make_shared<FusedAllSpMV<Scalar>>(s);
    g.start_then(all);
    g.then_finish(all);
  } else if (style == Style::FusedPull) {
    OpPtr ready = std::make_shared<XReady<Scalar>>(s);
    OpPtr yr = std::make_shared<PullSpMV<Scalar>>(s);
    OpPtr done = std::make_shared<XFree<Scalar>>(s);
    g.start_then(yl);
    g.start_then(ready);
    g.then(ready, yr);
    g.then(yl, yr);
    g.then(yr, done);
    g.then_finish(done);
  } else if (style == Style::Fused) {
    OpPtr put = std::make_shared<PackPut<Scalar>>(s, style);
    OpPtr yr = std::make_shared<RemoteSpMV<Scalar>>(s, true, true, style);
    g.start_then(yl);
    g.start_then(put);
    g.then(put, yr);
    g.then(yl, yr);
    g.then_finish(yr);
  } else {
    OpPtr put = std::make_shared<PackPut<Scalar>>(s, style);
    OpPtr yr = std::make_shared<RemoteSpMV<Scalar>>(s, false, true, style);
    OpPtr y = std::make_shared<VectorAdd<Scalar>>(s);
    g.start_then(yl);
    g.start_then(put);
    g.then(put, yr);
    g.then(yl, y);
    g.then(yr, y);
    g.then_finish(y);
  }
}

template <typename Scalar>
std::shared_ptr<ChoiceOp> make_spmv_choice(const std::shared_ptr<RowPartSpmv<Scalar>> &s,
                                           bool withExperimental) {
  using Style = typename RowPartSpmv<Scalar>::Style;
  std::vector<OpPtr> alts;
  alts.push_back(std::make_shared<SpMV<Scalar>>("spmv_reference", s, Style::Reference));
  alts.push_back(std::make_shared<SpMV<Scalar>>("spmv_fused", s, Style::Fused));
  alts.push_back(std::make_shared<SpMV<Scalar>>("spmv_fused_separate", s, Style::FusedSeparate));
  alts.push_back(std::make_shared<SpMV<Scalar>>("spmv_fused_all", s, Style::FusedAll));
  if (withExperimental) alts.push_back(std::make_shared<SpMV<Scalar>>("spmv_fused_pull", s, Style::FusedPull));
  return std::make_shared<ListChoiceOp>("spmv", alts);
}
template std::shared_ptr<ChoiceOp> make_spmv_choice<double>(const std::shared_ptr<RowPartSpmv<double>> &, bool);
template std::shared_ptr<ChoiceOp> make_spmv_choice<float>(const std::shared_ptr<RowPartSpmv<float>> &, bool);

template class RowPartSpmv<double>;
template class RowPartSpmv<float>;
template class SpMV<double>;
template class SpMV<float>;

} // namespace tenzing
