// Row-partitioned distributed SpMV: set-up, device data and the ops the search schedules.
//
// Reference: include/tenzing/spmv/row_part_spmv.cuh:18-445 (set-up),
// ops_spmv.cuh:61-436 (ops and the SpMV compound), src/spmv/ops_spmv.cu.
// No MPI: every rank builds its own row block, the "which x entries does each
// rank need" exchange runs over the Communicator, and x entries travel as
// kernel stores into the owner's receive buffer in a SymmetricHeap.
// y = A_local * lx + A_remote * rx is computed correctly (the reference
// overwrites the local product with the remote one and its VectorAdd is
// empty).  fp64 and fp32 are supported (reference: fp32 only, D12).
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "tenzing/cuda/ops_cuda.hpp"
#include "tenzing/cuda/symmetric_heap.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/operation_compound.hpp"
#include "tenzing/spmv/csr_mat.hpp"

namespace tenzing {

// device CSR (plain device allocations, owned)
template <typename Scalar> struct DeviceCsr {
  int rows = 0, cols = 0;
  int64_t nnz = 0;
  std::shared_ptr<void> rowPtr, colInd, val;
  const int *row_ptr() const { return static_cast<const int *>(rowPtr.get()); }
  const int *col_ind() const { return static_cast<const int *>(colInd.get()); }
  const Scalar *values() const { return static_cast<const Scalar *>(val.get()); }
  int nnz_per_row() const { return rows ? int((nnz + rows - 1) / rows) : 1; }
};

template <typename Scalar> class RowPartSpmv {
public:
  // a block of x entries exchanged with one peer
  struct Param {
    int peer = 0;
    int displ = 0;    // offset in my send list (send) / in my rx (recv)
    int count = 0;
    int dstDispl = 0; // send only: offset in the peer's rx
  };
  // FusedPull (validated against the oracle on 2 and 8 B200s, not timed there) is built on request
  // only and is not among the alternatives of make_spmv_choice unless asked for
  enum class Style { Reference, Fused, FusedSeparate, FusedAll, FusedPull };
  static constexpr int kFlagSets = 6; // one per style + the acknowledgements of FusedPull

  // myRows: the rows [get_partition(n, rank, size)) of the global n x n matrix, global column ids
  RowPartSpmv(std::shared_ptr<Communicator> comm, const CsrMat<int, Scalar> &myRows, int64_t n,
              SymmetricHeap::Memory memory = SymmetricHeap::Memory::Device);

  int rank() const { return heap_->rank(); }
  int size() const { return heap_->size(); }
  int64_t global_rows() const { return n_; }
  Range row_range() const { return range_; }
  SymmetricHeap &heap() const { return *heap_; }
  bool null_device() const { return heap_->memory() == SymmetricHeap::Memory::Host; }

  const DeviceCsr<Scalar> &lA() const { return lA_; }
  const DeviceCsr<Scalar> &rA() const { return rA_; }
  const SplitMat<int, Scalar> &host_split() const { return split_; }
  Scalar *lx(int rank) const { return heap_->at<Scalar>(rank, lxOff_); }
  Scalar *lx() const { return lx(rank()); }
  Scalar *rx(int rank, int parity) const { return heap_->at<Scalar>(rank, rxOff_[parity & 1]); }
  Scalar *ly() const { return heap_->at<Scalar>(rank(), lyOff_); }
  Scalar *ry() const { return heap_->at<Scalar>(rank(), ryOff_); }
  Scalar *x_send_buf() const { return heap_->at<Scalar>(rank(), sendOff_); }
  const int *x_send_idx() const { return static_cast<const int *>(xSendIdx_.get()); }
  int num_send() const { return numSend_; }
  int num_remote_cols() const { return int(split_.globals.size()); }
  const std::vector<Param> &send_params() const { return sendParams_; }
  const std::vector<Param> &recv_params() const { return recvParams_; }
  // ranks I exchange with in either direction (the handshake is kept symmetric)
  const std::vector<int> &peers() const { return peers_; }
  // flag word that `src` publishes on rank `rank` for a style
  uint64_t *flag(int rank, Style style, int src) const {
    return heap_->at<uint64_t>(rank, flagOff_) + size_t(int(style)) * size_t(size()) + size_t(src);
  }
  // second flag set of the pull style: "rank `src` has finished reading rank `rank`'s x"
  uint64_t *ack_flag(int rank, int src) const {
    return heap_->at<uint64_t>(rank, flagOff_) + size_t(kFlagSets - 1) * size_t(size()) + size_t(src);
  }
  // pull style: address of every remote column in its owner's x (peer pointers), device array
  const Scalar *const *x_ptrs() const { return static_cast<const Scalar *const *>(xPtrs_.get()); }
  uint32_t *counter() const { return static_cast<uint32_t *>(counter_.get()); }
  // bytes of x this rank receives per product
  size_t recv_bytes() const { return size_t(num_remote_cols()) * sizeof(Scalar); }
  // algorithmic HBM bytes of one product on this rank (matrix + x + y streams)
  size_t hbm_bytes() const;

private:
  std::unique_ptr<SymmetricHeap> heap_;
  int64_t n_;
  Range range_;
  SplitMat<int, Scalar> split_;
  DeviceCsr<Scalar> lA_, rA_;
  std::shared_ptr<void> xSendIdx_, counter_, xPtrs_;
  int numSend_ = 0;
  std::vector<Param> sendParams_, recvParams_;
  std::vector<int> peers_;
  size_t lxOff_ = 0, rxOff_[2] = {0, 0}, lyOff_ = 0, ryOff_ = 0, sendOff_ = 0, flagOff_ = 0;
};

// The compound op "spmv" (reference: ops_spmv.cuh:306-436).  Reference style
// expands to the reference's 9 vertices
//   Start -> {yl, PostRecv, Pack}; Pack -> PostSend; {PostSend, PostRecv} -> {WaitSend, WaitRecv};
//   WaitRecv -> yr; {yl, yr} -> y; {y, WaitSend} -> Finish
// with copy-engine peer copies as the transport.  Fused:
//   Start -> {yl, PackPut}; {yl, PackPut} -> yr (waits on flags, accumulates into y) -> Finish
// FusedSeparate keeps yl and yr independent and adds them in `y`.  FusedAll is one op: a
// single kernel that puts, multiplies locally, waits for the peers and accumulates the remote part.
// FusedPull moves nothing ahead of time: XReady tells every rank "my x is final", yrPull waits
// for those flags and reads the remote x entries straight from their owners over NVLink while
// it multiplies (the gather is fused into the off-diagonal product), then acknowledges;
// XFree waits for the acknowledgements, after which x may be overwritten:
//   Start -> {yl, XReady}; {yl, XReady} -> yrPull -> XFree -> Finish
template <typename Scalar> class SpMV : public CompoundOp {
  std::string name_;
  Graph<OpBase> graph_;

public:
  using Style = typename RowPartSpmv<Scalar>::Style;
  SpMV(const std::string &name, const std::shared_ptr<RowPartSpmv<Scalar>> &spmv,
       Style style = Style::Fused);
  std::string name() const override { return name_; }
  Graph<OpBase> graph() const override { return graph_; }
  CLONE_DEF(SpMV)
};

// ChoiceOp "spmv" over the four default styles (+ the pull style on request)
template <typename Scalar>
std::shared_ptr<ChoiceOp> make_spmv_choice(const std::shared_ptr<RowPartSpmv<Scalar>> &spmv,
                                           bool withExperimental = false);

// Root-held matrix -> one row block per rank (reference: send_matrix / receive_matrix,
// row_part_spmv.cuh:18-90, four MPI_Isend + Probe/Recv per peer).  Rank `root` passes the whole
// matrix, the others pass nullptr; every rank gets its get_partition() block with global
// column ids.  Blocks travel as broadcasts (set-up only), so any Communicator will do.
template <typename Scalar>
CsrMat<int, Scalar> distribute_rows(Communicator &comm, const CsrMat<int, Scalar> *whole, int root = 0);

// convenience: band matrix block generated in place on every rank
template <typename Scalar>
std::shared_ptr<RowPartSpmv<Scalar>>
make_band_spmv(std::shared_ptr<Communicator> comm, int64_t n, int64_t bw, int nnzPerRow,
               uint64_t seed = 0, SymmetricHeap::Memory memory = SymmetricHeap::Memory::Device);

extern template class RowPartSpmv<double>;
extern template class RowPartSpmv<float>;
extern template class SpMV<double>;
extern template class SpMV<float>;

} // namespace tenzing
