// 3-D halo (ghost-cell) exchange workload.
//
// Reference: include/tenzing/halo_exchange/{ops_halo_exchange,dim,cuda_memory}.hpp,
// src/halo_exchange/ops_halo_exchange.cu:33-257 (graph), :519-699 (kernels).
// Intended behaviour is implemented, not the shipped defects:
// all 26 neighbours (faces, edges, corners; 6 or 18 selectable), correct
// message extents, kernels that actually index by thread, no device-wide sync
// inside an op, and a graph builder that does not throw.
//
// Data plane: every rank's grid, staging buffers and flag words live in a
// SymmetricHeap, so "send" is a kernel storing into the neighbour's memory
// over NVLink and "receive complete" is a flag word written by the neighbour.
// These implementation styles are offered to the search as alternatives:
//   Reference  pack -> isend -> (irecv) -> wait -> unpack, host driven: the
//              reference's op structure with copy-engine peer copies standing
//              in for CUDA-aware MPI (the baseline, not the product)
//   Staged     pack fused with the put into the neighbour's staging slot;
//              flag wait fused with the unpack
//   Direct     put straight into the neighbour's ghost cells, one wait op
//   FusedAll   one kernel for all direct puts, one wait for all of them
//   FusedStaged one kernel packs+puts all regions, one kernel waits+unpacks all
//   FusedGroups the same in three direction groups (x-related / y / z), so the unpack of one
//              group can overlap the puts of the next
//   FusedMixed regions with an x component (24-byte rows) go through staging so they cross
//              NVLink as dense aligned bursts; all other regions (4 KB rows) are put straight
//              into the neighbour's ghost cells and need no unpack
//   FusedPipelined one put kernel that publishes every region's flag as soon as that region
//              is complete and one unpack kernel that may run at the same time and follows
//              region by region (thin regions first): transfer and unpack overlap
//   FusedPull  receiver driven: after a tiny "my interior is ready" signal, one kernel LOADS every
//              region straight from the neighbours' grids over NVLink into its own ghost cells
//              (no staging, no unpack, all stores local) and acknowledges; one wait for the acks
#pragma once

#include <array>
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "tenzing/cuda/ops_cuda.hpp"
#include "tenzing/cuda/symmetric_heap.hpp"
#include "tenzing/flag_synchronizer.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/operation_compound.hpp"
#include "tenzing_b200.h"

namespace tenzing {

template <typename T> struct Dim3 {
  T x{}, y{}, z{};
  Dim3() = default;
  Dim3(T x_, T y_, T z_) : x(x_), y(y_), z(z_) {}
  Dim3 operator+(const Dim3 &o) const { return {T(x + o.x), T(y + o.y), T(z + o.z)}; }
  Dim3 operator-(const Dim3 &o) const { return {T(x - o.x), T(y - o.y), T(z - o.z)}; }
  Dim3 operator-() const { return {T(-x), T(-y), T(-z)}; }
  Dim3 operator*(const Dim3 &o) const { return {T(x * o.x), T(y * o.y), T(z * o.z)}; }
  bool operator==(const Dim3 &o) const { return x == o.x && y == o.y && z == o.z; }
  bool operator!=(const Dim3 &o) const { return !(*this == o); }
  T flatten() const { return x * y * z; }
  T operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};

// shared_ptr to device memory released with cudaFree (reference: cuda_memory.hpp:13-26)
struct CudaFreeDeleter {
  void operator()(void *p) const;
};
template <typename T> std::shared_ptr<T> cuda_make_shared(size_t n) {
  void *p = nullptr;
  CUDA_RUNTIME(cudaMalloc(&p, n * sizeof(T)));
  return std::shared_ptr<T>(static_cast<T *>(p), CudaFreeDeleter());
}

namespace HaloExchange {

enum class StorageOrder { QXYZ, XYZQ };
enum class Style { Reference, Staged, Direct, FusedAll, FusedStaged, FusedGroups, FusedMixed, FusedPipelined, FusedPull };
const char *style_name(Style s);

// MPI-style message tag of a direction (reference: ops_halo_exchange.cu:16-27)
inline int dir_to_tag(const Dim3<int> &d) { return 9 * (d.z + 1) + 3 * (d.y + 1) + (d.x + 1); }
// "dx<d>_dy<d>_dz<d>" as in the reference's op names
std::string dir_suffix(const Dim3<int> &d);
// the exchange directions: 6 (faces), 18 (+edges) or 26 (+corners)
std::vector<Dim3<int>> directions(int neighbors);
// near-cubic rank grid from descending prime factors (reference: halo_run_strategy.hpp:80-94)
Dim3<int> rank_grid(int size);

struct Args {
  int nQ = 3;                  // quantities per cell
  int nX = 512, nY = 512, nZ = 512; // interior cells per rank
  int nGhost = 3;              // ghost radius
  size_t pitch = 128;          // XYZQ: rows are padded to a multiple of this many bytes
  StorageOrder storageOrder = StorageOrder::XYZQ;
  int neighbors = 26;
  int elemSize = 8;            // 8 = fp64 (default), 4 = fp32
  Dim3<int> rankGrid;          // ranks per dimension; periodic in every dimension
  std::function<Dim3<int>(int)> rankToCoord;       // defaults follow rankGrid, x fastest
  std::function<int(const Dim3<int> &)> coordToRank; // periodic wrap
  // fill rankGrid / rankToCoord / coordToRank for `size` ranks
  void set_default_topology(int size);
};

// Owns the symmetric allocation of one rank: grid + staging + flags.
class Domain {
public:
  Domain(std::shared_ptr<Communicator> comm, const Args &args,
         SymmetricHeap::Memory memory = SymmetricHeap::Memory::Device);

  const Args &args() const { return args_; }
  SymmetricHeap &heap() const { return *heap_; }
  int rank() const { return heap_->rank(); }
  int size() const { return heap_->size(); }

  // grid geometry, ghost cells included
  Dim3<int> extent() const { return ext_; }          // cells per dimension incl. ghosts
  size_t row_elems() const { return rowElems_; }     // padded row length in elements
  size_t x_origin() const { return xOrigin_; }       // elements before cell x = 0 in every row
  size_t grid_elems() const { return gridElems_; }
  size_t grid_bytes() const { return gridElems_ * size_t(args_.elemSize); }
  void *grid() const { return heap_->at<char>(rank(), gridOff_); }
  void *grid_of(int rank) const { return heap_->at<char>(rank, gridOff_); }
  // element index of cell (x,y,z) (ghost-inclusive coordinates) of quantity q
  size_t index(int x, int y, int z, int q) const;
  // element strides {x, y, z, q}
  std::array<int64_t, 4> strides() const;

  // region geometry for a direction
  Dim3<int> send_offset(const Dim3<int> &d) const; // interior cells adjacent to face d
  Dim3<int> recv_offset(const Dim3<int> &d) const; // ghost cells on side d
  Dim3<int> region_extent(const Dim3<int> &d) const;
  size_t region_elems(const Dim3<int> &d) const;   // incl. nQ
  int neighbor(const Dim3<int> &d) const;          // rank on side d (periodic)

  // bytes leaving this rank per exchange (all directions)
  size_t bytes_per_exchange() const;

  // staging slots: index by direction tag; recv staging is double buffered by epoch parity
  void *send_staging(int rank, const Dim3<int> &d) const;
  void *recv_staging(int rank, int parity, const Dim3<int> &side) const;
  // flag words: one set per style so alternatives never share a word; within a
  // set, word i belongs to dirs()[i], so the first dirs().size() words are exactly
  // the flags an exchange uses
  uint64_t *flag(int rank, Style style, const Dim3<int> &side) const;
  uint64_t *flags_base(int rank, Style style) const;
  // second flag set of the pull style ("I have read your data" acknowledgements)
  uint64_t *ack_flag(int rank, const Dim3<int> &side) const;
  uint64_t *ack_flags_base(int rank) const;
  // third set, shared by the styles that put straight into ghost cells: word `side` of rank r
  // is published by r's neighbour on that side when it allows r to overwrite the ghost cells it
  // holds for r ("I have consumed what you gave me last time")
  uint64_t *ready_flag(int rank, const Dim3<int> &side) const;
  int dir_index(const Dim3<int> &d) const;

  const std::vector<Dim3<int>> &dirs() const { return dirs_; }

private:
  Args args_;
  std::unique_ptr<SymmetricHeap> heap_;
  std::vector<Dim3<int>> dirs_;
  Dim3<int> ext_;
  size_t rowElems_ = 0, gridElems_ = 0, xOrigin_ = 0;
  size_t gridOff_ = 0, sendOff_ = 0, recvOff_[2] = {0, 0}, flagOff_ = 0;
  std::array<int, 27> dirIndex_{};
  std::array<size_t, 27> slotOff_{}; // element offset of each direction's slot inside a staging area
  size_t stagingElems_ = 0;
};

// A GPU op that runs one box-copy plan per epoch parity (see tenzing_b200.h).
class BoxCopyOp : public GpuOp, public HasFlags {
  struct Plan {
    TzbBoxPlanInfo info{};
    std::shared_ptr<void> dev;     // device image
    std::shared_ptr<void> scratch; // per-box counters
  };
  std::string name_;
  std::shared_ptr<std::array<Plan, 2>> plans_; // shared between clones
  // host-memory domains keep the boxes themselves: Platform::host_execution() runs them
  std::shared_ptr<std::array<std::vector<TzbBox>, 2>> hostBoxes_;
  int elemSize_ = 8;
  bool ordered_ = false;
  bool null_ = false;
  bool issueOrderOnly_ = false;
  bool publishesFirst_ = false; // some box grants a permission (TzbBox::ready) before any wait
  std::vector<uint64_t> published_, awaited_; // flag channels (heap offsets), see set_heap()

public:
  // boxes[parity]: the copies to perform when (epoch & 1) == parity
  // ordered: boxes are processed in the given order and publish their flags one by one
  BoxCopyOp(const std::string &name, const std::vector<TzbBox> &even,
            const std::vector<TzbBox> &odd, int elemSize, bool nullDevice = false,
            bool ordered = false);
  void set_issue_order_only(bool v) { issueOrderOnly_ = v; }
  // translate the boxes' flag pointers into heap offsets so FlagSynchronizer can reason about them
  void set_flag_channels(std::vector<uint64_t> published, std::vector<uint64_t> awaited,
                         bool publishesFirst = false) {
    published_ = std::move(published);
    awaited_ = std::move(awaited);
    publishesFirst_ = publishesFirst;
  }
  bool publishes_before_waiting() const override { return publishesFirst_; }
  std::vector<uint64_t> flags_published() const override { return published_; }
  std::vector<uint64_t> flags_awaited() const override { return awaited_; }
  bool issue_order_only() const override { return issueOrderOnly_; }
  std::string name() const override { return name_; }
  CLONE_DEF(BoxCopyOp)
  void run(cudaStream_t) override { THROW_RUNTIME("BoxCopyOp needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t stream) override;
  size_t payload_bytes() const { return (*plans_)[0].info.payloadBytes; }
  int num_items() const { return (*plans_)[0].info.nItems; }
};

// names follow the reference: he_pack_<dir>, he_isend_<dir>, he_irecv_<dir>,
// he_waitrecv_<dir>, he_unpack_<dir>, he_wait_sends; fused ops: he_packput_<dir>,
// he_waitunpack_<dir>, he_put_<dir>, he_wait_puts, he_put_all, he_finish_all
//
// Adds the ops of one style between `preds` and `succs` (use g.start()/g.finish()
// for the whole-graph case).  Every op that publishes data to a neighbour
// precedes every op that waits for a neighbour, so any order the search picks
// is deadlock free when all ranks execute the same sequence.
void add_to_graph(Graph<OpBase> &g, const std::shared_ptr<Domain> &dom, const std::vector<OpPtr> &preds,
                  const std::vector<OpPtr> &succs, Style style = Style::Staged);

// the exchange as one compound op of a given style
std::shared_ptr<CompoundOp> make_compound(const std::shared_ptr<Domain> &dom, Style style);
// ChoiceOp "halo_exchange" over the compound ops of the given styles
std::shared_ptr<ChoiceOp> make_choice(const std::shared_ptr<Domain> &dom,
                                      const std::vector<Style> &styles = {Style::Reference,
                                                                          Style::Staged,
                                                                          Style::Direct,
                                                                          Style::FusedAll,
                                                                          Style::FusedStaged,
                                                                          Style::FusedGroups,
                                                                          Style::FusedMixed,
                                                                          Style::FusedPipelined,
                                                                          Style::FusedPull});

} // namespace HaloExchange
} // namespace tenzing
