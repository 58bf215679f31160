#include "tenzing/halo/halo_exchange.hpp"
#include "tenzing/host_exec.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "tenzing/numeric.hpp"

namespace tenzing {

void CudaFreeDeleter::operator()(void *p) const {
  if (p) cudaFree(p);
}

namespace HaloExchange {

const char *style_name(Style s) {
  switch (s) {
  case Style::Reference: return "reference";
  case Style::Staged: return "staged";
  case Style::Direct: return "direct";
  case Style::FusedAll: return "fused_all";
  case Style::FusedStaged: return "fused_staged";
  case Style::FusedGroups: return "fused_groups";
  case Style::FusedMixed: return "fused_mixed";
  case Style::FusedPipelined: return "fused_pipelined";
  case Style::FusedPull: return "fused_pull";
  }
  return "?";
}

std::string dir_suffix(const Dim3<int> &d) {
  return "dx" + std::to_string(d.x) + "_dy" + std::to_string(d.y) + "_dz" + std::to_string(d.z);
}

std::vector<Dim3<int>> directions(int neighbors) {
  if (neighbors != 6 && neighbors != 18 && neighbors != 26)
    THROW_RUNTIME("halo neighbors must be 6, 18 or 26");
  std::vector<Dim3<int>> r;
  for (int dz = -1; dz <= 1; ++dz)
    for (int dy = -1; dy <= 1; ++dy)
      for (int dx = -1; dx <= 1; ++dx) {
        const int nz = (dx != 0) + (dy != 0) + (dz != 0);
        if (nz == 0) continue;
        if (neighbors == 6 && nz != 1) continue;
        if (neighbors == 18 && nz > 2) continue;
        r.emplace_back(dx, dy, dz);
      }
  return r;
}

Dim3<int> rank_grid(int size) {
  Dim3<int> g(1, 1, 1);
  for (int f : prime_factors(size)) {
    // give the factor to the currently smallest dimension, preferring x, then y, then z
    if (g.x <= g.y && g.x <= g.z) g.x *= f;
    else if (g.y <= g.z) g.y *= f;
    else g.z *= f;
  }
  return g;
}

void Args::set_default_topology(int size) {
  if (rankGrid.flatten() != size) rankGrid = rank_grid(size);
  const Dim3<int> rg = rankGrid;
  rankToCoord = [rg](int r) { return Dim3<int>(r % rg.x, (r / rg.x) % rg.y, r / (rg.x * rg.y)); };
  coordToRank = [rg](const Dim3<int> &c) {
    auto wrap = [](int v, int n) { return ((v % n) + n) % n; };
    return wrap(c.x, rg.x) + wrap(c.y, rg.y) * rg.x + wrap(c.z, rg.z) * rg.x * rg.y;
  };
}

// ---------------------------------------------------------------- Domain
namespace {
constexpr int kNumStyles = 9; // + two extra flag sets: acknowledgements of the pull style, permissions of the direct puts
size_t round_up_sz(size_t v, size_t a) { return (v + a - 1) / a * a; }
} // namespace

Domain::Domain(std::shared_ptr<Communicator> comm, const Args &args, SymmetricHeap::Memory memory)
    : args_(args) {
  if (!comm) comm = std::make_shared<SelfComm>();
  if (!args_.rankToCoord || !args_.coordToRank) args_.set_default_topology(comm->size());
  if (args_.elemSize != 4 && args_.elemSize != 8) THROW_RUNTIME("elemSize must be 4 or 8");
  if (args_.nX < 1 || args_.nY < 1 || args_.nZ < 1 || args_.nQ < 1)
    THROW_RUNTIME("halo domain needs at least one cell and one quantity per rank (nX=" << args_.nX
                  << " nY=" << args_.nY << " nZ=" << args_.nZ << " nQ=" << args_.nQ << ")");
  // a ghost layer wider than the interior would need cells of the neighbour's neighbour
  if (args_.nGhost < 0 || args_.nGhost > std::min(args_.nX, std::min(args_.nY, args_.nZ)))
    THROW_RUNTIME("ghost radius " << args_.nGhost << " must lie in [0, smallest interior extent]");
  if (args_.rankGrid.flatten() != 0 && args_.rankGrid.flatten() != comm->size()) // 0: custom topology functions
    THROW_RUNTIME("rank grid " << args_.rankGrid.x << "x" << args_.rankGrid.y << "x" << args_.rankGrid.z
                               << " does not have " << comm->size() << " ranks");
  dirs_ = directions(args_.neighbors);
  const int g = args_.nGhost;
  ext_ = Dim3<int>(args_.nX + 2 * g, args_.nY + 2 * g, args_.nZ + 2 * g);
  const size_t es = size_t(args_.elemSize);
  const size_t pitch = args_.pitch ? args_.pitch : es;
  if (pitch % es) THROW_RUNTIME("pitch must be a multiple of the element size");
  // Rows are laid out so that the first INTERIOR cell of every row is 16-byte aligned: the
  // 4 KB interior runs (y/z faces, staging copies) can then be moved with 16-byte accesses on
  // both sides and cross NVLink as whole sectors, and each 24-byte ghost / send piece of an
  // x-face sits inside one 32-byte sector, next to its partner in the same 128-byte line.
  const size_t cellElems = args_.storageOrder == StorageOrder::XYZQ ? 1 : size_t(args_.nQ);
  xOrigin_ = 0;
  while (((xOrigin_ + size_t(g) * cellElems) * es) % 16) ++xOrigin_;
  const size_t rowUnpadded = (xOrigin_ + size_t(ext_.x) * cellElems) * es;
  rowElems_ = round_up_sz(rowUnpadded, pitch) / es;
  gridElems_ = rowElems_ * size_t(ext_.y) * size_t(ext_.z) *
               (args_.storageOrder == StorageOrder::XYZQ ? size_t(args_.nQ) : size_t(1));

  // staging slots, one per direction tag, each 256 B aligned
  size_t off = 0;
  slotOff_.fill(0);
  dirIndex_.fill(-1);
  for (size_t i = 0; i < dirs_.size(); ++i) dirIndex_[size_t(dir_to_tag(dirs_[i]))] = int(i);
  for (const auto &d : dirs_) {
    slotOff_[size_t(dir_to_tag(d))] = off;
    off += round_up_sz(region_elems(d) * es, 256) / es;
  }
  stagingElems_ = off;

  const size_t stagingBytes = stagingElems_ * es;
  const size_t flagBytes = sizeof(uint64_t) * 27 * (kNumStyles + 2);
  const size_t total = round_up_sz(grid_bytes(), 256) + 3 * round_up_sz(stagingBytes, 256) +
                       round_up_sz(flagBytes, 256) + 4096;
  heap_.reset(new SymmetricHeap(comm, total, memory));
  gridOff_ = heap_->alloc(grid_bytes());
  sendOff_ = heap_->alloc(stagingBytes);
  recvOff_[0] = heap_->alloc(stagingBytes);
  recvOff_[1] = heap_->alloc(stagingBytes);
  flagOff_ = heap_->alloc(flagBytes);
}

std::array<int64_t, 4> Domain::strides() const {
  if (args_.storageOrder == StorageOrder::XYZQ) {
    const int64_t sy = int64_t(rowElems_), sz = sy * ext_.y, sq = sz * ext_.z;
    return {1, sy, sz, sq};
  }
  const int64_t sy = int64_t(rowElems_), sz = sy * ext_.y;
  return {int64_t(args_.nQ), sy, sz, 1};
}

size_t Domain::index(int x, int y, int z, int q) const {
  const auto s = strides();
  return xOrigin_ + size_t(x * s[0] + y * s[1] + z * s[2] + q * s[3]);
}

Dim3<int> Domain::region_extent(const Dim3<int> &d) const {
  const int g = args_.nGhost;
  return Dim3<int>(d.x ? g : args_.nX, d.y ? g : args_.nY, d.z ? g : args_.nZ);
}

size_t Domain::region_elems(const Dim3<int> &d) const {
  const Dim3<int> e = region_extent(d);
  return size_t(e.x) * size_t(e.y) * size_t(e.z) * size_t(args_.nQ);
}

Dim3<int> Domain::send_offset(const Dim3<int> &d) const {
  const int g = args_.nGhost;
  auto off = [g](int dir, int n) { return dir > 0 ? n : g; }; // last g interior cells / first g / all
  return Dim3<int>(off(d.x, args_.nX), off(d.y, args_.nY), off(d.z, args_.nZ));
}

Dim3<int> Domain::recv_offset(const Dim3<int> &d) const {
  const int g = args_.nGhost;
  auto off = [g](int dir, int n) { return dir > 0 ? n + g : (dir < 0 ? 0 : g); };
  return Dim3<int>(off(d.x, args_.nX), off(d.y, args_.nY), off(d.z, args_.nZ));
}

int Domain::neighbor(const Dim3<int> &d) const {
  return args_.coordToRank(args_.rankToCoord(rank()) + d);
}

size_t Domain::bytes_per_exchange() const {
  size_t n = 0;
  for (const auto &d : dirs_) n += region_elems(d);
  return n * size_t(args_.elemSize);
}

void *Domain::send_staging(int rank, const Dim3<int> &d) const {
  return heap_->at<char>(rank, sendOff_) + slotOff_[size_t(dir_to_tag(d))] * size_t(args_.elemSize);
}

void *Domain::recv_staging(int rank, int parity, const Dim3<int> &side) const {
  return heap_->at<char>(rank, recvOff_[parity & 1]) +
         slotOff_[size_t(dir_to_tag(side))] * size_t(args_.elemSize);
}

int Domain::dir_index(const Dim3<int> &d) const {
  const int i = dirIndex_[size_t(dir_to_tag(d))];
  if (i < 0) THROW_RUNTIME("direction " << dir_suffix(d) << " is not exchanged");
  return i;
}

uint64_t *Domain::flags_base(int rank, Style style) const {
  return heap_->at<uint64_t>(rank, flagOff_) + size_t(int(style)) * 27;
}

uint64_t *Domain::flag(int rank, Style style, const Dim3<int> &side) const {
  return flags_base(rank, style) + dir_index(side);
}

uint64_t *Domain::ack_flags_base(int rank) const {
  return heap_->at<uint64_t>(rank, flagOff_) + size_t(kNumStyles) * 27;
}

uint64_t *Domain::ack_flag(int rank, const Dim3<int> &side) const {
  return ack_flags_base(rank) + dir_index(side);
}

uint64_t *Domain::ready_flag(int rank, const Dim3<int> &side) const {
  return heap_->at<uint64_t>(rank, flagOff_) + size_t(kNumStyles + 1) * 27 + dir_index(side);
}

// ---------------------------------------------------------------- BoxCopyOp
BoxCopyOp::BoxCopyOp(const std::string &name, const std::vector<TzbBox> &even,
                     const std::vector<TzbBox> &odd, int elemSize, bool nullDevice, bool ordered)
    : name_(name), plans_(std::make_shared<std::array<Plan, 2>>()), elemSize_(elemSize),
      ordered_(ordered), null_(nullDevice) {
  if (null_) {
    hostBoxes_ = std::make_shared<std::array<std::vector<TzbBox>, 2>>();
    (*hostBoxes_)[0] = even;
    (*hostBoxes_)[1] = odd;
  }
  auto make_plan = ordered ? tzb_boxcopy_plan_ordered : tzb_boxcopy_plan;
  const std::vector<TzbBox> *lists[2] = {&even, &odd};
  for (int p = 0; p < 2; ++p) {
    Plan &plan = (*plans_)[size_t(p)];
    const auto &boxes = *lists[p];
    if (p == 1 && boxes.size() == even.size() &&
        (boxes.empty() || std::memcmp(boxes.data(), even.data(), sizeof(TzbBox) * boxes.size()) == 0)) {
      plan = (*plans_)[0]; // parity-independent op: share one plan
      continue;
    }
    CUDA_RUNTIME(cudaError_t(make_plan(boxes.data(), int(boxes.size()), elemSize, &plan.info, nullptr)));
    std::vector<char> image(plan.info.planBytes);
    CUDA_RUNTIME(cudaError_t(
        make_plan(boxes.data(), int(boxes.size()), elemSize, &plan.info, image.data())));
    if (null_) continue;
    void *dev = nullptr, *scratch = nullptr;
    CUDA_RUNTIME(cudaMalloc(&dev, plan.info.planBytes));
    plan.dev = std::shared_ptr<void>(dev, CudaFreeDeleter());
    CUDA_RUNTIME(cudaMemcpy(dev, image.data(), plan.info.planBytes, cudaMemcpyHostToDevice));
    CUDA_RUNTIME(cudaMalloc(&scratch, plan.info.scratchBytes));
    plan.scratch = std::shared_ptr<void>(scratch, CudaFreeDeleter());
    CUDA_RUNTIME(cudaMemset(scratch, 0, plan.info.scratchBytes));
  }
  if (!null_) CUDA_RUNTIME(cudaDeviceSynchronize());
}

void BoxCopyOp::run(Platform &plat, cudaStream_t stream) {
  if (null_ && plat.host_execution()) {
    const auto &boxes = (*hostBoxes_)[size_t(plat.epoch() & 1)];
    host::boxcopy(boxes.data(), int(boxes.size()), elemSize_, plat.epoch(), ordered_);
    return;
  }
  if (null_ || plat.is_null()) return;
  const Plan &plan = (*plans_)[size_t(plat.epoch() & 1)];
  CUDA_RUNTIME(cudaError_t(
      tzb_boxcopy_launch(plan.dev.get(), &plan.info, plat.epoch(), plan.scratch.get(), stream)));
}

// ---------------------------------------------------------------- box geometry
namespace {

enum class Side { Grid, Dense };

// the copy of region (offset `off`, extent `e`) between a grid-shaped and/or
// dense buffer; strides in elements
TzbBox make_box(const Domain &dom, const void *src, Side srcSide, const Dim3<int> &srcOff, void *dst,
                Side dstSide, const Dim3<int> &dstOff, const Dim3<int> &e) {
  const Args &a = dom.args();
  const auto gs = dom.strides(); // {x, y, z, q}
  const size_t es = size_t(a.elemSize);
  TzbBox b;
  std::memset(&b, 0, sizeof(b));
  auto grid_base = [&](const void *p, const Dim3<int> &o) {
    return static_cast<const char *>(p) +
           (dom.x_origin() + size_t(o.x * gs[0] + o.y * gs[1] + o.z * gs[2])) * es;
  };
  b.src = srcSide == Side::Grid ? grid_base(src, srcOff) : static_cast<const char *>(src);
  b.dst = const_cast<char *>(dstSide == Side::Grid ? grid_base(dst, dstOff) : static_cast<const char *>(dst));
  if (dstSide == Side::Grid) {
    // what lies next to a destination run inside its row: padding if the run starts at the
    // first / ends at the last cell of the row (ghost cells of the x faces), live cells otherwise
    const int64_t cell = a.storageOrder == StorageOrder::XYZQ ? 1 : a.nQ;
    if (dstOff.x == 0) b.dstSlackLo = int32_t(dom.x_origin());
    if (dstOff.x + e.x == dom.extent().x)
      b.dstSlackHi = int32_t(int64_t(dom.row_elems()) - int64_t(dom.x_origin()) - int64_t(dom.extent().x) * cell);
  }
  if (a.storageOrder == StorageOrder::XYZQ) {
    // run = x, level1 = y, level2 = z, level3 = q
    b.run = e.x, b.n1 = e.y, b.n2 = e.z, b.n3 = a.nQ;
    const int64_t dense[3] = {int64_t(e.x), int64_t(e.x) * e.y, int64_t(e.x) * e.y * e.z};
    const int64_t grid[3] = {gs[1], gs[2], gs[3]};
    const int64_t *ss = srcSide == Side::Grid ? grid : dense;
    const int64_t *ds = dstSide == Side::Grid ? grid : dense;
    b.s1 = ss[0], b.s2 = ss[1], b.s3 = ss[2];
    b.d1 = ds[0], b.d2 = ds[1], b.d3 = ds[2];
  } else {
    // run = x*q (cells of a row are contiguous with their quantities), level1 = y, level2 = z
    b.run = e.x * a.nQ, b.n1 = e.y, b.n2 = e.z, b.n3 = 1;
    const int64_t dense[3] = {int64_t(e.x) * a.nQ, int64_t(e.x) * a.nQ * e.y, 0};
    const int64_t grid[3] = {gs[1], gs[2], 0};
    const int64_t *ss = srcSide == Side::Grid ? grid : dense;
    const int64_t *ds = dstSide == Side::Grid ? grid : dense;
    b.s1 = ss[0], b.s2 = ss[1], b.s3 = ss[2];
    b.d1 = ds[0], b.d2 = ds[1], b.d3 = ds[2];
  }
  return b;
}

// shared by the host-driven reference-structure ops: the two internal streams
// that play the role of the MPI progress engine
struct RefTransport {
  static constexpr int kSendStreams = 4; // peer copies are pipelined over several copy streams
  cudaStream_t send[kSendStreams] = {nullptr, nullptr, nullptr, nullptr};
  cudaStream_t recv = nullptr;
  int next = 0;
  bool null = false;
  explicit RefTransport(bool nullDevice) : null(nullDevice) {
    if (null) return;
    for (auto &s : send) CUDA_RUNTIME(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    CUDA_RUNTIME(cudaStreamCreateWithFlags(&recv, cudaStreamNonBlocking));
  }
  ~RefTransport() {
    for (auto s : send)
      if (s) cudaStreamDestroy(s);
    if (recv) cudaStreamDestroy(recv);
  }
  cudaStream_t next_send_stream() {
    cudaStream_t s = send[next];
    next = (next + 1) % kSendStreams;
    return s;
  }
};

// "MPI_Isend" of a packed buffer: copy-engine peer copy + flag publish, host posted
class RefIsend : public CpuOp {
  std::string name_;
  std::shared_ptr<Domain> dom_;
  std::shared_ptr<RefTransport> tr_;
  Dim3<int> d_;

public:
  RefIsend(const std::string &name, std::shared_ptr<Domain> dom, std::shared_ptr<RefTransport> tr,
           const Dim3<int> &d)
      : name_(name), dom_(std::move(dom)), tr_(std::move(tr)), d_(d) {}
  std::string name() const override { return name_; }
  CLONE_DEF(RefIsend)
  void run(Platform &plat) override {
    const int peer = dom_->neighbor(d_);
    const Dim3<int> side = -d_;
    const size_t bytes = dom_->region_elems(d_) * size_t(dom_->args().elemSize);
    if (tr_->null && plat.host_execution()) {
      host::copy_bytes(dom_->recv_staging(peer, int(plat.epoch() & 1), side),
                       dom_->send_staging(dom_->rank(), d_), bytes);
      host::flag_signal(dom_->flag(peer, Style::Reference, side), plat.epoch());
      return;
    }
    if (tr_->null || plat.is_null()) return;
    cudaStream_t st = tr_->next_send_stream();
    CUDA_RUNTIME(cudaMemcpyAsync(dom_->recv_staging(peer, int(plat.epoch() & 1), side),
                                 dom_->send_staging(dom_->rank(), d_), bytes,
                                 cudaMemcpyDeviceToDevice, st));
    TzbFlagList l;
    l.n = 1;
    l.flag[0] = dom_->flag(peer, Style::Reference, side);
    CUDA_RUNTIME(cudaError_t(tzb_flags_signal(&l, plat.epoch(), st)));
  }
};

// "MPI_Wait" on a receive: the host blocks until the neighbour's flag arrived
class RefWaitRecv : public CpuOp {
  std::string name_;
  std::shared_ptr<Domain> dom_;
  std::shared_ptr<RefTransport> tr_;
  Dim3<int> side_;

public:
  RefWaitRecv(const std::string &name, std::shared_ptr<Domain> dom,
              std::shared_ptr<RefTransport> tr, const Dim3<int> &side)
      : name_(name), dom_(std::move(dom)), tr_(std::move(tr)), side_(side) {}
  std::string name() const override { return name_; }
  CLONE_DEF(RefWaitRecv)
  void run(Platform &plat) override {
    if (tr_->null && plat.host_execution()) {
      host::flags_wait(dom_->flag(dom_->rank(), Style::Reference, side_), 1, plat.epoch());
      return;
    }
    if (tr_->null || plat.is_null()) return;
    CUDA_RUNTIME(cudaError_t(tzb_flags_wait(dom_->flag(dom_->rank(), Style::Reference, side_), 1,
                                            plat.epoch(), tr_->recv)));
    CUDA_RUNTIME(cudaStreamSynchronize(tr_->recv));
  }
};

// "MPI_Waitall" on the sends
class RefWaitSends : public CpuOp {
  std::shared_ptr<RefTransport> tr_;

public:
  explicit RefWaitSends(std::shared_ptr<RefTransport> tr) : tr_(std::move(tr)) {}
  std::string name() const override { return "he_wait_sends"; }
  CLONE_DEF(RefWaitSends)
  void run(Platform &plat) override {
    if (tr_->null || plat.is_null()) return;
    for (auto s : tr_->send) CUDA_RUNTIME(cudaStreamSynchronize(s));
  }
};

// GPU op that waits for a set of local flags (completion of direct puts)
class WaitFlagsOp : public GpuOp, public HasFlags {
  std::string name_;
  const uint64_t *flags_;
  int n_;
  bool null_;
  std::vector<uint64_t> awaited_;

public:
  WaitFlagsOp(const std::string &name, const uint64_t *flags, int n, bool nullDevice)
      : name_(name), flags_(flags), n_(n), null_(nullDevice) {}
  void set_channels(const SymmetricHeap &heap) {
    awaited_.clear();
    for (int i = 0; i < n_; ++i) awaited_.push_back(uint64_t(heap.offset_of(flags_ + i)));
  }
  std::vector<uint64_t> flags_published() const override { return {}; }
  std::vector<uint64_t> flags_awaited() const override { return awaited_; }
  std::string name() const override { return name_; }
  CLONE_DEF(WaitFlagsOp)
  void run(cudaStream_t) override { THROW_RUNTIME("WaitFlagsOp needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t stream) override {
    if (null_ && plat.host_execution()) {
      host::flags_wait(flags_, n_, plat.epoch());
      return;
    }
    if (null_ || plat.is_null()) return;
    CUDA_RUNTIME(cudaError_t(tzb_flags_wait(flags_, n_, plat.epoch(), stream)));
  }
};

// GPU op that publishes the epoch to a list of (peer) flags
class SignalFlagsOp : public GpuOp, public HasFlags {
  std::string name_;
  TzbFlagList list_;
  bool null_;
  std::vector<uint64_t> published_;

public:
  SignalFlagsOp(const std::string &name, const std::vector<uint64_t *> &flags, const SymmetricHeap &heap,
                bool nullDevice)
      : name_(name), null_(nullDevice) {
    if (flags.size() > 32) THROW_RUNTIME("at most 32 flags per signal op");
    list_.n = int(flags.size());
    for (size_t i = 0; i < flags.size(); ++i) {
      list_.flag[i] = flags[i];
      published_.push_back(uint64_t(heap.offset_of(flags[i])));
    }
  }
  std::string name() const override { return name_; }
  CLONE_DEF(SignalFlagsOp)
  void run(cudaStream_t) override { THROW_RUNTIME("SignalFlagsOp needs the platform epoch"); }
  void run(Platform &plat, cudaStream_t stream) override {
    if (null_ && plat.host_execution()) {
      host::flags_signal(list_, plat.epoch());
      return;
    }
    if (null_ || plat.is_null()) return;
    CUDA_RUNTIME(cudaError_t(tzb_flags_signal(&list_, plat.epoch(), stream)));
  }
  std::vector<uint64_t> flags_published() const override { return published_; }
  std::vector<uint64_t> flags_awaited() const override { return {}; }
};

struct Builder {
  std::shared_ptr<Domain> dom;
  bool null;
  int es;

  explicit Builder(std::shared_ptr<Domain> d)
      : dom(std::move(d)), null(dom->heap().memory() == SymmetricHeap::Memory::Host),
        es(dom->args().elemSize) {}

  const void *my_grid() const { return dom->grid(); }

  // pack region d of my grid into a dense buffer
  TzbBox pack_box(const Dim3<int> &d, void *dst) const {
    return make_box(*dom, my_grid(), Side::Grid, dom->send_offset(d), dst, Side::Dense, {},
                    dom->region_extent(d));
  }
  // unpack a dense buffer into my ghost cells on side e
  TzbBox unpack_box(const Dim3<int> &e, const void *src) const {
    return make_box(*dom, src, Side::Dense, {}, dom->grid(), Side::Grid, dom->recv_offset(e),
                    dom->region_extent(e));
  }
  // my send region d -> neighbour's ghost cells on side -d.  Ghost cells are not double
  // buffered, so the put may only start once the neighbour has said that it is done with what
  // they held (it says so at the start of ITS exchange of this epoch, in the box that goes the
  // other way); before waiting for that, this box grants the same permission to the neighbour.
  TzbBox direct_box(const Dim3<int> &d) const {
    TzbBox b = make_box(*dom, my_grid(), Side::Grid, dom->send_offset(d), dom->grid_of(dom->neighbor(d)),
                        Side::Grid, dom->recv_offset(-d), dom->region_extent(d));
    b.dstRemote = dom->neighbor(d) != dom->rank();
    // TZB_NO_READY=1: measurement knob (what does the handshake cost?) - not for production,
    // see tests/test_host_execution.py::test_consecutive_exchanges_without_any_barrier_between_them
    static const bool handshake = !(std::getenv("TZB_NO_READY") && std::atoi(std::getenv("TZB_NO_READY")) != 0);
    if (handshake) {
      b.ready = dom->ready_flag(dom->neighbor(d), -d);
      b.wait = dom->ready_flag(dom->rank(), d);
    }
    return b;
  }
  // pack + put into the neighbour's staging slot (one box per parity)
  TzbBox packput_box(const Dim3<int> &d, int parity, Style style) const {
    const int peer = dom->neighbor(d);
    TzbBox b = pack_box(d, dom->recv_staging(peer, parity, -d));
    b.dstRemote = peer != dom->rank();
    b.signal = dom->flag(peer, style, -d);
    return b;
  }
  TzbBox waitunpack_box(const Dim3<int> &e, int parity, Style style) const {
    TzbBox b = unpack_box(e, dom->recv_staging(dom->rank(), parity, e));
    b.wait = dom->flag(dom->rank(), style, e);
    return b;
  }

  // flag channels (heap offsets) of a box list, for the FlagSynchronizer
  void annotate(BoxCopyOp &op, const std::vector<TzbBox> &boxes) const {
    std::vector<uint64_t> pub, wait;
    bool first = false;
    for (const TzbBox &bx : boxes) {
      if (bx.signal) pub.push_back(uint64_t(dom->heap().offset_of(bx.signal)));
      if (bx.ready) pub.push_back(uint64_t(dom->heap().offset_of(bx.ready))), first = true;
      if (bx.wait) wait.push_back(uint64_t(dom->heap().offset_of(bx.wait)));
    }
    op.set_flag_channels(pub, wait, first);
  }
  std::shared_ptr<BoxCopyOp> op(const std::string &name, const std::vector<TzbBox> &even,
                                const std::vector<TzbBox> &odd, bool ordered = false) const {
    auto o = std::make_shared<BoxCopyOp>(name, even, odd, es, null, ordered);
    annotate(*o, even); // both parities use the same flag words
    return o;
  }
  std::shared_ptr<BoxCopyOp> op(const std::string &name, const std::vector<TzbBox> &boxes) const {
    return op(name, boxes, boxes);
  }
};

void connect(Graph<OpBase> &g, const std::vector<OpPtr> &from, const OpPtr &to) {
  for (const auto &p : from) g.add_edge(p, to);
}
void connect(Graph<OpBase> &g, const OpPtr &from, const std::vector<OpPtr> &to) {
  for (const auto &s : to) g.add_edge(from, s);
}

} // namespace

void add_to_graph(Graph<OpBase> &g, const std::shared_ptr<Domain> &dom,
                  const std::vector<OpPtr> &preds, const std::vector<OpPtr> &succs, Style style) {
  const Builder b(dom);
  const auto &dirs = dom->dirs();

  if (style == Style::Reference) {
    auto tr = std::make_shared<RefTransport>(b.null);
    OpPtr waitSends = std::make_shared<RefWaitSends>(tr);
    std::vector<OpPtr> posts, waits;
    for (const auto &d : dirs) {
      const std::string sfx = dir_suffix(d);
      OpPtr pack = b.op("he_pack_" + sfx, {b.pack_box(d, dom->send_staging(dom->rank(), d))});
      OpPtr isend = std::make_shared<RefIsend>("he_isend_" + sfx, dom, tr, d);
      OpPtr irecv = std::make_shared<NoOp>("he_irecv_" + sfx);
      OpPtr wait = std::make_shared<RefWaitRecv>("he_waitrecv_" + sfx, dom, tr, d);
      OpPtr unpack = b.op("he_unpack_" + sfx,
                          {b.unpack_box(d, dom->recv_staging(dom->rank(), 0, d))},
                          {b.unpack_box(d, dom->recv_staging(dom->rank(), 1, d))});
      connect(g, preds, pack);
      g.add_edge(pack, isend);
      g.add_edge(isend, waitSends);
      connect(g, preds, irecv);
      g.add_edge(irecv, wait);
      g.add_edge(wait, unpack);
      connect(g, unpack, succs);
      g.add_edge(irecv, waitSends);
      posts.push_back(isend);
      posts.push_back(irecv);
      waits.push_back(wait);
    }
    connect(g, waitSends, succs);
    // all posts before any wait
    for (const auto &p : posts)
      for (const auto &w : waits) g.add_edge(p, w);
    return;
  }

  if (style == Style::Staged) {
    std::vector<OpPtr> puts, waits;
    for (const auto &d : dirs) {
      const std::string sfx = dir_suffix(d);
      puts.push_back(b.op("he_packput_" + sfx, {b.packput_box(d, 0, style)}, {b.packput_box(d, 1, style)}));
      waits.push_back(b.op("he_waitunpack_" + sfx, {b.waitunpack_box(d, 0, style)},
                           {b.waitunpack_box(d, 1, style)}));
    }
    for (const auto &p : puts) {
      connect(g, preds, p);
      for (const auto &w : waits) g.add_edge(p, w);
    }
    for (const auto &w : waits) connect(g, w, succs);
    return;
  }

  if (style == Style::Direct) {
    auto waitOp = std::make_shared<WaitFlagsOp>(
        "he_wait_puts", dom->flags_base(dom->rank(), style), int(dirs.size()), b.null);
    waitOp->set_channels(dom->heap());
    OpPtr wait = waitOp;
    // one put op per direction: the permissions cannot travel inside the puts (the put of
    // direction d would wait for the neighbour's put of direction -d, which every rank may have
    // queued BEHIND it), so a tiny op grants all of them up front
    std::vector<uint64_t *> ready;
    for (const auto &d : dirs) ready.push_back(dom->ready_flag(dom->neighbor(d), -d));
    OpPtr readyOp = std::make_shared<SignalFlagsOp>("he_ready", ready, dom->heap(), b.null);
    connect(g, preds, readyOp);
    for (const auto &d : dirs) {
      TzbBox box = b.direct_box(d);
      box.ready = nullptr;
      box.signal = dom->flag(dom->neighbor(d), style, -d);
      OpPtr put = b.op("he_put_" + dir_suffix(d), {box});
      g.add_edge(readyOp, put);
      g.add_edge(put, wait);
    }
    connect(g, wait, succs);
    return;
  }

  if (style == Style::FusedAll) {
    std::vector<TzbBox> boxes;
    for (const auto &d : dirs) {
      TzbBox box = b.direct_box(d);
      box.signal = dom->flag(dom->neighbor(d), style, -d);
      boxes.push_back(box);
    }
    OpPtr put = b.op("he_put_all", boxes);
    auto waitOp = std::make_shared<WaitFlagsOp>(
        "he_finish_all", dom->flags_base(dom->rank(), style), int(dirs.size()), b.null);
    waitOp->set_channels(dom->heap());
    OpPtr wait = waitOp;
    connect(g, preds, put);
    g.add_edge(put, wait);
    connect(g, wait, succs);
    return;
  }

  if (style == Style::FusedStaged) {
    std::vector<TzbBox> put[2], fin[2];
    for (int p = 0; p < 2; ++p)
      for (const auto &d : dirs) {
        put[p].push_back(b.packput_box(d, p, style));
        fin[p].push_back(b.waitunpack_box(d, p, style));
      }
    OpPtr putOp = b.op("he_packput_all", put[0], put[1]);
    OpPtr finOp = b.op("he_waitunpack_all", fin[0], fin[1]);
    connect(g, preds, putOp);
    g.add_edge(putOp, finOp);
    connect(g, finOp, succs);
    return;
  }
  if (style == Style::FusedGroups) {
    // group 0: every direction with an x component (thin rows), 1: y faces and yz edges, 2: z faces
    auto group_of = [](const Dim3<int> &d) { return d.x != 0 ? 0 : (d.y != 0 ? 1 : 2); };
    const char *gname[3] = {"x", "y", "z"};
    for (int grp = 0; grp < 3; ++grp) {
      std::vector<TzbBox> put[2], fin[2];
      for (int p = 0; p < 2; ++p)
        for (const auto &d : dirs) {
          if (group_of(d) != grp) continue;
          put[p].push_back(b.packput_box(d, p, style));
          fin[p].push_back(b.waitunpack_box(d, p, style));
        }
      if (put[0].empty()) continue;
      // ordered plans: these kernels run beside each other and must share the SMs
      OpPtr putOp = b.op(std::string("he_packput_") + gname[grp], put[0], put[1], true);
      OpPtr finOp = b.op(std::string("he_waitunpack_") + gname[grp], fin[0], fin[1], true);
      connect(g, preds, putOp);
      // only its own puts precede a group's unpack: ordered plans leave half of every SM free
      // (see tzb_boxcopy_launch), so a later put can always run beside a spinning unpack
      g.add_edge(putOp, finOp);
      connect(g, finOp, succs);
    }
    return;
  }
  if (style == Style::FusedMixed) {
    std::vector<TzbBox> thinPut[2], thinFin[2], thick;
    std::vector<const uint64_t *> thickFlags;
    for (const auto &d : dirs) {
      if (d.x != 0) {
        for (int p = 0; p < 2; ++p) {
          thinPut[p].push_back(b.packput_box(d, p, style));
          thinFin[p].push_back(b.waitunpack_box(d, p, style));
        }
      } else {
        TzbBox box = b.direct_box(d);
        box.signal = dom->flag(dom->neighbor(d), style, -d);
        thick.push_back(box);
      }
    }
    // completion of the direct puts: a box-copy plan of empty boxes that only wait
    std::vector<TzbBox> thickWait;
    for (const auto &d : dirs) {
      if (d.x != 0) continue;
      TzbBox w;
      std::memset(&w, 0, sizeof(w));
      w.wait = dom->flag(dom->rank(), style, d);
      thickWait.push_back(w);
    }
    // The permissions for the direct puts are granted by BOTH put kernels (the same stores
    // twice): whichever of them the schedule runs first gets them out at the very start of the
    // exchange, so a thick put that runs second finds them waiting instead of stalling until
    // every partner has reached its own thick put.
    for (const TzbBox &t : thick) {
      if (!t.ready) continue;
      TzbBox grant;
      std::memset(&grant, 0, sizeof(grant));
      grant.ready = t.ready;
      thinPut[0].push_back(grant);
      thinPut[1].push_back(grant);
    }
    OpPtr putThin = b.op("he_packput_thin", thinPut[0], thinPut[1]);
    OpPtr finThin = b.op("he_waitunpack_thin", thinFin[0], thinFin[1]);
    connect(g, preds, putThin);
    g.add_edge(putThin, finThin);
    connect(g, finThin, succs);
    if (!thick.empty()) {
      OpPtr putThick = b.op("he_put_thick", thick);
      OpPtr finThick = b.op("he_wait_thick", thickWait);
      connect(g, preds, putThick);
      g.add_edge(putThick, finThick);
      connect(g, finThick, succs);
    }
    return;
  }
  if (style == Style::FusedPipelined) {
    // thin regions first: their unpack is the slow part, so it gets the longest overlap
    std::vector<Dim3<int>> order;
    for (const auto &d : dirs)
      if (d.x != 0) order.push_back(d);
    for (const auto &d : dirs)
      if (d.x == 0) order.push_back(d);
    std::vector<TzbBox> put[2], fin[2];
    for (int p = 0; p < 2; ++p)
      for (const auto &d : order) {
        put[p].push_back(b.packput_box(d, p, style));
        // the neighbour on side e sends the region it calls -e: follow the PUT order
        fin[p].push_back(b.waitunpack_box(-d, p, style));
      }
    auto putOp = b.op("he_packput_pipe", put[0], put[1], true);
    auto finOp = b.op("he_waitunpack_pipe", fin[0], fin[1], true);
    // The unpack may run beside the put only when every flag it spins on is published by
    // ANOTHER GPU.  Where a neighbour is this rank itself (a rank-grid dimension of 1) the
    // flags come from this GPU's own put kernel, and nothing guarantees that kernel a place
    // on the SMs next to a spinning unpack: keep the real event edge there.
    bool selfNeighbour = false;
    for (const auto &d : dirs) selfNeighbour = selfNeighbour || dom->neighbor(d) == dom->rank();
    finOp->set_issue_order_only(!selfNeighbour);
    connect(g, preds, OpPtr(putOp));
    g.add_edge(OpPtr(putOp), OpPtr(finOp)); // issue order only: the kernels may overlap
    connect(g, OpPtr(putOp), succs);
    connect(g, OpPtr(finOp), succs);
    return;
  }
  if (style == Style::FusedPull) {
    // 1. tell every neighbour that my interior may be read for this epoch
    std::vector<uint64_t *> ready;
    for (const auto &d : dirs) ready.push_back(dom->flag(dom->neighbor(d), style, -d));
    OpPtr readyOp = std::make_shared<SignalFlagsOp>("he_ready", ready, dom->heap(), b.null);
    // 2. load every region from the neighbour's grid into my ghost cells; acknowledge
    std::vector<TzbBox> boxes;
    for (const auto &e : dirs) {
      const int nb = dom->neighbor(e);
      // the neighbour on my side e: its interior cells next to ITS face -e are my ghost cells e
      TzbBox box = make_box(*dom, dom->grid_of(nb), Side::Grid, dom->send_offset(-e), dom->grid(),
                            Side::Grid, dom->recv_offset(e), dom->region_extent(e));
      box.wait = dom->flag(dom->rank(), style, e);
      box.signal = dom->ack_flag(nb, -e);
      boxes.push_back(box);
    }
    OpPtr pull = b.op("he_pull_all", boxes);
    // 3. my interior may change again once every neighbour has read it
    auto waitOp = std::make_shared<WaitFlagsOp>("he_wait_acks", dom->ack_flags_base(dom->rank()),
                                                int(dirs.size()), b.null);
    waitOp->set_channels(dom->heap());
    OpPtr wait = waitOp;
    connect(g, preds, readyOp);
    g.add_edge(readyOp, pull);
    g.add_edge(pull, wait);
    connect(g, wait, succs);
    return;
  }
  THROW_RUNTIME("unknown halo style");
}

namespace {
class HaloCompound : public CompoundOp {
  std::string name_;
  Graph<OpBase> graph_;

public:
  HaloCompound(const std::string &name, const std::shared_ptr<Domain> &dom, Style style)
      : name_(name) {
    add_to_graph(graph_, dom, {graph_.start()}, {graph_.finish()}, style);
  }
  std::string name() const override { return name_; }
  Graph<OpBase> graph() const override { return graph_; }
  CLONE_DEF(HaloCompound)
};
} // namespace

std::shared_ptr<CompoundOp> make_compound(const std::shared_ptr<Domain> &dom, Style style) {
  return std::make_shared<HaloCompound>(std::string("halo_") + style_name(style), dom, style);
}

std::shared_ptr<ChoiceOp> make_choice(const std::shared_ptr<Domain> &dom,
                                      const std::vector<Style> &styles) {
  std::vector<OpPtr> alts;
  for (Style s : styles) alts.push_back(make_compound(dom, s));
  return std::make_shared<ListChoiceOp>("halo_exchange", alts);
}

} // namespace HaloExchange
} // namespace tenzing
