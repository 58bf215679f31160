// Python bindings for the workload libraries (halo exchange, SpMV) and the raw
// kernel entry points (used by the numerics tests).
#include <pybind11/functional.h>
#include <pybind11/pybind11.h>
#include "op_cast.hpp"
#include <pybind11/stl.h>

#include "tenzing/halo/halo_exchange.hpp"
#include "tenzing/spmv/array.hpp"
#include "tenzing/tenzing.hpp"
#include "tenzing_b200.h"

namespace py = pybind11;
using namespace tenzing;

void bind_spmv(py::module_ &m); // bindings_spmv.cpp

namespace {

void check(int err, const char *what) {
  if (err != 0)
    throw std::runtime_error(std::string(what) + " failed: " +
                             cudaGetErrorString(static_cast<cudaError_t>(err)));
}
template <typename T> T *ptr(uintptr_t p) { return reinterpret_cast<T *>(p); }

HaloExchange::Style style_from(const std::string &s) {
  using HaloExchange::Style;
  for (Style st : {Style::Reference, Style::Staged, Style::Direct, Style::FusedAll, Style::FusedStaged, Style::FusedGroups, Style::FusedMixed, Style::FusedPipelined, Style::FusedPull})
    if (s == HaloExchange::style_name(st)) return st;
  throw std::runtime_error("unknown halo style '" + s + "'");
}

} // namespace

void bind_workloads(py::module_ &m) {
  // ---- raw kernels (device pointers as ints, stream as int) ----
  py::module_ k = m.def_submodule("kernels");
  k.def("spmv_csr_f64", [](int rows, uintptr_t rp, uintptr_t ci, uintptr_t v, uintptr_t x, uintptr_t y,
                           bool acc, uintptr_t s) {
    check(tzb_spmv_csr_f64(rows, ptr<int>(rp), ptr<int>(ci), ptr<double>(v), ptr<double>(x),
                           ptr<double>(y), acc, ptr<void>(s)), "spmv_csr_f64");
  });
  k.def("spmv_csr_f32", [](int rows, uintptr_t rp, uintptr_t ci, uintptr_t v, uintptr_t x, uintptr_t y,
                           bool acc, uintptr_t s) {
    check(tzb_spmv_csr_f32(rows, ptr<int>(rp), ptr<int>(ci), ptr<float>(v), ptr<float>(x),
                           ptr<float>(y), acc, ptr<void>(s)), "spmv_csr_f32");
  });
  k.def("spmv_csr_wait_f64", [](int rows, uintptr_t rp, uintptr_t ci, uintptr_t v, uintptr_t x,
                                uintptr_t y, bool acc, uintptr_t flags, int nFlags, uint64_t epoch,
                                uintptr_t s) {
    check(tzb_spmv_csr_wait_f64(rows, ptr<int>(rp), ptr<int>(ci), ptr<double>(v), ptr<double>(x),
                                ptr<double>(y), acc, ptr<uint64_t>(flags), nFlags, epoch,
                                ptr<void>(s)), "spmv_csr_wait_f64");
  });
  k.def("gather_f64", [](int n, uintptr_t idx, uintptr_t src, uintptr_t dst, uintptr_t s) {
    check(tzb_gather_f64(n, ptr<int>(idx), ptr<double>(src), ptr<double>(dst), ptr<void>(s)), "gather_f64");
  });
  k.def("gather_f32", [](int n, uintptr_t idx, uintptr_t src, uintptr_t dst, uintptr_t s) {
    check(tzb_gather_f32(n, ptr<int>(idx), ptr<float>(src), ptr<float>(dst), ptr<void>(s)), "gather_f32");
  });
  k.def("vector_add_f64", [](int n, uintptr_t a, uintptr_t y, uintptr_t s) {
    check(tzb_vector_add_f64(n, ptr<double>(a), ptr<double>(y), ptr<void>(s)), "vector_add_f64");
  });
  k.def("gather_put_f64", [](const std::vector<int> &displ, const std::vector<uintptr_t> &dst,
                             const std::vector<uintptr_t> &flags, uintptr_t idx, uintptr_t src,
                             uint64_t epoch, uintptr_t counter, uintptr_t s) {
    TzbGatherPutArgs a;
    std::memset(&a, 0, sizeof(a));
    a.nPeers = int(dst.size());
    if (a.nPeers > 16 || displ.size() != dst.size() + 1 || flags.size() != dst.size())
      throw std::runtime_error("gather_put: bad argument sizes");
    for (size_t i = 0; i < displ.size(); ++i) a.displ[i] = displ[i];
    for (size_t i = 0; i < dst.size(); ++i) {
      a.peerDst[i] = ptr<void>(dst[i]);
      a.peerFlag[i] = ptr<uint64_t>(flags[i]);
    }
    check(tzb_gather_put_f64(&a, ptr<int>(idx), ptr<double>(src), epoch, ptr<uint32_t>(counter),
                             ptr<void>(s)), "gather_put_f64");
  });
  k.def("multicast_put", [](int n, uintptr_t src, uintptr_t mcDst, uintptr_t mcFlag, uint64_t epoch,
                            uintptr_t counter, bool f64, uintptr_t s) {
    if (f64)
      check(tzb_multicast_put_f64(n, ptr<double>(src), ptr<double>(mcDst), ptr<uint64_t>(mcFlag), epoch,
                                  ptr<uint32_t>(counter), ptr<void>(s)), "multicast_put_f64");
    else
      check(tzb_multicast_put_f32(n, ptr<float>(src), ptr<float>(mcDst), ptr<uint64_t>(mcFlag), epoch,
                                  ptr<uint32_t>(counter), ptr<void>(s)), "multicast_put_f32");
  });
  k.def("spmv_csr_wait", [](int rows, uintptr_t rp, uintptr_t ci, uintptr_t val, uintptr_t x, uintptr_t y,
                            bool accumulate, uintptr_t flags, int nFlags, uint64_t epoch, bool f64, uintptr_t s) {
    if (f64)
      check(tzb_spmv_csr_wait_f64(rows, ptr<int>(rp), ptr<int>(ci), ptr<double>(val), ptr<double>(x),
                                  ptr<double>(y), accumulate, ptr<uint64_t>(flags), nFlags, epoch, ptr<void>(s)),
            "spmv_csr_wait_f64");
    else
      check(tzb_spmv_csr_wait_f32(rows, ptr<int>(rp), ptr<int>(ci), ptr<float>(val), ptr<float>(x),
                                  ptr<float>(y), accumulate, ptr<uint64_t>(flags), nFlags, epoch, ptr<void>(s)),
            "spmv_csr_wait_f32");
  });
  k.def("copy_bytes", [](uintptr_t dst, uintptr_t src, size_t bytes, uintptr_t s) {
    check(tzb_copy_bytes(ptr<void>(dst), ptr<void>(src), bytes, ptr<void>(s)), "copy_bytes");
  });
  k.def("flags_wait", [](uintptr_t flags, int n, uint64_t epoch, uintptr_t s) {
    check(tzb_flags_wait(ptr<uint64_t>(flags), n, epoch, ptr<void>(s)), "flags_wait");
  });
  k.def("flags_signal", [](const std::vector<uintptr_t> &flags, uint64_t epoch, uintptr_t s) {
    TzbFlagList l;
    l.n = int(flags.size());
    if (l.n > 32) throw std::runtime_error("flags_signal: at most 32 flags");
    for (int i = 0; i < l.n; ++i) l.flag[i] = ptr<uint64_t>(flags[size_t(i)]);
    check(tzb_flags_signal(&l, epoch, ptr<void>(s)), "flags_signal");
  });
  k.def("l2_flush", [](uintptr_t buf, size_t bytes, unsigned value, uintptr_t s) {
    check(tzb_l2_flush(ptr<void>(buf), bytes, value, ptr<void>(s)), "l2_flush");
  });
  k.def("launch_count", [] { return tzb_launch_count(); });
  k.def("set_wait_timeout_ms", [](unsigned ms) { check(tzb_set_wait_timeout_ms(ms), "set_wait_timeout_ms"); });
  // cudaLimitMaxL2FetchGranularity hint: returns the value the device reports afterwards
  k.def("set_l2_fetch_granularity", [](size_t bytes) {
    cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, bytes);
    if (e != cudaSuccess) (void)cudaGetLastError();
    size_t v = 0;
    cudaDeviceGetLimit(&v, cudaLimitMaxL2FetchGranularity);
    return std::make_pair(int(e), v);
  });

  // ---- symmetric heap ----
  py::class_<SymmetricHeap, std::shared_ptr<SymmetricHeap>>(m, "SymmetricHeap")
      // Collective (other ranks may be threads of this process), so the GIL is released - but
      // only around the C++ constructor: a call_guard on py::init would also cover pybind's
      // registration of the new instance, which must happen under the GIL.
      .def(py::init([](std::shared_ptr<Communicator> comm, size_t bytes, bool host) {
             py::gil_scoped_release rel;
             return std::make_shared<SymmetricHeap>(std::move(comm), bytes,
                                                    host ? SymmetricHeap::Memory::Host
                                                         : SymmetricHeap::Memory::Device);
           }),
           py::arg("comm"), py::arg("bytes"), py::arg("host") = false)
      .def("rank", &SymmetricHeap::rank)
      .def("size", &SymmetricHeap::size)
      .def("bytes", &SymmetricHeap::bytes)
      .def("local_ptr", [](const SymmetricHeap &h) { return reinterpret_cast<uintptr_t>(h.local()); })
      .def("peer_ptr", [](const SymmetricHeap &h, int r) { return reinterpret_cast<uintptr_t>(h.peer(r)); })
      .def("alloc", &SymmetricHeap::alloc, py::arg("bytes"), py::arg("align") = 256)
      .def("zero", &SymmetricHeap::zero, py::arg("offset") = 0, py::arg("bytes") = size_t(-1));

  // ---- host-posted P2P ops (the MPI-shaped op library) ----
  py::class_<P2PTransport, std::shared_ptr<P2PTransport>>(m, "P2PTransport")
      .def(py::init<bool>(), py::arg("null") = false);
  py::class_<Request, RequestPtr>(m, "Request").def(py::init<>()).def("wait", &Request::wait);
  m.def("Isend", [](uintptr_t buf, size_t bytes, uintptr_t peerBuf, uintptr_t peerFlag, RequestPtr req,
                    std::shared_ptr<P2PTransport> tr, const std::string &name) {
    Isend::Args a;
    a.buf = ptr<void>(buf), a.bytes = bytes, a.peerBuf = ptr<void>(peerBuf);
    a.peerFlag = ptr<uint64_t>(peerFlag), a.request = std::move(req);
    return std::static_pointer_cast<BoundOp>(std::make_shared<Isend>(a, std::move(tr), name));
  }, py::arg("buf"), py::arg("bytes"), py::arg("peer_buf"), py::arg("peer_flag"), py::arg("request"),
     py::arg("transport"), py::arg("name"));
  m.def("Irecv", [](uintptr_t flag, RequestPtr req, std::shared_ptr<P2PTransport> tr, const std::string &name) {
    Irecv::Args a;
    a.flag = ptr<uint64_t>(flag), a.request = std::move(req);
    return std::static_pointer_cast<BoundOp>(std::make_shared<Irecv>(a, std::move(tr), name));
  }, py::arg("flag"), py::arg("request"), py::arg("transport"), py::arg("name"));
  // variants that own their request (reference: OwningIsend / OwningIrecv); returns (op, request)
  m.def("OwningIsend", [](uintptr_t buf, size_t bytes, uintptr_t peerBuf, uintptr_t peerFlag,
                          std::shared_ptr<P2PTransport> tr, const std::string &name) {
    Isend::Args a;
    a.buf = ptr<void>(buf), a.bytes = bytes, a.peerBuf = ptr<void>(peerBuf);
    a.peerFlag = ptr<uint64_t>(peerFlag);
    auto op = std::make_shared<OwningIsend>(a, std::move(tr), name);
    return std::make_pair(std::static_pointer_cast<BoundOp>(op), op->request());
  }, py::arg("buf"), py::arg("bytes"), py::arg("peer_buf"), py::arg("peer_flag"), py::arg("transport"),
     py::arg("name"));
  m.def("OwningIrecv", [](uintptr_t flag, std::shared_ptr<P2PTransport> tr, const std::string &name) {
    Irecv::Args a;
    a.flag = ptr<uint64_t>(flag);
    auto op = std::make_shared<OwningIrecv>(a, std::move(tr), name);
    return std::make_pair(std::static_pointer_cast<BoundOp>(op), op->request());
  }, py::arg("flag"), py::arg("transport"), py::arg("name"));
  m.def("Wait", [](RequestPtr req, const std::string &name) {
    return std::static_pointer_cast<BoundOp>(std::make_shared<Wait>(std::move(req), name));
  });
  m.def("MultiWait", [](const std::vector<RequestPtr> &reqs, const std::string &name) {
    auto w = std::make_shared<MultiWait>(name);
    for (const auto &r : reqs) w->add_request(r);
    return std::static_pointer_cast<BoundOp>(w);
  });
  m.def("OwningWaitall", [](const std::vector<RequestPtr> &reqs, const std::string &name) {
    auto w = std::make_shared<OwningWaitall>(name);
    for (const auto &r : reqs) w->add_request(r);
    return std::static_pointer_cast<BoundOp>(w);
  });
  m.def("Ialltoallv", [](uintptr_t sendbuf, const std::vector<size_t> &offs, const std::vector<size_t> &bytes,
                         const std::vector<uintptr_t> &peerBufs, const std::vector<uintptr_t> &peerFlags,
                         RequestPtr req, std::shared_ptr<P2PTransport> tr, const std::string &name) {
    Ialltoallv::Args a;
    a.sendbuf = ptr<void>(sendbuf), a.sendOffsets = offs, a.sendBytes = bytes, a.request = std::move(req);
    for (uintptr_t p : peerBufs) a.peerBufs.push_back(ptr<void>(p));
    for (uintptr_t p : peerFlags) a.peerFlags.push_back(ptr<uint64_t>(p));
    return std::static_pointer_cast<BoundOp>(std::make_shared<Ialltoallv>(a, std::move(tr), name));
  });

  // ---- NUMA-aware pinned staging ----
  py::class_<PinnedPool>(m, "PinnedPool")
      .def(py::init<int>(), py::arg("cuda_device") = -1)
      .def("numa_node", &PinnedPool::numa_node)
      .def("bytes_allocated", &PinnedPool::bytes_allocated)
      .def("get", [](PinnedPool &p, size_t bytes) { return reinterpret_cast<uintptr_t>(p.get(bytes)); })
      .def("put", [](PinnedPool &p, uintptr_t ptr, size_t bytes) { p.put(reinterpret_cast<void *>(ptr), bytes); });

  // ---- halo exchange ----
  py::module_ h = m.def_submodule("halo");
  using HaloExchange::Args;
  using HaloExchange::Domain;
  py::class_<Args>(h, "Args")
      .def(py::init<>())
      .def_readwrite("nQ", &Args::nQ)
      .def_readwrite("nX", &Args::nX)
      .def_readwrite("nY", &Args::nY)
      .def_readwrite("nZ", &Args::nZ)
      .def_readwrite("nGhost", &Args::nGhost)
      .def_readwrite("pitch", &Args::pitch)
      .def_readwrite("neighbors", &Args::neighbors)
      .def_readwrite("elemSize", &Args::elemSize)
      .def_property("storageOrder",
                    [](const Args &a) { return a.storageOrder == HaloExchange::StorageOrder::XYZQ ? "XYZQ" : "QXYZ"; },
                    [](Args &a, const std::string &s) {
                      if (s == "XYZQ") a.storageOrder = HaloExchange::StorageOrder::XYZQ;
                      else if (s == "QXYZ") a.storageOrder = HaloExchange::StorageOrder::QXYZ;
                      else throw std::runtime_error("storageOrder must be XYZQ or QXYZ");
                    })
      .def_property("rankGrid",
                    [](const Args &a) { return std::vector<int>{a.rankGrid.x, a.rankGrid.y, a.rankGrid.z}; },
                    [](Args &a, const std::vector<int> &g) { a.rankGrid = Dim3<int>(g.at(0), g.at(1), g.at(2)); });
  py::class_<Domain, std::shared_ptr<Domain>>(h, "Domain")
      .def(py::init([](std::shared_ptr<Communicator> comm, const Args &args, bool host) {
             py::gil_scoped_release rel; // collective; see SymmetricHeap above
             return std::make_shared<Domain>(std::move(comm), args,
                                             host ? SymmetricHeap::Memory::Host
                                                  : SymmetricHeap::Memory::Device);
           }),
           py::arg("comm"), py::arg("args"), py::arg("host") = false)
      .def("rank", &Domain::rank)
      .def("size", &Domain::size)
      .def("extent", [](const Domain &d) { auto e = d.extent(); return std::vector<int>{e.x, e.y, e.z}; })
      .def("row_elems", &Domain::row_elems)
      .def("x_origin", &Domain::x_origin)
      .def("grid_elems", &Domain::grid_elems)
      .def("grid_bytes", &Domain::grid_bytes)
      .def("grid_ptr", [](const Domain &d) { return reinterpret_cast<uintptr_t>(d.grid()); })
      .def("strides", [](const Domain &d) { auto s = d.strides(); return std::vector<int64_t>(s.begin(), s.end()); })
      .def("index", &Domain::index)
      .def("bytes_per_exchange", &Domain::bytes_per_exchange)
      .def("rank_coord", [](const Domain &d) { auto c = d.args().rankToCoord(d.rank()); return std::vector<int>{c.x, c.y, c.z}; })
      .def("rank_grid", [](const Domain &d) { auto c = d.args().rankGrid; return std::vector<int>{c.x, c.y, c.z}; })
      .def("neighbor", [](const Domain &d, int dx, int dy, int dz) { return d.neighbor(Dim3<int>(dx, dy, dz)); })
      // staging buffers of THIS rank (dense, one slot per direction): what the reference-style
      // pack ops fill and unpack ops read; used by transports written in Python (torch.distributed)
      .def("region_elems", [](const Domain &d, int dx, int dy, int dz) { return d.region_elems(Dim3<int>(dx, dy, dz)); })
      .def("send_staging_ptr", [](const Domain &d, int dx, int dy, int dz) {
        return reinterpret_cast<uintptr_t>(d.send_staging(d.rank(), Dim3<int>(dx, dy, dz)));
      })
      .def("recv_staging_ptr", [](const Domain &d, int parity, int dx, int dy, int dz) {
        return reinterpret_cast<uintptr_t>(d.recv_staging(d.rank(), parity, Dim3<int>(dx, dy, dz)));
      })
      .def("dirs", [](const Domain &d) {
        std::vector<std::vector<int>> r;
        for (const auto &x : d.dirs()) r.push_back({x.x, x.y, x.z});
        return r;
      });
  h.def("rank_grid", [](int size) { auto g = HaloExchange::rank_grid(size); return std::vector<int>{g.x, g.y, g.z}; });
  h.def("dir_to_tag", [](int dx, int dy, int dz) { return HaloExchange::dir_to_tag(Dim3<int>(dx, dy, dz)); });
  h.def("styles", [] {
    return std::vector<std::string>{"reference", "staged", "direct", "fused_all", "fused_staged", "fused_groups", "fused_mixed", "fused_pipelined", "fused_pull"};
  });
  h.def("add_to_graph", [](Graph<OpBase> &g, const std::shared_ptr<Domain> &dom,
                           const std::vector<OpPtr> &preds, const std::vector<OpPtr> &succs,
                           const std::string &style) {
    HaloExchange::add_to_graph(g, dom, preds, succs, style_from(style));
  }, py::arg("graph"), py::arg("domain"), py::arg("preds"), py::arg("succs"), py::arg("style") = "staged");
  h.def("make_compound", [](const std::shared_ptr<Domain> &dom, const std::string &style) {
    return HaloExchange::make_compound(dom, style_from(style));
  });
  h.def("make_choice", [](const std::shared_ptr<Domain> &dom, const std::vector<std::string> &styles) {
    std::vector<HaloExchange::Style> s;
    for (const auto &x : styles) s.push_back(style_from(x));
    return HaloExchange::make_choice(dom, s);
  }, py::arg("domain"), py::arg("styles") = std::vector<std::string>{"reference", "staged", "direct", "fused_all", "fused_staged", "fused_groups", "fused_mixed", "fused_pipelined", "fused_pull"});

  bind_spmv(m);
}
