// pybind11 front end: tenzing_b200._C
//
// The reference has no Python bindings (README roadmap, README.md:56-58); this
// module exposes the C++ framework 1:1 - ops, graph, sequence, state, decisions,
// synchronizer, benchmarkers, both solvers - and lets Python supply ops
// (callables), a communicator (torch.distributed) and a benchmarker.
#include <pybind11/functional.h>
#include <pybind11/pybind11.h>
#include "op_cast.hpp"
#include <pybind11/stl.h>

#include "tenzing/cuda/device_barrier.hpp"
#include "tenzing/cuda/graph_replay.hpp"
#include "tenzing/host_exec.hpp"
#include "tenzing/tenzing.hpp"

namespace py = pybind11;
using namespace tenzing;

void bind_workloads(py::module_ &m); // bindings_workloads.cpp

namespace {

// ---- Python-implemented pieces ----
class PyCpuOp : public CpuOp {
  std::string name_;
  py::object fn_;

public:
  PyCpuOp(const std::string &name, py::object fn) : name_(name), fn_(std::move(fn)) {}
  ~PyCpuOp() override {
    py::gil_scoped_acquire gil;
    fn_ = py::object();
  }
  std::string name() const override { return name_; }
  CLONE_DEF(PyCpuOp)
  void run(Platform &) override {
    py::gil_scoped_acquire gil;
    if (!fn_.is_none()) fn_();
  }
};

class PyGpuOp : public GpuOp {
  std::string name_;
  py::object fn_;

public:
  PyGpuOp(const std::string &name, py::object fn) : name_(name), fn_(std::move(fn)) {}
  ~PyGpuOp() override {
    py::gil_scoped_acquire gil;
    fn_ = py::object();
  }
  std::string name() const override { return name_; }
  CLONE_DEF(PyGpuOp)
  void run(cudaStream_t stream) override {
    py::gil_scoped_acquire gil;
    if (!fn_.is_none()) fn_(reinterpret_cast<uintptr_t>(stream));
  }
};

// forwards the control plane to a Python object (torch.distributed wrapper)
class PyComm : public Communicator {
  py::object obj_;
  int rank_, size_; // fixed for the life of the group; cached so late callers need no Python

public:
  explicit PyComm(py::object obj) : obj_(std::move(obj)) {
    rank_ = obj_.attr("rank")().cast<int>();
    size_ = obj_.attr("size")().cast<int>();
  }
  ~PyComm() override {
    py::gil_scoped_acquire gil;
    obj_ = py::object();
  }
  int rank() const override { return rank_; }
  int size() const override { return size_; }
  void barrier() override {
    py::gil_scoped_acquire gil;
    obj_.attr("barrier")();
  }
  void bcast(void *buf, size_t bytes, int root) override {
    py::gil_scoped_acquire gil;
    py::bytes in(static_cast<const char *>(buf), bytes);
    std::string out = obj_.attr("bcast_bytes")(in, root).cast<std::string>();
    if (out.size() != bytes) THROW_RUNTIME("PyComm.bcast_bytes returned the wrong length");
    std::memcpy(buf, out.data(), bytes);
  }
  void allreduce_max(double *vals, size_t n) override {
    py::gil_scoped_acquire gil;
    std::vector<double> v(vals, vals + n);
    std::vector<double> out = obj_.attr("allreduce_max")(v).cast<std::vector<double>>();
    if (out.size() != n) THROW_RUNTIME("PyComm.allreduce_max returned the wrong length");
    std::copy(out.begin(), out.end(), vals);
  }
  void allgather(const void *in, void *out, size_t bytesPerRank) override {
    py::gil_scoped_acquire gil;
    py::bytes b(static_cast<const char *>(in), bytesPerRank);
    std::vector<std::string> parts = obj_.attr("allgather_bytes")(b).cast<std::vector<std::string>>();
    for (size_t r = 0; r < parts.size(); ++r) {
      if (parts[r].size() != bytesPerRank) THROW_RUNTIME("PyComm.allgather_bytes: bad part size");
      std::memcpy(static_cast<char *>(out) + r * bytesPerRank, parts[r].data(), bytesPerRank);
    }
  }
  std::string kind() const override { return "python"; }
};

class PyBenchmarker : public Benchmarker {
  py::object fn_;

public:
  explicit PyBenchmarker(py::object fn) : fn_(std::move(fn)) {}
  ~PyBenchmarker() override {
    py::gil_scoped_acquire gil;
    fn_ = py::object();
  }
  Benchmark::Result benchmark(Sequence<BoundOp> &order, Platform &, const Benchmark::Opts &) override {
    py::gil_scoped_acquire gil;
    py::object r = fn_(order);
    if (py::isinstance<Benchmark::Result>(r)) return r.cast<Benchmark::Result>();
    const double t = r.cast<double>();
    Benchmark::Result res;
    res.pct01 = res.pct10 = res.pct50 = res.pct90 = res.pct99 = t;
    return res;
  }
};

std::vector<int> op_streams(const OpBase &op) {
  std::vector<int> r;
  if (const auto *hs = dynamic_cast<const HasStream *>(&op))
    for (const Stream &s : hs->get_streams()) r.push_back(s.id());
  return r;
}
std::vector<int> op_events(const OpBase &op) {
  std::vector<int> r;
  if (const auto *he = dynamic_cast<const HasEvent *>(&op))
    for (const Event &e : he->get_events()) r.push_back(e.id());
  return r;
}

using G = Graph<OpBase>;
using Seq = Sequence<BoundOp>;

// reproduce::Header -> dict (None if the line was not a header)
py::object header_to_py(const reproduce::Header &h) {
  if (!h.valid) return py::none();
  py::dict d;
  d["args"] = h.args;
  d["hash"] = h.hash;
  d["major"] = h.major;
  d["minor"] = h.minor;
  d["patch"] = h.patch;
  d["version"] = h.version();
  d["same_build"] = h.same_build();
  return d;
}

} // namespace

PYBIND11_MODULE(_C, m) {
  m.doc() = "tenzing-b200 native core";
  m.attr("__version__") = reproduce::version_string();

  // ---- L0 ----
  m.def("init", [] { tenzing::init(); });
  m.def("version_string", &reproduce::version_string);
  m.def("reproduce_header", &reproduce::header_json);
  m.def("parse_reproduce_header",
        [](const std::string &line) { return header_to_py(reproduce::parse_header(line)); });
  m.def("set_verbosity", &log::set_verbosity);
  m.def("set_host_wait_timeout", &host::set_wait_timeout_seconds, py::arg("seconds"),
        "seconds a host-executed flag wait may take before it raises (dead-lock detection)");
  m.def("set_log_rank", &log::set_rank);
  m.def("avg", [](const std::vector<double> &v) { return avg(v); });
  m.def("med", [](const std::vector<double> &v) { return med(v); });
  m.def("var", [](const std::vector<double> &v) { return var(v); });
  m.def("stddev", [](const std::vector<double> &v) { return stddev(v); });
  m.def("corr", [](const std::vector<double> &a, const std::vector<double> &b) { return corr(a, b); });
  m.def("prime_factors", [](long n) { return prime_factors<long>(n); });
  m.def("round_up", [](long x, long s) { return round_up<long>(x, s); });
  m.def("runs_test", &randomness::runs_test);
  m.def("runs_test_z", &randomness::runs_test_z);
  m.def("compound_test", &randomness::compound_test);
  m.def("bind_to_local_memory", &numa::bind_to_local_memory);
  m.def("numa_node_of_pci_device", &numa::node_of_pci_device);
  m.def("numa_cpus_of_node", &numa::cpus_of_node);
  m.def("counters", [] {
    py::dict d;
    const auto &mc = counters::mcts();
    d["mcts.SELECT_TIME"] = double(mc.SELECT_TIME);
    d["mcts.EXPAND_TIME"] = double(mc.EXPAND_TIME);
    d["mcts.ROLLOUT_TIME"] = double(mc.ROLLOUT_TIME);
    d["mcts.REDUNDANT_SYNC_TIME"] = double(mc.REDUNDANT_SYNC_TIME);
    d["mcts.RMAP_TIME"] = double(mc.RMAP_TIME);
    d["mcts.BENCHMARK_TIME"] = double(mc.BENCHMARK_TIME);
    d["mcts.BACKPROP_TIME"] = double(mc.BACKPROP_TIME);
    d["mcts.ITERS"] = uint64_t(mc.ITERS);
    const auto &dc = counters::dfs();
    d["dfs.ENUMERATE_TIME"] = double(dc.ENUMERATE_TIME);
    d["dfs.BENCHMARK_TIME"] = double(dc.BENCHMARK_TIME);
    d["dfs.STATES_VISITED"] = uint64_t(dc.STATES_VISITED);
    d["dfs.STATES_PRUNED"] = uint64_t(dc.STATES_PRUNED);
    return d;
  });
  m.def("reset_counters", &counters::reset);
  // parse + dump through the built-in JSON value (format tests)
  m.def("json_roundtrip", [](const std::string &text) { return Json::parse(text).dump(); });

  py::class_<Equivalence>(m, "Equivalence")
      .def("__bool__", [](const Equivalence &e) { return bool(e); })
      .def("__str__", &Equivalence::str)
      .def("stream_map",
           [](const Equivalence &e) {
             std::map<int, int> r;
             for (const auto &kv : e.streams().forward()) r[kv.first.id()] = kv.second.id();
             return r;
           })
      .def("event_map", [](const Equivalence &e) {
        std::map<int, int> r;
        for (const auto &kv : e.events().forward()) r[kv.first.id()] = kv.second.id();
        return r;
      });

  // ---- communicators ----
  py::class_<Communicator, std::shared_ptr<Communicator>>(m, "Communicator")
      .def("rank", &Communicator::rank)
      .def("size", &Communicator::size)
      .def("kind", &Communicator::kind)
      .def("barrier", &Communicator::barrier, py::call_guard<py::gil_scoped_release>())
      .def("bcast_string", &Communicator::bcast_string, py::call_guard<py::gil_scoped_release>())
      .def("bcast_bytes",
           [](Communicator &c, py::bytes b, int root) {
             std::string s = b;
             {
               py::gil_scoped_release rel;
               s = c.bcast_string(s, root);
             }
             return py::bytes(s);
           })
      .def("max_over_ranks", &Communicator::max_over_ranks, py::call_guard<py::gil_scoped_release>())
      .def("allgather_bytes", [](Communicator &c, py::bytes b) {
        std::string s = b;
        std::vector<std::string> parts;
        {
          py::gil_scoped_release rel;
          parts = c.allgather_string(s);
        }
        std::vector<py::bytes> out;
        for (auto &p : parts) out.emplace_back(p);
        return out;
      });
  py::class_<SelfComm, Communicator, std::shared_ptr<SelfComm>>(m, "SelfComm").def(py::init<>());
  py::class_<SocketComm, Communicator, std::shared_ptr<SocketComm>>(m, "SocketComm")
      .def(py::init([](int rank, int size, const std::string &addr, int port, double timeout) {
             py::gil_scoped_release rel; // blocks until every rank has connected
             return std::make_shared<SocketComm>(rank, size, addr, port, timeout);
           }),
           py::arg("rank"), py::arg("size"), py::arg("addr") = "127.0.0.1", py::arg("port") = 29517,
           py::arg("timeout") = 120.0)
      .def_static("from_env", &SocketComm::from_env, py::arg("port_offset") = 17);
  py::class_<PyComm, Communicator, std::shared_ptr<PyComm>>(m, "PyComm").def(py::init<py::object>());
  // run fn(rank, comm) on `size` threads sharing a LoopbackWorld
  m.def("loopback_run", [](int size, py::function fn) {
    py::gil_scoped_release rel;
    LoopbackWorld::run(size, [&](int rank, std::shared_ptr<Communicator> comm) {
      py::gil_scoped_acquire gil;
      try {
        fn(rank, comm);
      } catch (py::error_already_set &e) {
        // a rank released from a collective because ANOTHER rank failed: keep it recognisable,
        // so that the failure reported to the caller is the root cause
        if (std::string(e.what()).find("LoopbackWorld: another rank failed") != std::string::npos)
          throw LoopbackAborted();
        throw;
      }
    });
  });

  // ---- platform ----
  py::class_<Platform>(m, "Platform")
      .def_static(
          "make_n_streams",
          [](int n, std::shared_ptr<Communicator> comm, bool null) {
            return new Platform(Platform::make_n_streams(
                n, std::move(comm), null ? Platform::Device::Null : Platform::Device::Cuda));
          },
          py::arg("n"), py::arg("comm") = nullptr, py::arg("null") = false,
          py::return_value_policy::take_ownership)
      .def_static(
          "make_null",
          [](int n, std::shared_ptr<Communicator> comm) {
            return new Platform(Platform::make_null(n, std::move(comm)));
          },
          py::arg("n"), py::arg("comm") = nullptr, py::return_value_policy::take_ownership)
      .def("is_null", &Platform::is_null)
      .def("num_streams", &Platform::num_streams)
      .def("new_stream", [](Platform &p) { return p.new_stream().id(); })
      .def("adopt_stream",
           [](Platform &p, uintptr_t s) { return p.adopt_stream(reinterpret_cast<cudaStream_t>(s)).id(); })
      .def("ensure_streams", &Platform::ensure_streams)
      .def("cuda_stream",
           [](const Platform &p, int s) { return reinterpret_cast<uintptr_t>(p.cuda_stream(Stream(s))); })
      .def("streams",
           [](const Platform &p) {
             std::vector<int> r;
             for (const Stream &s : p.streams_) r.push_back(s.id());
             return r;
           })
      .def("provision_events", &Platform::provision_events)
      .def("comm", &Platform::comm_ptr)
      .def("epoch", &Platform::epoch)
      .def("next_epoch", &Platform::next_epoch)
      .def("set_host_execution", &Platform::set_host_execution, py::arg("on") = true,
           "Null platform + host-memory workloads: execute the ops' host restatement instead of "
           "skipping them (tenzing/host_exec.hpp)")
      .def("host_execution", &Platform::host_execution)
      .def("enable_host_streams", &Platform::enable_host_streams,
           "host execution with one worker thread per stream and CUDA's event semantics "
           "(tenzing/host_streams.hpp): ops on different streams run concurrently unless ordered")
      .def("enable_device_barrier",
           [](Platform &p) {
             if (!p.is_null() && !p.device_barrier())
               p.set_device_barrier(std::make_shared<DeviceBarrier>(p.comm_ptr()));
           },
           py::call_guard<py::gil_scoped_release>(),
           "collective: line up measurement starts with a GPU-side all-rank barrier")
      .def("synchronize", &Platform::synchronize, py::call_guard<py::gil_scoped_release>());

  // ---- ops ----
  py::class_<OpBase, std::shared_ptr<OpBase>>(m, "OpBase")
      .def("name", &OpBase::name)
      .def("desc", &OpBase::desc)
      .def("kind", &OpBase::kind)
      .def("key", &OpBase::key)
      .def("unbound_key", &OpBase::unbound_key)
      .def("uid", &OpBase::uid)
      .def("json", [](const OpBase &o) { return o.json().dump(); })
      .def("clone", &OpBase::clone)
      .def("eq", [](const OpBase &a, const OpBase &b) { return a.eq(b); })
      .def("lt", [](const OpBase &a, const OpBase &b) { return a.lt(b); })
      .def("tag", &OpBase::tag)
      .def("streams", &op_streams)
      .def("events", &op_events)
      .def("__eq__", [](const OpBase &a, const OpBase &b) { return a.eq(b); })
      .def("__hash__", [](const OpBase &a) { return a.uid(); })
      .def("__repr__", [](const OpBase &a) { return "<" + a.kind() + " " + a.desc() + ">"; });
  py::class_<BoundOp, OpBase, std::shared_ptr<BoundOp>>(m, "BoundOp")
      .def("run", &BoundOp::run, py::call_guard<py::gil_scoped_release>());
  py::class_<CpuOp, BoundOp, std::shared_ptr<CpuOp>>(m, "CpuOp");
  py::class_<Start, CpuOp, std::shared_ptr<Start>>(m, "Start").def(py::init<>());
  py::class_<Finish, CpuOp, std::shared_ptr<Finish>>(m, "Finish").def(py::init<>());
  py::class_<NoOp, CpuOp, std::shared_ptr<NoOp>>(m, "NoOp").def(py::init<const std::string &>());
  py::class_<PyCpuOp, CpuOp, std::shared_ptr<PyCpuOp>>(m, "PyCpuOp")
      .def(py::init<const std::string &, py::object>(), py::arg("name"), py::arg("fn") = py::none());
  py::class_<GpuOp, OpBase, std::shared_ptr<GpuOp>>(m, "GpuOp");
  py::class_<PyGpuOp, GpuOp, std::shared_ptr<PyGpuOp>>(m, "PyGpuOp")
      .def(py::init<const std::string &, py::object>(), py::arg("name"), py::arg("fn") = py::none());
  py::class_<LambdaGpuOp, GpuOp, std::shared_ptr<LambdaGpuOp>>(m, "EmptyGpuOp")
      .def(py::init([](const std::string &n) { return std::make_shared<LambdaGpuOp>(n); }));
  py::class_<BoundGpuOp, BoundOp, std::shared_ptr<BoundGpuOp>>(m, "BoundGpuOp")
      .def(py::init([](const OpPtr &op, int s) {
        // ops defined in C++ workload libraries arrive typed as OpBase
        auto g = std::dynamic_pointer_cast<GpuOp>(op);
        if (!g) throw std::runtime_error("BoundGpuOp: " + op->desc() + " is not a GpuOp");
        return std::make_shared<BoundGpuOp>(g, Stream(s));
      }))
      .def("stream", [](const BoundGpuOp &b) { return b.stream().id(); })
      .def("unbound", &BoundGpuOp::unbound);
  py::class_<ChoiceOp, OpBase, std::shared_ptr<ChoiceOp>>(m, "ChoiceOpBase")
      .def("choices", &ChoiceOp::choices);
  py::class_<ListChoiceOp, ChoiceOp, std::shared_ptr<ListChoiceOp>>(m, "ChoiceOp")
      .def(py::init<const std::string &, const std::vector<OpPtr> &>());
  py::class_<CompoundOp, OpBase, std::shared_ptr<CompoundOp>>(m, "CompoundOpBase")
      .def("graph", &CompoundOp::graph);
  py::class_<GraphCompoundOp, CompoundOp, std::shared_ptr<GraphCompoundOp>>(m, "CompoundOp")
      .def(py::init<const std::string &, const G &>());

  py::class_<CudaEventRecord, BoundOp, std::shared_ptr<CudaEventRecord>>(m, "CudaEventRecord")
      .def(py::init([](int e, int s, const std::string &n) {
             return std::make_shared<CudaEventRecord>(Event(e), Stream(s), n);
           }),
           py::arg("event"), py::arg("stream"), py::arg("name") = "CER-anon")
      .def("event", [](const CudaEventRecord &o) { return o.event().id(); })
      .def("stream", [](const CudaEventRecord &o) { return o.stream().id(); });
  py::class_<CudaStreamWaitEvent, BoundOp, std::shared_ptr<CudaStreamWaitEvent>>(m, "CudaStreamWaitEvent")
      .def(py::init([](int s, int e, const std::string &n) {
             return std::make_shared<CudaStreamWaitEvent>(Stream(s), Event(e), n);
           }),
           py::arg("stream"), py::arg("event"), py::arg("name") = "CSWE-anon")
      .def("event", [](const CudaStreamWaitEvent &o) { return o.event().id(); })
      .def("stream", [](const CudaStreamWaitEvent &o) { return o.stream().id(); });
  py::class_<CudaEventSync, BoundOp, std::shared_ptr<CudaEventSync>>(m, "CudaEventSync")
      .def(py::init([](int e, const std::string &n) {
             return std::make_shared<CudaEventSync>(Event(e), n);
           }),
           py::arg("event"), py::arg("name") = "CES-anon")
      .def("event", [](const CudaEventSync &o) { return o.event().id(); });
  py::class_<StreamWait, BoundOp, std::shared_ptr<StreamWait>>(m, "StreamWait")
      .def(py::init([](int waitee, int waiter, int e) {
        return std::make_shared<StreamWait>(Stream(waitee), Stream(waiter), Event(e));
      }));
  py::class_<StreamSync, BoundOp, std::shared_ptr<StreamSync>>(m, "StreamSync")
      .def(py::init([](int s) { return std::make_shared<StreamSync>(Stream(s)); }))
      .def("stream", [](const StreamSync &o) { return o.stream().id(); });

  m.def("keep_uniques", [](std::vector<BoundOpPtr> v) {
    keep_uniques(v);
    return v;
  });
  m.def("make_platform_variations", &make_platform_variations);
  m.def("unbound_contains", &unbound_contains);
  m.def("contains", &contains);

  // ---- graph ----
  py::class_<G>(m, "Graph")
      .def(py::init<>())
      .def(py::init<const G &>())
      .def("start", &G::start)
      .def("finish", &G::finish)
      .def("start_then", [](G &g, const OpPtr &op) { return g.start_then(op); })
      .def("then", [](G &g, const OpPtr &a, const OpPtr &b) { return g.then(a, b); })
      .def("then_finish", &G::then_finish)
      .def("add_edge", [](G &g, const OpPtr &a, const OpPtr &b) { g.add_edge(a, b); })
      .def("vertex_size", &G::vertex_size)
      .def("edge_size", &G::edge_size)
      .def("validate", &G::validate,
           "'' if the structural invariants hold (mirrored edges, indexes, acyclic, every vertex on a "
           "Start..Finish path), else the first violation")
      .def("preds_find_name", &G::preds_find_name)
      .def("succs_find_name", &G::succs_find_name)
      .def("contains", &G::contains)
      .def("contains_unbound", &G::contains_unbound)
      .def("find", &G::find)
      .def("find_unbound", &G::find_unbound)
      .def("find_name", &G::find_name)
      .def("vertices", &G::vertices)
      .def("start_vertices", &G::start_vertices)
      .def("finish_vertices", &G::finish_vertices)
      .def("succs", &G::succs)
      .def("preds", &G::preds)
      .def("clone", &G::clone)
      .def("replace", &G::replace)
      .def("clone_but_replace", &G::clone_but_replace, py::arg("dst"), py::arg("src"))
      .def("clone_but_expand", &G::clone_but_expand)
      .def("erase", &G::erase)
      .def("erase_edge_only", &G::erase_edge_only)
      .def("frontier", [](const G &g, const std::vector<OpPtr> &done) { return g.frontier(done); })
      .def("frontier", [](const G &g, const Seq &done) { return g.frontier(done.vector()); })
      .def("graphviz", &G::graphviz)
      .def("dump_graphviz", &G::dump_graphviz)
      .def("__str__", &G::str);
  m.def("use_streams", [](const G &g, const std::vector<int> &s) {
    std::vector<Stream> v;
    for (int i : s) v.emplace_back(i);
    return use_streams(g, v);
  });
  m.def("use_streams2", [](const G &g, const std::vector<int> &s) {
    std::vector<Stream> v;
    for (int i : s) v.emplace_back(i);
    return use_streams2(g, v);
  });
  m.def("apply_assignment", [](const G &g, const std::vector<std::pair<std::string, int>> &a) {
    std::vector<std::pair<std::string, Stream>> v;
    for (const auto &kv : a) v.emplace_back(kv.first, Stream(kv.second));
    return apply_assignment(g, v);
  });
  m.def("is_equivalent_stream_mapping", &is_equivalent_stream_mapping);
  m.def("insert_synchronization", &insert_synchronization);
  m.def("graph_equivalence", [](const G &a, const G &b) { return get_equivalence(a, b); });

  // ---- sequence ----
  py::class_<Seq>(m, "Sequence")
      .def(py::init<>())
      .def(py::init([](const std::vector<BoundOpPtr> &v) { return Seq(v); }))
      .def("__len__", &Seq::size)
      .def("__getitem__",
           [](const Seq &s, long i) {
             if (i < 0) i += long(s.size());
             if (i < 0 || size_t(i) >= s.size()) throw py::index_error();
             return s[size_t(i)];
           })
      .def("__iter__", [](const Seq &s) { return py::make_iterator(s.begin(), s.end()); },
           py::keep_alive<0, 1>())
      .def("push_back", [](Seq &s, const BoundOpPtr &op) { s.push_back(op); })
      .def("ops", [](const Seq &s) { return s.vector(); })
      .def("names",
           [](const Seq &s) {
             std::vector<std::string> r;
             for (const auto &o : s) r.push_back(o->name());
             return r;
           })
      .def("contains", [](const Seq &s, const OpBase &o) { return s.contains(o); })
      .def("contains_unbound", [](const Seq &s, const OpBase &o) { return s.contains_unbound(o); })
      .def("new_unique_event", [](const Seq &s) { return s.new_unique_event().id(); })
      .def("desc", &Seq::desc, py::arg("delim") = ", ")
      .def("canonical_key", [](const Seq &s) { return canonical_key(s); })
      .def("to_json", [](const Seq &s, const G &g) { return to_json(s, g).dump(); })
      .def("run", [](Seq &s, Platform &p) { run_sequence(s, p); },
           py::call_guard<py::gil_scoped_release>())
      .def("__repr__", [](const Seq &s) { return "<Sequence " + s.desc() + ">"; });
  m.def("sequence_from_json", [](const std::string &text, const G &g) {
    Seq s;
    from_json(Json::parse(text), g, s);
    return s;
  });
  m.def("op_from_json", [](const std::string &text, const G &g) {
    BoundOpPtr op;
    from_json(Json::parse(text), g, op);
    return op;
  });
  m.def("sequence_equivalence", [](const Seq &a, const Seq &b) { return get_equivalence(a, b); });
  m.def("bcast_sequence",
        [](const Seq &s, const G &g, Communicator &c, int root) { return bcast(s, g, c, root); },
        py::arg("seq"), py::arg("graph"), py::arg("comm"), py::arg("root") = 0,
        py::call_guard<py::gil_scoped_release>());
  m.def("csv_row", &csv_row);

  // ---- decisions / state / synchronizer ----
  py::class_<Decision, DecisionPtr>(m, "Decision")
      .def("desc", &Decision::desc)
      .def("kind",
           [](const Decision &d) {
             switch (d.kind()) {
             case Decision::Kind::ExecuteOp: return "ExecuteOp";
             case Decision::Kind::ExpandOp: return "ExpandOp";
             case Decision::Kind::ChooseOp: return "ChooseOp";
             case Decision::Kind::AssignOpStream: return "AssignOpStream";
             }
             return "?";
           })
      .def("__repr__", [](const Decision &d) { return "<Decision " + d.desc() + ">"; });
  py::class_<ExecuteOp, Decision, std::shared_ptr<ExecuteOp>>(m, "ExecuteOp")
      .def(py::init<const BoundOpPtr &>())
      .def_readonly("op", &ExecuteOp::op);
  py::class_<ExpandOp, Decision, std::shared_ptr<ExpandOp>>(m, "ExpandOp")
      .def(py::init<const std::shared_ptr<CompoundOp> &>())
      .def_readonly("op", &ExpandOp::op);
  py::class_<ChooseOp, Decision, std::shared_ptr<ChooseOp>>(m, "ChooseOp")
      .def(py::init<const std::shared_ptr<ChoiceOp> &, const OpPtr &>())
      .def_readonly("orig", &ChooseOp::orig)
      .def_readonly("replacement", &ChooseOp::replacement);
  py::class_<AssignOpStream, Decision, std::shared_ptr<AssignOpStream>>(m, "AssignOpStream")
      .def(py::init([](const GpuOpPtr &op, int s) { return std::make_shared<AssignOpStream>(op, Stream(s)); }))
      .def_readonly("op", &AssignOpStream::op)
      .def_property_readonly("stream", [](const AssignOpStream &a) { return a.stream.id(); });

  py::class_<SDP::State>(m, "State")
      .def(py::init<const G &>())
      .def(py::init<const G &, const Seq &>())
      .def("graph", &SDP::State::graph, py::return_value_policy::copy)
      .def("sequence", &SDP::State::sequence, py::return_value_policy::copy)
      .def("get_syncs_before_op", &SDP::State::get_syncs_before_op)
      .def("get_decisions", &SDP::State::get_decisions, py::arg("plat"), py::arg("quiet") = true)
      .def("apply", &SDP::State::apply)
      .def("frontier", &SDP::State::frontier, py::arg("plat"), py::arg("quiet") = true)
      .def("is_terminal", &SDP::State::is_terminal)
      .def("canonical_key", [](const SDP::State &st) { return py::bytes(st.canonical_key()); },
           "opaque bytes: equal for states that differ only in the numbering of streams and events "
           "(comparable within one process only)")
      .def("canonical_hash",
           [](const SDP::State &st) {
             const StateKey k = st.canonical_hash();
             return py::make_tuple(k.a, k.b);
           },
           "128-bit digest of the same canonical form (what the solvers de-duplicate on)");
  m.def("state_equivalence",
        [](const SDP::State &a, const SDP::State &b) { return SDP::get_equivalence(a, b); });

  py::class_<EventSynchronizer>(m, "EventSynchronizer")
      .def_static("is_synced",
                  [](const BoundOpPtr &op, const G &g, const Seq &path) {
                    return EventSynchronizer::is_synced(op, g, path);
                  })
      .def_static("make_syncs",
                  [](const BoundOpPtr &op, const G &g, const Seq &path, bool quiet) {
                    return EventSynchronizer::make_syncs(op, g, path, quiet);
                  },
                  py::arg("op"), py::arg("graph"), py::arg("path"), py::arg("quiet") = true);

  // cross-rank flag dependencies (device-side synchronisation between GPUs)
  py::class_<FlagSynchronizer>(m, "FlagSynchronizer")
      .def_static("is_deadlock_free", &FlagSynchronizer::is_deadlock_free)
      .def_static("check",
                  [](const Seq &s) {
                    std::vector<std::tuple<size_t, std::string, uint64_t>> out;
                    for (const auto &v : FlagSynchronizer::check(s)) out.emplace_back(v.position, v.op, v.channel);
                    return out;
                  })
      .def_static("add_edges", &FlagSynchronizer::add_edges)
      .def_static("is_ready", [](const BoundOp &op, const Seq &path) { return FlagSynchronizer::is_ready(op, path); });

  // ---- schedule ----
  m.def("remove_redundant_syncs", [](Seq &s) { return Schedule::remove_redundant_syncs(s); });
  m.def("make_schedules",
        [](const G &g, size_t maxCount) {
          std::vector<Seq> r;
          for (auto &s : make_schedules(g, maxCount)) r.push_back(std::move(s.order));
          return r;
        },
        py::arg("graph"), py::arg("max_count") = 0);
  m.def("make_schedules_random",
        [](const G &g, size_t n, uint64_t seed) {
          std::vector<Seq> r;
          for (auto &s : make_schedules_random(g, n, seed)) r.push_back(std::move(s.order));
          return r;
        },
        py::arg("graph"), py::arg("n"), py::arg("seed") = 0);
  m.def("count_schedules", &count_schedules);

  // ---- benchmarkers ----
  py::class_<Benchmark::Result>(m, "BenchResult")
      .def(py::init<>())
      .def_readwrite("pct01", &Benchmark::Result::pct01)
      .def_readwrite("pct10", &Benchmark::Result::pct10)
      .def_readwrite("pct50", &Benchmark::Result::pct50)
      .def_readwrite("pct90", &Benchmark::Result::pct90)
      .def_readwrite("pct99", &Benchmark::Result::pct99)
      .def_readwrite("stddev", &Benchmark::Result::stddev)
      .def_static("from_samples", &Benchmark::Result::from_samples)
      .def("__repr__", [](const Benchmark::Result &r) {
        std::ostringstream ss;
        ss << "<BenchResult pct10=" << r.pct10 << " pct50=" << r.pct50 << " stddev=" << r.stddev << ">";
        return ss.str();
      });
  py::class_<Benchmark::Opts>(m, "BenchOpts")
      .def(py::init<>())
      .def(py::init([](size_t nIters, size_t maxRetries, double targetSecs, bool check) {
             Benchmark::Opts o;
             o.nIters = nIters;
             o.maxRetries = maxRetries;
             o.targetSecs = targetSecs;
             o.checkRandomness = check;
             return o;
           }),
           py::arg("n_iters") = 1000, py::arg("max_retries") = 10, py::arg("target_secs") = 0.01,
           py::arg("check_randomness") = true)
      .def_readwrite("n_iters", &Benchmark::Opts::nIters)
      .def_readwrite("max_retries", &Benchmark::Opts::maxRetries)
      .def_readwrite("target_secs", &Benchmark::Opts::targetSecs)
      .def_readwrite("check_randomness", &Benchmark::Opts::checkRandomness);
  py::class_<Benchmarker>(m, "Benchmarker")
      .def("benchmark",
           [](Benchmarker &b, Seq &s, Platform &p, const Benchmark::Opts &o) {
             return b.benchmark(s, p, o);
           },
           py::arg("seq"), py::arg("plat"), py::arg("opts") = Benchmark::Opts(),
           py::call_guard<py::gil_scoped_release>());
  py::class_<EmpiricalBenchmarker, Benchmarker>(m, "EmpiricalBenchmarker")
      .def(py::init<>())
      // several sequences measured interleaved, in a fresh random order per iteration
      .def("benchmark_batch",
           [](EmpiricalBenchmarker &b, const std::vector<Seq> &seqs, Platform &p, const Benchmark::Opts &o) {
             std::vector<Schedule> scheds(seqs.size());
             for (size_t i = 0; i < seqs.size(); ++i) scheds[i].order = seqs[i];
             return b.benchmark(scheds, p, o);
           },
           py::arg("seqs"), py::arg("plat"), py::arg("opts") = Benchmark::Opts(),
           py::call_guard<py::gil_scoped_release>());
  py::class_<DeviceBenchmarker, EmpiricalBenchmarker>(m, "DeviceBenchmarker").def(py::init<>());
  py::class_<CsvBenchmarker, Benchmarker>(m, "CsvBenchmarker")
      .def(py::init<const std::string &, const G &>())
      .def_static("from_text", [](const std::string &t, const G &g) {
        return new CsvBenchmarker(CsvBenchmarker::from_text(t, g));
      }, py::return_value_policy::take_ownership)
      .def("num_rows", [](const CsvBenchmarker &b) { return b.rows().size(); })
      .def("producer", [](const CsvBenchmarker &b) -> py::object {
        return header_to_py(b.producer());
      });
  py::class_<PyBenchmarker, Benchmarker>(m, "PyBenchmarker").def(py::init<py::object>());
  py::class_<CachingBenchmarker, Benchmarker>(m, "CachingBenchmarker")
      .def(py::init<Benchmarker &, bool>(), py::arg("inner"), py::arg("reuse") = true,
           py::keep_alive<1, 2>())
      .def("preload", [](CachingBenchmarker &b, const Seq &s, const Benchmark::Result &r) { b.preload(s, r); })
      .def("preload_csv", [](CachingBenchmarker &b, const CsvBenchmarker &c) { b.preload(c); })
      .def("hits", &CachingBenchmarker::hits)
      .def("misses", &CachingBenchmarker::misses)
      .def("size", &CachingBenchmarker::size)
      .def("pending", &CachingBenchmarker::pending);
  py::class_<SimBenchmarker, Benchmarker>(m, "SimBenchmarker")
      .def(py::init([](const std::unordered_map<std::string, double> &gpu,
                       const std::unordered_map<std::string, double> &cpu, double defaultGpu,
                       double defaultCpu, double launchCost, double contention) {
             SimBenchmarker::Model mod;
             mod.gpuSeconds = gpu, mod.cpuSeconds = cpu;
             mod.defaultGpuSeconds = defaultGpu, mod.defaultCpuSeconds = defaultCpu;
             mod.launchCost = launchCost, mod.contention = contention;
             return new SimBenchmarker(mod);
           }),
           py::arg("gpu_seconds") = std::unordered_map<std::string, double>{},
           py::arg("cpu_seconds") = std::unordered_map<std::string, double>{},
           py::arg("default_gpu_seconds") = 10e-6, py::arg("default_cpu_seconds") = 0.0,
           py::arg("launch_cost") = 3e-6, py::arg("contention") = 0.0)
      .def("simulate", &SimBenchmarker::simulate)
      .def("timeline", [](const SimBenchmarker &b, const Seq &s) {
        std::vector<std::pair<double, double>> out;
        for (const auto &sp : b.timeline(s)) out.emplace_back(sp.start, sp.end);
        return out;
      });

  // device-timed steps with an L2 flush between them (bench.py): returns seconds per step
  m.def(
      "device_time_steps",
      [](Seq &seq, Platform &plat, int steps, uintptr_t flushPtr, size_t flushBytes) {
        std::vector<double> out;
        if (plat.is_null()) return out;
        cudaEvent_t start, stop;
        std::vector<cudaEvent_t> joins(size_t(plat.num_streams()));
        CUDA_RUNTIME(cudaEventCreate(&start));
        CUDA_RUNTIME(cudaEventCreate(&stop));
        for (auto &e : joins) CUDA_RUNTIME(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        cudaStream_t s0 = plat.cuda_stream(plat.streams_[0]);
        // Ranks are lined up on the GPUs, not on the hosts: a host barrier leaves tens of
        // microseconds of skew, which a device-timed exchange would count as waiting time.
        if (plat.comm().size() > 1 && !plat.device_barrier())
          plat.set_device_barrier(std::make_shared<DeviceBarrier>(plat.comm_ptr()));
        for (int i = 0; i < steps; ++i) {
          plat.synchronize();
          plat.comm().barrier();
          // The L2 flush is queued in front of the step and NOT waited for: while it runs the
          // host enqueues the device barrier, the start event and the step's kernels, so the
          // timed region begins with its work already queued (as in the reference's
          // back-to-back protocol, src/benchmarker.cpp:93-97) instead of with an idle GPU
          // waiting for the first launch to arrive.
          if (flushPtr && flushBytes)
            CUDA_RUNTIME(cudaError_t(tzb_l2_flush(reinterpret_cast<void *>(flushPtr), flushBytes, unsigned(i), s0)));
          else
            CUDA_RUNTIME(cudaError_t(tzb_delay_us(30, s0)));
          if (plat.device_barrier()) plat.device_barrier()->enqueue(s0);
          CUDA_RUNTIME(cudaEventRecord(start, s0));
          for (int k = 1; k < plat.num_streams(); ++k)
            CUDA_RUNTIME(cudaStreamWaitEvent(plat.cuda_stream(plat.streams_[size_t(k)]), start, 0));
          run_sequence(seq, plat);
          for (int k = 1; k < plat.num_streams(); ++k) {
            CUDA_RUNTIME(cudaEventRecord(joins[size_t(k)], plat.cuda_stream(plat.streams_[size_t(k)])));
            CUDA_RUNTIME(cudaStreamWaitEvent(s0, joins[size_t(k)], 0));
          }
          CUDA_RUNTIME(cudaEventRecord(stop, s0));
          CUDA_RUNTIME(cudaEventSynchronize(stop));
          float ms = 0;
          CUDA_RUNTIME(cudaEventElapsedTime(&ms, start, stop));
          out.push_back(double(ms) * 1e-3);
        }
        cudaEventDestroy(start);
        cudaEventDestroy(stop);
        for (auto &e : joins) cudaEventDestroy(e);
        return out;
      },
      py::arg("seq"), py::arg("plat"), py::arg("steps"), py::arg("flush_ptr") = 0,
      py::arg("flush_bytes") = 0, py::call_guard<py::gil_scoped_release>());

  // ---- CUDA-graph replay of a sequence ----
  py::class_<GraphReplay>(m, "GraphReplay")
      .def(py::init([](Seq &seq, Platform &plat) { return new GraphReplay(seq, plat); }), py::arg("seq"),
           py::arg("plat"), py::keep_alive<1, 3>())
      .def_static("supported", [](const Seq &seq) {
        std::string why;
        const bool ok = GraphReplay::supported(seq, &why);
        return std::make_pair(ok, why);
      })
      .def("launch", &GraphReplay::launch, py::call_guard<py::gil_scoped_release>())
      .def("kernel_nodes", &GraphReplay::kernel_nodes);
  // device time of single replays, measured exactly as device_time_steps measures single steps
  m.def(
      "device_time_graph_steps",
      [](GraphReplay &replay, Platform &plat, int steps, uintptr_t flushPtr, size_t flushBytes) {
        std::vector<double> out;
        cudaEvent_t start, stop;
        CUDA_RUNTIME(cudaEventCreate(&start));
        CUDA_RUNTIME(cudaEventCreate(&stop));
        cudaStream_t s0 = replay.stream();
        if (plat.comm().size() > 1 && !plat.device_barrier())
          plat.set_device_barrier(std::make_shared<DeviceBarrier>(plat.comm_ptr()));
        for (int i = 0; i < steps; ++i) {
          plat.synchronize();
          plat.comm().barrier();
          if (flushPtr && flushBytes)
            CUDA_RUNTIME(cudaError_t(tzb_l2_flush(reinterpret_cast<void *>(flushPtr), flushBytes, unsigned(i), s0)));
          else
            CUDA_RUNTIME(cudaError_t(tzb_delay_us(30, s0)));
          if (plat.device_barrier()) plat.device_barrier()->enqueue(s0);
          CUDA_RUNTIME(cudaEventRecord(start, s0));
          replay.launch();
          CUDA_RUNTIME(cudaEventRecord(stop, s0));
          CUDA_RUNTIME(cudaEventSynchronize(stop));
          float ms = 0;
          CUDA_RUNTIME(cudaEventElapsedTime(&ms, start, stop));
          out.push_back(double(ms) * 1e-3);
        }
        cudaEventDestroy(start);
        cudaEventDestroy(stop);
        return out;
      },
      py::arg("replay"), py::arg("plat"), py::arg("steps"), py::arg("flush_ptr") = 0,
      py::arg("flush_bytes") = 0, py::call_guard<py::gil_scoped_release>());

  // ---- DFS ----
  py::module_ d = m.def_submodule("dfs");
  py::class_<dfs::Opts>(d, "Opts")
      .def(py::init<>())
      .def_readwrite("max_seqs", &dfs::Opts::maxSeqs)
      .def_readwrite("remove_redundant_syncs", &dfs::Opts::removeRedundantSyncs)
      .def_readwrite("bench_opts", &dfs::Opts::benchOpts);
  py::class_<dfs::SimResult>(d, "SimResult")
      .def_readonly("seq", &dfs::SimResult::seq)
      .def_readonly("bench_result", &dfs::SimResult::benchResult);
  py::class_<dfs::Result>(d, "Result")
      .def_readonly("sim_results", &dfs::Result::simResults)
      .def("csv", &dfs::Result::csv)
      .def("best", &dfs::Result::best);
  d.def("get_all_sequences", &dfs::get_all_sequences, py::arg("graph"), py::arg("plat"),
        py::arg("max_seqs") = -1, py::arg("remove_redundant_syncs") = false,
        py::call_guard<py::gil_scoped_release>());
  d.def("explore", &dfs::explore, py::arg("graph"), py::arg("plat"), py::arg("benchmarker"),
        py::arg("opts") = dfs::Opts(), py::call_guard<py::gil_scoped_release>());

  // ---- MCTS ----
  py::module_ mc = m.def_submodule("mcts");
  py::class_<mcts::Opts>(mc, "Opts")
      .def(py::init<>())
      .def_readwrite("n_iters", &mcts::Opts::nIters)
      .def_readwrite("dump_tree", &mcts::Opts::dumpTree)
      .def_readwrite("dump_tree_prefix", &mcts::Opts::dumpTreePrefix)
      .def_readwrite("expand_rollout", &mcts::Opts::expandRollout)
      .def_readwrite("seed", &mcts::Opts::seed)
      .def_readwrite("remove_redundant_syncs", &mcts::Opts::removeRedundantSyncs)
      .def_readwrite("max_tree_nodes", &mcts::Opts::maxTreeNodes)
      .def_readwrite("bench_opts", &mcts::Opts::benchOpts);
  py::class_<mcts::SimResult>(mc, "SimResult")
      .def_readonly("path", &mcts::SimResult::path)
      .def_readonly("bench_result", &mcts::SimResult::benchResult);
  py::class_<mcts::Result>(mc, "Result")
      .def_readonly("sim_results", &mcts::Result::simResults)
      .def("csv", &mcts::Result::csv)
      .def("best", &mcts::Result::best);
  py::class_<mcts::Strategy>(mc, "Strategy").def("name", &mcts::Strategy::name);
  mc.def("make_strategy", [](const std::string &n) { return mcts::make_strategy(n); });
  mc.def("strategy_names", &mcts::strategy_names);
  mc.def("histogram", &mcts::histogram, py::arg("v"), py::arg("t_min"), py::arg("t_max"),
         py::arg("n_bins") = 10);
  mc.def("explore",
         [](const G &g, Platform &plat, Benchmarker &b, const std::string &strategy,
            const mcts::Opts &opts) {
           auto s = mcts::make_strategy(strategy);
           return mcts::explore(g, plat, b, *s, opts);
         },
         py::arg("graph"), py::arg("plat"), py::arg("benchmarker"), py::arg("strategy") = "FastMin",
         py::arg("opts") = mcts::Opts(), py::call_guard<py::gil_scoped_release>());
  mc.def("get_frontier", &mcts::get_frontier);

  // tree-level access for tests / notebooks
  py::class_<mcts::Node>(mc, "Node")
      .def(py::init<const G &>())
      .def_readonly("n", &mcts::Node::n_)
      .def_readonly("fully_visited", &mcts::Node::fullyVisited_)
      .def_readonly("expanded", &mcts::Node::expanded_)
      .def("op", [](const mcts::Node &n) { return n.op_; })
      .def("parent", [](mcts::Node &n) { return n.parent_; }, py::return_value_policy::reference)
      .def("children",
           [](mcts::Node &n) {
             py::list l;
             // references into the tree: valid while the root object is alive
             for (auto &c : n.children_) l.append(py::cast(c.get(), py::return_value_policy::reference));
             return l;
           })
      .def("times", [](const mcts::Node &n) { return n.state_.times; })
      .def("size", &mcts::Node::size)
      .def("unvisited_size", &mcts::Node::unvisited_size)
      .def("fully_visited_size", &mcts::Node::fully_visited_size)
      .def("is_leaf", &mcts::Node::is_leaf)
      .def("is_terminal", &mcts::Node::is_terminal)
      .def("get_sequence", &mcts::Node::get_sequence)
      .def("ensure_children", &mcts::Node::ensure_children)
      .def("expand", &mcts::Node::expand, py::return_value_policy::reference)
      .def("root", [](mcts::Node &n) -> mcts::Node & { return n.root(); },
           py::return_value_policy::reference)
      .def("graphviz", [](const mcts::Node &n) { return mcts::graphviz(n); })
      .def("desc", &mcts::Node::desc);
  // one full MCTS iteration on an explicit tree with a Python cost function
  mc.def("iterate", [](mcts::Node &root, const Platform &plat, mcts::Strategy &strategy,
                       py::function cost, uint64_t seed, bool expandRollout) {
    std::mt19937_64 rng(seed);
    mcts::Node &sel = root.select(strategy, rng);
    mcts::Node &child = sel.expand(plat);
    auto rr = child.get_rollout(plat, rng, expandRollout);
    Schedule::remove_redundant_syncs(rr.sequence);
    const double t = cost(rr.sequence).cast<double>();
    Benchmark::Result br;
    br.pct01 = br.pct10 = br.pct50 = br.pct90 = br.pct99 = t;
    rr.backpropStart->backprop(strategy, br);
    return rr.sequence;
  });

  bind_workloads(m);
}
