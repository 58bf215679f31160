// Python bindings for the distributed SpMV workload.
#include <pybind11/pybind11.h>
#include "op_cast.hpp"
#include <pybind11/stl.h>

#include "tenzing/spmv/mtx_reader.hpp"
#include "tenzing/spmv/row_part_spmv.hpp"
#include "tenzing/tenzing.hpp"

namespace py = pybind11;
using namespace tenzing;

namespace {

template <typename T> py::bytes as_bytes(const std::vector<T> &v) {
  return py::bytes(reinterpret_cast<const char *>(v.data()), v.size() * sizeof(T));
}

template <typename Scalar> py::dict csr_dict(const CsrMat<int, Scalar> &m) {
  py::dict d;
  d["rows"] = m.num_rows();
  d["cols"] = m.num_cols();
  d["row_ptr"] = as_bytes(m.rowPtr_);
  d["col_ind"] = as_bytes(m.colInd_);
  d["val"] = as_bytes(m.val_);
  return d;
}

template <typename Scalar>
typename RowPartSpmv<Scalar>::Style spmv_style(const std::string &s) {
  using Style = typename RowPartSpmv<Scalar>::Style;
  if (s == "reference") return Style::Reference;
  if (s == "fused") return Style::Fused;
  if (s == "fused_separate") return Style::FusedSeparate;
  if (s == "fused_all") return Style::FusedAll;
  if (s == "fused_pull") return Style::FusedPull;
  throw std::runtime_error("unknown spmv style '" + s + "'");
}

template <typename Scalar> void bind_one(py::module_ &m, const char *suffix) {
  using R = RowPartSpmv<Scalar>;
  const std::string cls = std::string("RowPartSpmv") + suffix;
  py::class_<R, std::shared_ptr<R>>(m, cls.c_str())
      .def("rank", &R::rank)
      .def("size", &R::size)
      .def("global_rows", &R::global_rows)
      .def("row_range", [](const R &r) { return std::make_pair(r.row_range().lb, r.row_range().ub); })
      .def("local_rows", [](const R &r) { return r.row_range().extent(); })
      .def("num_remote_cols", &R::num_remote_cols)
      .def("num_send", &R::num_send)
      .def("local_nnz", [](const R &r) { return r.lA().nnz; })
      .def("remote_nnz", [](const R &r) { return r.rA().nnz; })
      .def("lx_ptr", [](const R &r) { return reinterpret_cast<uintptr_t>(r.lx()); })
      .def("ly_ptr", [](const R &r) { return reinterpret_cast<uintptr_t>(r.ly()); })
      .def("ry_ptr", [](const R &r) { return reinterpret_cast<uintptr_t>(r.ry()); })
      .def("rx_ptr", [](const R &r, int parity) { return reinterpret_cast<uintptr_t>(r.rx(r.rank(), parity)); })
      .def("send_buf_ptr", [](const R &r) { return reinterpret_cast<uintptr_t>(r.x_send_buf()); })
      .def("remote_globals", [](const R &r) { return as_bytes(r.host_split().globals); })
      .def("local_csr", [](const R &r) { return csr_dict(r.host_split().local); })
      .def("remote_csr", [](const R &r) { return csr_dict(r.host_split().remote); })
      .def("send_params", [](const R &r) {
        std::vector<std::vector<int>> v;
        for (const auto &p : r.send_params()) v.push_back({p.peer, p.displ, p.count, p.dstDispl});
        return v;
      })
      .def("recv_params", [](const R &r) {
        std::vector<std::vector<int>> v;
        for (const auto &p : r.recv_params()) v.push_back({p.peer, p.displ, p.count});
        return v;
      })
      .def("peers", &R::peers)
      .def("recv_bytes", &R::recv_bytes)
      .def("hbm_bytes", &R::hbm_bytes);
  const std::string mk = std::string("make_band_spmv") + suffix;
  m.def(mk.c_str(),
        [](std::shared_ptr<Communicator> comm, int64_t n, int64_t bw, int nnzPerRow, uint64_t seed,
           bool host) {
          return make_band_spmv<Scalar>(std::move(comm), n, bw, nnzPerRow, seed,
                                        host ? SymmetricHeap::Memory::Host
                                             : SymmetricHeap::Memory::Device);
        },
        py::arg("comm"), py::arg("n"), py::arg("bw"), py::arg("nnz_per_row") = 10,
        py::arg("seed") = 0, py::arg("host") = false, py::call_guard<py::gil_scoped_release>());
  // matrix from a MatrixMarket file: rank 0 reads it and hands every rank its row block
  const std::string mtx = std::string("make_mtx_spmv") + suffix;
  m.def(mtx.c_str(),
        [](std::shared_ptr<Communicator> comm, const std::string &path, bool host) {
          if (!comm) comm = std::make_shared<SelfComm>();
          CsrMat<int, Scalar> whole;
          // rank 0 reads; a bad file must fail on every rank, not hang the others in a collective
          int64_t n = -1;
          std::string error;
          if (comm->rank() == 0) {
            try {
              whole = MtxReader<int, Scalar>(path).read_csr();
              if (whole.num_rows() != whole.num_cols())
                THROW_RUNTIME(path << ": " << whole.num_rows() << " x " << whole.num_cols()
                                   << " - the distributed SpMV needs a square matrix");
              n = whole.num_rows();
            } catch (const std::exception &e) {
              error = e.what();
            }
          }
          comm->bcast(&n, sizeof(n), 0);
          if (n < 0) {
            error = comm->bcast_string(error, 0);
            throw std::runtime_error(error);
          }
          CsrMat<int, Scalar> mine = distribute_rows<Scalar>(*comm, comm->rank() == 0 ? &whole : nullptr, 0);
          return std::make_shared<R>(comm, mine, n,
                                     host ? SymmetricHeap::Memory::Host : SymmetricHeap::Memory::Device);
        },
        py::arg("comm"), py::arg("path"), py::arg("host") = false,
        py::call_guard<py::gil_scoped_release>());
  const std::string rows = std::string("random_band_rows") + suffix;
  m.def(rows.c_str(),
        [](int64_t n, int64_t bw, int nnzPerRow, int64_t rb, int64_t re, uint64_t seed) {
          return csr_dict(random_band_rows<int, Scalar>(n, bw, nnzPerRow, rb, re, seed));
        },
        py::arg("n"), py::arg("bw"), py::arg("nnz_per_row"), py::arg("row_begin"),
        py::arg("row_end"), py::arg("seed") = 0);
  const std::string comp = std::string("make_spmv") + suffix;
  m.def(comp.c_str(),
        [](const std::shared_ptr<R> &r, const std::string &style, const std::string &name) {
          return std::static_pointer_cast<CompoundOp>(
              std::make_shared<SpMV<Scalar>>(name, r, spmv_style<Scalar>(style)));
        },
        py::arg("spmv"), py::arg("style") = "fused", py::arg("name") = "spmv");
  const std::string ch = std::string("make_spmv_choice") + suffix;
  m.def(ch.c_str(),
        [](const std::shared_ptr<R> &r, bool withExperimental) {
          return make_spmv_choice<Scalar>(r, withExperimental);
        },
        py::arg("spmv"), py::arg("with_experimental") = false);
}

} // namespace

void bind_spmv(py::module_ &m) {
  py::module_ s = m.def_submodule("spmv");
  bind_one<double>(s, "F64");
  bind_one<float>(s, "F32");
  s.def("get_partition", [](int64_t domain, int i, int n) {
    const Range r = get_partition(domain, i, n);
    return std::make_pair(r.lb, r.ub);
  });
  s.def("get_owner", &get_owner);
  s.def("styles", [] { return std::vector<std::string>{"reference", "fused", "fused_separate", "fused_all"}; });
  // built and host-executed, but their kernels have not run on hardware yet
  s.def("experimental_styles", [] { return std::vector<std::string>{"fused_pull"}; });
}
