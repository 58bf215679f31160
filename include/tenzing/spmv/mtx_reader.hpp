// MatrixMarket coordinate-format reader (the reference typedefs one from
// cwpearson/mm and never uses it, tenzing-dfs/examples/spmv.cu:35-37; here the
// drivers accept --matrix-file).  Supports real/integer/pattern, general/symmetric.
#pragma once

#include <fstream>
#include <sstream>
#include <string>

#include "tenzing/macro_at.hpp"
#include "tenzing/spmv/csr_mat.hpp"

namespace tenzing {

template <typename Ordinal, typename Scalar> class MtxReader {
  std::string path_;

public:
  explicit MtxReader(const std::string &path) : path_(path) {}
  CooMat<Ordinal, Scalar> read_coo() const {
    std::ifstream is(path_);
    if (!is) THROW_RUNTIME("cannot open " << path_);
    std::string line;
    if (!std::getline(is, line) || line.rfind("%%MatrixMarket", 0) != 0)
      THROW_RUNTIME(path_ << ": missing MatrixMarket banner");
    const bool pattern = line.find("pattern") != std::string::npos;
    const bool symmetric = line.find("symmetric") != std::string::npos;
    if (line.find("coordinate") == std::string::npos) THROW_RUNTIME(path_ << ": only coordinate format");
    while (std::getline(is, line) && !line.empty() && line[0] == '%') {
    }
    long long m = 0, n = 0, nnz = -1;
    if (!(std::istringstream(line) >> m >> n >> nnz) || m <= 0 || n <= 0 || nnz < 0)
      THROW_RUNTIME(path_ << ": bad size line '" << line << "'");
    CooMat<Ordinal, Scalar> coo(m, n);
    for (long long k = 0; k < nnz; ++k) {
      long long i, j;
      double v = 1.0;
      is >> i >> j;
      if (!pattern) is >> v;
      if (!is) THROW_RUNTIME(path_ << ": truncated at entry " << k);
      if (i < 1 || i > m || j < 1 || j > n)
        THROW_RUNTIME(path_ << ": entry " << k << " (" << i << ", " << j << ") lies outside the " << m
                            << " x " << n << " matrix");
      coo.push_back(Ordinal(i - 1), Ordinal(j - 1), Scalar(v));
      if (symmetric && i != j) coo.push_back(Ordinal(j - 1), Ordinal(i - 1), Scalar(v));
    }
    return coo;
  }
  CsrMat<Ordinal, Scalar> read_csr() const { return CsrMat<Ordinal, Scalar>(read_coo()); }
};

} // namespace tenzing
