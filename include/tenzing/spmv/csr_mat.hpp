// Host-side sparse matrix containers, generators, partitioning and the
// local/remote column split used by the distributed SpMV workload.
//
// Reference: include/tenzing/spmv/{coo_mat,csr_mat,partition,split_mat,array,where}.hpp.
// Differences: generators are seeded and can produce just a block of rows
// (every rank builds its own block; no matrix scatter from a root), values are
// not all 1.0 so numerics tests mean something, and split_local_remote
// relabels remote columns once (the reference relabels twice).
#pragma once

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <map>
#include <numeric>
#include <vector>

#include "tenzing/macro_at.hpp"

namespace tenzing {

enum class Where { host, device };

// non-owning view usable on host (and on device when compiled by nvcc)
template <typename T> struct ArrayView {
  T *data_ = nullptr;
  int64_t size_ = 0;
  ArrayView() = default;
  ArrayView(T *d, int64_t n) : data_(d), size_(n) {}
#ifdef __CUDACC__
  __host__ __device__
#endif
  T &operator()(int64_t i) const {
#ifdef TENZING_RUNTIME_CHECK_BOUNDS
    // reference: spmv/array.hpp:36-43 (always on there; opt-in here, it costs registers)
    if (i < 0 || i >= size_) printf("ArrayView: index %lld out of range [0, %lld)\n", (long long)i, (long long)size_);
#endif
    return data_[i];
  }
#ifdef __CUDACC__
  __host__ __device__
#endif
  int64_t size() const { return size_; }
  T *data() const { return data_; }
};

template <typename Ordinal, typename Scalar> class CooMat {
public:
  struct Entry {
    Ordinal i, j;
    Scalar e;
  };

private:
  int64_t numRows_, numCols_;
  std::vector<Entry> data_;

public:
  CooMat(int64_t m, int64_t n) : numRows_(m), numCols_(n) {}
  int64_t num_rows() const { return numRows_; }
  int64_t num_cols() const { return numCols_; }
  int64_t nnz() const { return int64_t(data_.size()); }
  void push_back(Ordinal i, Ordinal j, Scalar e) { data_.push_back(Entry{i, j, e}); }
  const std::vector<Entry> &entries() const { return data_; }
  // sort by (row, col) and drop duplicate coordinates (first value wins)
  void sort_and_dedup() {
    std::sort(data_.begin(), data_.end(), [](const Entry &a, const Entry &b) {
      return a.i != b.i ? a.i < b.i : a.j < b.j;
    });
    data_.erase(std::unique(data_.begin(), data_.end(),
                            [](const Entry &a, const Entry &b) { return a.i == b.i && a.j == b.j; }),
                data_.end());
  }
};

// host CSR
template <typename Ordinal, typename Scalar> class CsrMat {
public:
  std::vector<Ordinal> rowPtr_; // num_rows + 1
  std::vector<Ordinal> colInd_;
  std::vector<Scalar> val_;
  int64_t numCols_ = 0;

  CsrMat() : rowPtr_(1, 0) {}
  CsrMat(int64_t rows, int64_t cols) : rowPtr_(size_t(rows) + 1, 0), numCols_(cols) {}
  explicit CsrMat(CooMat<Ordinal, Scalar> coo) : numCols_(coo.num_cols()) {
    coo.sort_and_dedup();
    rowPtr_.assign(size_t(coo.num_rows()) + 1, 0);
    for (const auto &e : coo.entries()) {
      ++rowPtr_[size_t(e.i) + 1];
      colInd_.push_back(e.j);
      val_.push_back(e.e);
    }
    std::partial_sum(rowPtr_.begin(), rowPtr_.end(), rowPtr_.begin());
  }
  int64_t num_rows() const { return int64_t(rowPtr_.size()) - 1; }
  int64_t num_cols() const { return numCols_; }
  int64_t nnz() const { return int64_t(colInd_.size()); }
  const Ordinal *row_ptr() const { return rowPtr_.data(); }
  const Ordinal *col_ind() const { return colInd_.data(); }
  const Scalar *val() const { return val_.data(); }

  // keep rows [begin, end), renumbered from 0 (columns unchanged)
  CsrMat retain_rows(int64_t begin, int64_t end) const {
    CsrMat r(end - begin, numCols_);
    const Ordinal lo = rowPtr_[size_t(begin)], hi = rowPtr_[size_t(end)];
    r.colInd_.assign(colInd_.begin() + lo, colInd_.begin() + hi);
    r.val_.assign(val_.begin() + lo, val_.begin() + hi);
    for (int64_t i = begin; i <= end; ++i) r.rowPtr_[size_t(i - begin)] = rowPtr_[size_t(i)] - lo;
    return r;
  }
  // y = A x on the host (the numerics oracle)
  std::vector<Scalar> multiply(const std::vector<Scalar> &x) const {
    std::vector<Scalar> y(size_t(num_rows()), Scalar(0));
    for (int64_t i = 0; i < num_rows(); ++i) {
      Scalar acc = 0;
      for (Ordinal k = rowPtr_[size_t(i)]; k < rowPtr_[size_t(i) + 1]; ++k)
        acc += val_[size_t(k)] * x[size_t(colInd_[size_t(k)])];
      y[size_t(i)] = acc;
    }
    return y;
  }
};

// ---- partitioning (reference: partition.hpp:10-75): remainder goes to the low ranks ----
struct Range {
  int64_t lb = 0, ub = 0; // [lb, ub)
  int64_t extent() const { return ub - lb; }
};
inline Range get_partition(int64_t domain, int i, int n) {
  const int64_t div = domain / n, rem = domain % n;
  Range r;
  r.lb = (i < rem) ? int64_t(i) * (div + 1) : rem * (div + 1) + (int64_t(i) - rem) * div;
  r.ub = r.lb + ((i < rem) ? div + 1 : div);
  return r;
}
inline int get_owner(int64_t domain, int64_t idx, int n) {
  const int64_t div = domain / n, rem = domain % n;
  if (idx < rem * (div + 1)) return int(idx / (div + 1));
  return int(rem + (idx - rem * (div + 1)) / std::max<int64_t>(div, 1));
}
template <typename Ordinal, typename Scalar>
std::vector<CsrMat<Ordinal, Scalar>> part_by_rows(const CsrMat<Ordinal, Scalar> &m, int parts) {
  std::vector<CsrMat<Ordinal, Scalar>> r;
  for (int p = 0; p < parts; ++p) {
    const Range rg = get_partition(m.num_rows(), p, parts);
    r.push_back(m.retain_rows(rg.lb, rg.ub));
  }
  return r;
}

// ---- generators ----
// counter-based hash -> reproducible per (seed, row, k) regardless of who generates the row
inline uint64_t mix64(uint64_t x) {
  x += 0x9e3779b97f4a7c15ULL;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ULL;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebULL;
  return x ^ (x >> 31);
}
inline uint64_t hash3(uint64_t seed, uint64_t a, uint64_t b) {
  return mix64(seed ^ mix64(a ^ mix64(b + 0x632be59bd9b4e019ULL)));
}

// Rows [rowBegin, rowEnd) of an n x n matrix with ~nnzPerRow entries per row
// whose columns lie in [row-bw, row+bw] (clipped).  bw >= n gives a uniformly
// random sparse matrix.  The block has rowEnd-rowBegin rows and n columns
// (global column ids).  Values are in [0.5, 1.5).
template <typename Ordinal, typename Scalar>
CsrMat<Ordinal, Scalar> random_band_rows(int64_t n, int64_t bw, int nnzPerRow, int64_t rowBegin,
                                         int64_t rowEnd, uint64_t seed = 0) {
  CsrMat<Ordinal, Scalar> m(rowEnd - rowBegin, n);
  m.colInd_.reserve(size_t((rowEnd - rowBegin) * nnzPerRow));
  m.val_.reserve(size_t((rowEnd - rowBegin) * nnzPerRow));
  std::vector<Ordinal> cols;
  for (int64_t i = rowBegin; i < rowEnd; ++i) {
    const int64_t lo = std::max<int64_t>(0, i - bw), hi = std::min<int64_t>(n - 1, i + bw);
    const uint64_t span = uint64_t(hi - lo + 1);
    cols.clear();
    for (int k = 0; k < nnzPerRow; ++k) cols.push_back(Ordinal(lo + int64_t(hash3(seed, uint64_t(i), uint64_t(k)) % span)));
    std::sort(cols.begin(), cols.end());
    cols.erase(std::unique(cols.begin(), cols.end()), cols.end());
    for (Ordinal c : cols) {
      m.colInd_.push_back(c);
      const uint64_t h = hash3(seed ^ 0xabcdefULL, uint64_t(i), uint64_t(c));
      m.val_.push_back(Scalar(0.5 + double(h >> 11) * (1.0 / 9007199254740992.0)));
    }
    m.rowPtr_[size_t(i - rowBegin) + 1] = Ordinal(m.colInd_.size());
  }
  return m;
}

// whole-matrix forms with the reference's signatures (csr_mat.hpp:301-370):
// nnz is the total entry budget
template <typename Ordinal, typename Scalar>
CsrMat<Ordinal, Scalar> random_band_matrix(int64_t n, int64_t bw, int64_t nnz, uint64_t seed = 0) {
  const int perRow = int(std::max<int64_t>(1, nnz / std::max<int64_t>(n, 1)));
  return random_band_rows<Ordinal, Scalar>(n, bw, perRow, 0, n, seed);
}
template <typename Ordinal, typename Scalar>
CsrMat<Ordinal, Scalar> random_matrix(int64_t m, int64_t n, int64_t nnz, uint64_t seed = 0) {
  if (m != n) THROW_RUNTIME("random_matrix: only square matrices are generated");
  return random_band_matrix<Ordinal, Scalar>(n, n, nnz, seed);
}

// ---- local / remote split (reference: split_mat.hpp:22-136) ----
template <typename Ordinal, typename Scalar> struct SplitMat {
  int64_t loff = 0;                 // global column of local column 0
  CsrMat<Ordinal, Scalar> local;    // columns shifted to 0
  CsrMat<Ordinal, Scalar> remote;   // columns compacted to 0..N in ascending global order
  std::vector<Ordinal> globals;     // remote column -> global column (ascending)
};

// `m` holds the rows owned by this rank with GLOBAL column ids; columns in
// [colBegin, colEnd) are local
template <typename Ordinal, typename Scalar>
SplitMat<Ordinal, Scalar> split_local_remote(const CsrMat<Ordinal, Scalar> &m, int64_t colBegin,
                                             int64_t colEnd) {
  SplitMat<Ordinal, Scalar> s;
  s.loff = colBegin;
  // ascending global order => entries from rank 0 come first, then rank 1, ...
  for (Ordinal c : m.colInd_)
    if (c < colBegin || c >= colEnd) s.globals.push_back(c);
  std::sort(s.globals.begin(), s.globals.end());
  s.globals.erase(std::unique(s.globals.begin(), s.globals.end()), s.globals.end());
  s.local = CsrMat<Ordinal, Scalar>(m.num_rows(), colEnd - colBegin);
  s.remote = CsrMat<Ordinal, Scalar>(m.num_rows(), int64_t(s.globals.size()));
  for (int64_t i = 0; i < m.num_rows(); ++i) {
    for (Ordinal k = m.rowPtr_[size_t(i)]; k < m.rowPtr_[size_t(i) + 1]; ++k) {
      const Ordinal c = m.colInd_[size_t(k)];
      if (c >= colBegin && c < colEnd) {
        s.local.colInd_.push_back(Ordinal(c - colBegin));
        s.local.val_.push_back(m.val_[size_t(k)]);
      } else {
        const auto it = std::lower_bound(s.globals.begin(), s.globals.end(), c);
        s.remote.colInd_.push_back(Ordinal(it - s.globals.begin()));
        s.remote.val_.push_back(m.val_[size_t(k)]);
      }
    }
    s.local.rowPtr_[size_t(i) + 1] = Ordinal(s.local.colInd_.size());
    s.remote.rowPtr_[size_t(i) + 1] = Ordinal(s.remote.colInd_.size());
  }
  return s;
}

} // namespace tenzing
