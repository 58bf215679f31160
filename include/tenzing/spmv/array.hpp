// Owning arrays in device or pinned host memory plus the NUMA-aware pinned staging pool.
//
// Reference: include/tenzing/spmv/array.hpp:19-195 (Array<Where,T>, ArrayView) and the dead
// src/numa.cpp.  ArrayView itself lives in csr_mat.hpp.
#pragma once

#include <cstddef>
#include <cstring>
#include <mutex>
#include <vector>

#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/numa.hpp"
#include "tenzing/spmv/csr_mat.hpp"

namespace tenzing {

template <Where W, typename T> class Array;

// device memory (cudaMalloc)
template <typename T> class Array<Where::device, T> {
  T *data_ = nullptr;
  int64_t size_ = 0;

public:
  Array() = default;
  explicit Array(int64_t n) { resize(n); }
  explicit Array(const std::vector<T> &host) {
    resize(int64_t(host.size()));
    if (size_) CUDA_RUNTIME(cudaMemcpy(data_, host.data(), size_t(size_) * sizeof(T), cudaMemcpyHostToDevice));
  }
  Array(const Array &) = delete;
  Array &operator=(const Array &) = delete;
  Array(Array &&o) noexcept : data_(o.data_), size_(o.size_) { o.data_ = nullptr, o.size_ = 0; }
  ~Array() {
    if (data_) cudaFree(data_);
  }
  void resize(int64_t n) {
    if (data_) CUDA_RUNTIME(cudaFree(data_));
    data_ = nullptr;
    size_ = n;
    if (n) CUDA_RUNTIME(cudaMalloc(reinterpret_cast<void **>(&data_), size_t(n) * sizeof(T)));
  }
  int64_t size() const { return size_; }
  T *data() const { return data_; }
  ArrayView<T> view() const { return ArrayView<T>(data_, size_); }
  std::vector<T> to_host() const {
    std::vector<T> h(static_cast<size_t>(size_));
    if (size_) CUDA_RUNTIME(cudaMemcpy(h.data(), data_, size_t(size_) * sizeof(T), cudaMemcpyDeviceToHost));
    return h;
  }
};

// pinned host memory (cudaHostAlloc)
template <typename T> class Array<Where::host, T> {
  T *data_ = nullptr;
  int64_t size_ = 0;

public:
  Array() = default;
  explicit Array(int64_t n) { resize(n); }
  Array(const Array &) = delete;
  Array &operator=(const Array &) = delete;
  Array(Array &&o) noexcept : data_(o.data_), size_(o.size_) { o.data_ = nullptr, o.size_ = 0; }
  ~Array() {
    if (data_) cudaFreeHost(data_);
  }
  void resize(int64_t n) {
    if (data_) CUDA_RUNTIME(cudaFreeHost(data_));
    data_ = nullptr;
    size_ = n;
    if (n) CUDA_RUNTIME(cudaHostAlloc(reinterpret_cast<void **>(&data_), size_t(n) * sizeof(T), cudaHostAllocDefault));
  }
  int64_t size() const { return size_; }
  T *data() const { return data_; }
  T &operator[](int64_t i) { return data_[i]; }
  const T &operator[](int64_t i) const { return data_[i]; }
  ArrayView<T> view() const { return ArrayView<T>(data_, size_); }
};

// Pinned staging buffers placed on the NUMA node next to the GPU: the calling thread is bound
// to the GPU's node first (numa::bind_to_local_memory), so the pages cudaHostAlloc pins are
// local to the PCIe root the GPU hangs off.  Buffers are recycled by size class.
class PinnedPool {
  struct Block {
    void *p;
    size_t bytes;
  };
  std::mutex mu_;
  std::vector<Block> free_;
  size_t allocated_ = 0;
  int node_ = -1;

public:
  explicit PinnedPool(int cudaDevice = -1) {
    if (cudaDevice < 0 && cudaGetDevice(&cudaDevice) != cudaSuccess) {
      (void)cudaGetLastError();
      return;
    }
    if (numa::bind_to_local_memory(cudaDevice)) {
      char busId[32] = {0};
      if (cudaDeviceGetPCIBusId(busId, sizeof(busId), cudaDevice) == cudaSuccess)
        node_ = numa::node_of_pci_device(busId);
    }
  }
  PinnedPool(const PinnedPool &) = delete;
  PinnedPool &operator=(const PinnedPool &) = delete;
  ~PinnedPool() {
    for (auto &b : free_) cudaFreeHost(b.p);
  }
  int numa_node() const { return node_; }
  size_t bytes_allocated() const { return allocated_; }
  void *get(size_t bytes) {
    size_t cls = 4096;
    while (cls < bytes) cls *= 2;
    {
      std::lock_guard<std::mutex> lk(mu_);
      for (size_t i = 0; i < free_.size(); ++i)
        if (free_[i].bytes == cls) {
          void *p = free_[i].p;
          free_.erase(free_.begin() + long(i));
          return p;
        }
    }
    void *p = nullptr;
    CUDA_RUNTIME(cudaHostAlloc(&p, cls, cudaHostAllocDefault));
    std::memset(p, 0, cls); // first touch on this (bound) thread
    allocated_ += cls;
    return p;
  }
  void put(void *p, size_t bytes) {
    size_t cls = 4096;
    while (cls < bytes) cls *= 2;
    std::lock_guard<std::mutex> lk(mu_);
    free_.push_back(Block{p, cls});
  }
};

} // namespace tenzing
