#include "tenzing/cuda/graph_replay.hpp"

#include <vector>

#include "tenzing/cuda/ops_cuda.hpp"
#include "tenzing/macro_at.hpp"
#include "tenzing/operation.hpp"
#include "tenzing_b200.h"

namespace tenzing {

namespace {
// host ops a replay leaves out: the graph launch orders everything behind Start, and waiting
// for the replay to finish is the caller's stream synchronisation
bool is_skipped(const BoundOp &op) {
  return dynamic_cast<const Start *>(&op) || dynamic_cast<const Finish *>(&op) ||
         dynamic_cast<const CudaEventSync *>(&op);
}
bool is_device_work(const BoundOp &op) {
  return dynamic_cast<const BoundGpuOp *>(&op) || dynamic_cast<const CudaEventRecord *>(&op) ||
         dynamic_cast<const CudaStreamWaitEvent *>(&op);
}
} // namespace

bool GraphReplay::supported(const Sequence<BoundOp> &seq, std::string *why) {
  bool hostSyncSeen = false;
  for (const auto &op : seq) {
    if (dynamic_cast<const CudaEventSync *>(op.get())) {
      hostSyncSeen = true;
      continue;
    }
    if (dynamic_cast<const Start *>(op.get()) || dynamic_cast<const Finish *>(op.get())) continue;
    if (!is_device_work(*op)) {
      if (why) *why = "op '" + op->name() + "' runs on the host";
      return false;
    }
    // a host synchronisation in the MIDDLE means some host op needs device results
    if (hostSyncSeen && dynamic_cast<const BoundGpuOp *>(op.get())) {
      if (why) *why = "kernel '" + op->name() + "' is issued after a host synchronisation";
      return false;
    }
  }
  return true;
}

GraphReplay::GraphReplay(const Sequence<BoundOp> &seq, Platform &plat) : plat_(plat), seq_(seq) {
  if (plat.is_null()) THROW_RUNTIME("graph replay needs a CUDA platform");
  std::string why;
  if (!supported(seq, &why)) THROW_RUNTIME("sequence cannot be replayed as a CUDA graph: " << why);
  CUDA_RUNTIME(cudaMalloc(reinterpret_cast<void **>(&epochDev_), sizeof(uint64_t)));
  const uint64_t saved = plat.epoch();
  plat.synchronize();
  try {
    capture_one();
    capture_one();
  } catch (...) {
    tzb_set_epoch_source(nullptr);
    plat.set_epoch(saved);
    for (auto &e : exec_)
      if (e) cudaGraphExecDestroy(e);
    cudaFree(epochDev_);
    throw;
  }
  // the captures advanced the host epoch twice without running anything
  plat.set_epoch(saved);
  devEpoch_ = saved;
  CUDA_RUNTIME(cudaMemcpy(epochDev_, &saved, sizeof(saved), cudaMemcpyHostToDevice));
}

void GraphReplay::capture_one() {
  const int ns = plat_.num_streams();
  cudaStream_t s0 = plat_.cuda_stream(plat_.streams_[0]);
  std::vector<cudaEvent_t> ev(size_t(ns), nullptr);
  for (auto &e : ev) CUDA_RUNTIME(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  cudaGraph_t graph = nullptr;
  tzb_set_epoch_source(epochDev_);
  cudaError_t err = cudaStreamBeginCapture(s0, cudaStreamCaptureModeThreadLocal);
  if (err == cudaSuccess) {
    try {
      CUDA_RUNTIME(cudaError_t(tzb_epoch_advance(epochDev_, s0)));
      // every other stream joins the capture behind the epoch update ...
      CUDA_RUNTIME(cudaEventRecord(ev[0], s0));
      for (int k = 1; k < ns; ++k)
        CUDA_RUNTIME(cudaStreamWaitEvent(plat_.cuda_stream(plat_.streams_[size_t(k)]), ev[0], 0));
      plat_.next_epoch();
      for (auto &op : seq_)
        if (!is_skipped(*op)) op->run(plat_);
      // ... and is joined back before the capture ends
      for (int k = 1; k < ns; ++k) {
        CUDA_RUNTIME(cudaEventRecord(ev[size_t(k)], plat_.cuda_stream(plat_.streams_[size_t(k)])));
        CUDA_RUNTIME(cudaStreamWaitEvent(s0, ev[size_t(k)], 0));
      }
    } catch (...) {
      cudaGraph_t dead = nullptr;
      cudaStreamEndCapture(s0, &dead);
      if (dead) cudaGraphDestroy(dead);
      tzb_set_epoch_source(nullptr);
      for (auto &e : ev) cudaEventDestroy(e);
      (void)cudaGetLastError();
      throw;
    }
    err = cudaStreamEndCapture(s0, &graph);
  }
  tzb_set_epoch_source(nullptr);
  for (auto &e : ev) cudaEventDestroy(e);
  if (err != cudaSuccess || !graph) {
    (void)cudaGetLastError();
    THROW_RUNTIME("stream capture failed: " << cudaGetErrorString(err));
  }
  const size_t parity = size_t(plat_.epoch() & 1);
  size_t n = 0;
  CUDA_RUNTIME(cudaGraphGetNodes(graph, nullptr, &n));
  std::vector<cudaGraphNode_t> nodes(n);
  if (n) CUDA_RUNTIME(cudaGraphGetNodes(graph, nodes.data(), &n));
  kernelNodes_[parity] = 0;
  for (auto node : nodes) {
    cudaGraphNodeType t;
    CUDA_RUNTIME(cudaGraphNodeGetType(node, &t));
    if (t == cudaGraphNodeTypeKernel) ++kernelNodes_[parity];
  }
  if (kernelNodes_[parity]) --kernelNodes_[parity]; // the epoch update is bookkeeping, not work
  err = cudaGraphInstantiate(&exec_[parity], graph, 0);
  cudaGraphDestroy(graph);
  if (err != cudaSuccess) {
    (void)cudaGetLastError();
    THROW_RUNTIME("cudaGraphInstantiate failed: " << cudaGetErrorString(err));
  }
}

GraphReplay::~GraphReplay() {
  for (auto &e : exec_)
    if (e) cudaGraphExecDestroy(e);
  if (epochDev_) cudaFree(epochDev_);
}

cudaStream_t GraphReplay::stream() const { return plat_.cuda_stream(plat_.streams_[0]); }

void GraphReplay::launch() {
  cudaStream_t s0 = stream();
  if (devEpoch_ != plat_.epoch()) {
    // the sequence (or another one) ran the ordinary way in between: bring the device word up to date
    devEpoch_ = plat_.epoch();
    CUDA_RUNTIME(cudaError_t(tzb_epoch_set(epochDev_, devEpoch_, s0)));
  }
  const uint64_t e = plat_.next_epoch();
  CUDA_RUNTIME(cudaGraphLaunch(exec_[size_t(e & 1)], s0));
  devEpoch_ = e;
}

} // namespace tenzing
