// Replay of a complete sequence as a CUDA graph.
//
// The reference's benchmark loop re-issues every op of a sequence from the host each iteration
// (src/benchmarker.cpp:93-97); on a B200 a kernel launch costs 5-8 us in a stream, which for a
// three-kernel schedule of a small problem is most of the iteration.  A sequence whose ops are
// all device work (kernels bound to streams, event records and stream waits; the host-side
// event synchronisation in front of Finish is replaced by the graph's own completion) can be
// captured once and replayed with one launch.
//
// Two graphs are captured, one per epoch parity, because the ops pick their staging buffers by
// (epoch & 1) on the host; the epoch VALUE, which the flag protocol needs fresh every time, is
// read by the kernels from a device word that the first node of each graph increments
// (tzb_set_epoch_source / tzb_epoch_advance in tenzing_b200.h).
#pragma once

#include <string>

#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/platform.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {

class GraphReplay {
  Platform &plat_;
  Sequence<BoundOp> seq_;
  cudaGraphExec_t exec_[2] = {nullptr, nullptr};
  size_t kernelNodes_[2] = {0, 0};
  uint64_t *epochDev_ = nullptr;
  uint64_t devEpoch_ = 0; // host mirror of *epochDev_

  void capture_one();

public:
  // can `seq` be replayed as a graph?  (why not, if not)
  static bool supported(const Sequence<BoundOp> &seq, std::string *why = nullptr);
  // captures the two graphs; the platform's events must be provisioned for seq.  Nothing runs.
  GraphReplay(const Sequence<BoundOp> &seq, Platform &plat);
  ~GraphReplay();
  GraphReplay(const GraphReplay &) = delete;
  GraphReplay &operator=(const GraphReplay &) = delete;

  // one execution under the platform's next epoch: a single cudaGraphLaunch on stream 0 of the
  // platform; does not synchronise (synchronise that stream, or record an event on it)
  void launch();
  cudaStream_t stream() const;
  // kernel nodes of one replay (for launch accounting)
  size_t kernel_nodes() const { return kernelNodes_[0]; }
};

} // namespace tenzing
