// Message-size sweep 1 KB - 256 MB between neighbouring ranks (BASELINE.json config 5):
// the fused path (one kernel copies the payload into the peer's buffer and publishes a flag;
// the peer's kernel waits for it) against the host-driven stand-in for CUDA-aware MPI
// (cudaMemcpyAsync peer copy + flag kernel, host wait).  Every rank sends to (rank+1) % size
// and receives from (rank-1+size) % size.  Times are device-timed, max over ranks.
#include <algorithm>
#include <cstdio>
#include <vector>

#include "tenzing/argparse.hpp"
#include "tenzing/cuda/device_barrier.hpp"
#include "tenzing/cuda/symmetric_heap.hpp"
#include "tenzing/driver.hpp"
#include "tenzing_b200.h"

using namespace tenzing;

int main(int argc, char **argv) {
  driver::Context ctx = driver::start(argc, argv);
  size_t maxBytes = size_t(256) << 20;
  int reps = 20;
  argparse::Parser p("halo message-size sweep");
  p.add_option(maxBytes, "--max-bytes", "");
  p.add_option(reps, "--reps", "");
  if (!p.parse(argc, argv) || !ctx.haveGpu) {
    std::fprintf(stderr, "%s(needs a GPU)\n", p.help().c_str());
    return 1;
  }
  try {
    SymmetricHeap heap(ctx.comm, 2 * maxBytes + 4096);
    const size_t srcOff = heap.alloc(maxBytes), dstOff = heap.alloc(maxBytes), flagOff = heap.alloc(256);
    const int next = (ctx.rank + 1) % ctx.size;
    cudaStream_t s;
    CUDA_RUNTIME(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CUDA_RUNTIME(cudaEventCreate(&e0));
    CUDA_RUNTIME(cudaEventCreate(&e1));
    void *devPlan = nullptr, *scratch = nullptr;
    CUDA_RUNTIME(cudaMalloc(&devPlan, 1 << 20));
    CUDA_RUNTIME(cudaMalloc(&scratch, 256));
    CUDA_RUNTIME(cudaMemset(scratch, 0, 256));
    uint64_t epoch = 0;
    // ranks start together on the GPUs (a host barrier alone leaves 15-25 us of skew, which at
    // 1 KB is the whole measurement), with the step's kernels already queued behind a short delay
    DeviceBarrier startLine(ctx.comm);
    if (ctx.rank == 0) std::printf("bytes,fused_us,fused_GBs,memcpy_us,memcpy_GBs\n");
    for (size_t bytes = 1024; bytes <= maxBytes; bytes *= 4) {
      // one contiguous region expressed as rows of 4 KB
      TzbBox box{};
      box.src = heap.at<char>(ctx.rank, srcOff);
      box.dst = heap.at<char>(next, dstOff);
      box.run = int(std::min<size_t>(bytes / 8, 512));
      box.n1 = int(bytes / 8 / size_t(box.run));
      box.n2 = box.n3 = 1;
      box.s1 = box.d1 = box.run;
      box.signal = heap.at<uint64_t>(next, flagOff);
      TzbBoxPlanInfo info;
      CUDA_RUNTIME(cudaError_t(tzb_boxcopy_plan(&box, 1, 8, &info, nullptr)));
      std::vector<char> image(info.planBytes);
      CUDA_RUNTIME(cudaError_t(tzb_boxcopy_plan(&box, 1, 8, &info, image.data())));
      if (info.planBytes > (1 << 20)) THROW_RUNTIME("plan too large");
      CUDA_RUNTIME(cudaMemcpy(devPlan, image.data(), info.planBytes, cudaMemcpyHostToDevice));
      double best[2] = {1e30, 1e30};
      for (int mode = 0; mode < 2; ++mode) {
        for (int r = 0; r < reps + 3; ++r) {
          ++epoch;
          CUDA_RUNTIME(cudaStreamSynchronize(s));
          ctx.comm->barrier();
          CUDA_RUNTIME(cudaError_t(tzb_delay_us(40, s)));
          if (ctx.size > 1) startLine.enqueue(s);
          CUDA_RUNTIME(cudaEventRecord(e0, s));
          if (mode == 0) {
            CUDA_RUNTIME(cudaError_t(tzb_boxcopy_launch(devPlan, &info, epoch, scratch, s)));
          } else {
            CUDA_RUNTIME(cudaMemcpyAsync(box.dst, box.src, bytes, cudaMemcpyDeviceToDevice, s));
            TzbFlagList l;
            l.n = 1;
            l.flag[0] = box.signal;
            CUDA_RUNTIME(cudaError_t(tzb_flags_signal(&l, epoch, s)));
          }
          CUDA_RUNTIME(cudaError_t(tzb_flags_wait(heap.at<uint64_t>(ctx.rank, flagOff), 1, epoch, s)));
          CUDA_RUNTIME(cudaEventRecord(e1, s));
          CUDA_RUNTIME(cudaEventSynchronize(e1));
          float ms = 0;
          CUDA_RUNTIME(cudaEventElapsedTime(&ms, e0, e1));
          double t = ctx.comm->max_over_ranks(double(ms) * 1e-3);
          if (r >= 3) best[mode] = std::min(best[mode], t);
        }
      }
      if (ctx.rank == 0)
        std::printf("%zu,%.2f,%.1f,%.2f,%.1f\n", bytes, best[0] * 1e6, bytes / best[0] / 1e9,
                    best[1] * 1e6, bytes / best[1] / 1e9);
    }
    ctx.comm->barrier();
  } catch (const std::exception &e) {
    STDERR("fatal: " << e.what());
    return 2;
  }
  return 0;
}
