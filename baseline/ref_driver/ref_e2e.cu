// End-to-end step of the reference, through the reference's own public API (baseline only).
//
// bench.py's end-to-end number is: copy this step's input from pinned host memory to the device,
// run one iteration, read the result back.  The reference ships no driver that does that, so this
// is the program a user of its library would write: the set-up of its SpMV example
// (tenzing-dfs/examples/spmv.cu: same matrix, same RowPartSpmv / SpMV ops, same two-stream
// Platform), a short run of its own DFS solver with its own EmpiricalBenchmarker to pick a
// schedule (the best of the first schedules it enumerates), and then that schedule executed
// step by step with the x block copied in before and the y block copied out after each step.
// Nothing of tenzing_b200 is involved.
//
//   ref-e2e <steps> <warmup> <maxSeqs>       one process per rank (MPI stand-in), prints on rank 0:
//   REF_E2E median_us=<..> p10_us=<..> p90_us=<..> h2d_bytes=<..> d2h_bytes=<..> schedule_pct10_us=<..>
#include "tenzing/benchmarker.hpp"
#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/init.hpp"
#include "tenzing/spmv/csr_mat.hpp"
#include "tenzing/spmv/ops_spmv.cuh"
#include "tenzing/spmv/row_part_spmv.cuh"
#include "tenzing/spmv/where.hpp"

#include "tenzing/dfs/dfs.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

using Index = int;
using Value = float; // the reference's SpMV ops accept float only
using Clock = std::chrono::steady_clock;

namespace {
struct World {
  int rank = 0, size = 1;
};
World join_world(int &argc, char **&argv) {
  World w;
  MPI_Init(&argc, &argv);
  MPI_Comm_rank(MPI_COMM_WORLD, &w.rank);
  MPI_Comm_size(MPI_COMM_WORLD, &w.size);
  int nDev = 0;
  CUDA_RUNTIME(cudaGetDeviceCount(&nDev));
  CUDA_RUNTIME(cudaSetDevice(w.rank % nDev));
  CUDA_RUNTIME(cudaFree(0));
  return w;
}
} // namespace

int main(int argc, char **argv) {
  const int steps = argc > 1 ? std::atoi(argv[1]) : 20;
  const int warmup = argc > 2 ? std::atoi(argv[2]) : 5;
  const int maxSeqs = argc > 3 ? std::atoi(argv[3]) : 48;
  tenzing::init(argc, argv);
  const World world = join_world(argc, argv);
  const int rank = world.rank;

  // the problem of the reference's example: 150000 rows, 10 non-zeros per row, band m / ranks
  const int rowsTotal = 150000;
  CsrMat<Where::host, Index, Value> whole;
  if (rank == 0) whole = random_band_matrix<Index, Value>(rowsTotal, rowsTotal / world.size, rowsTotal * 10);
  RowPartSpmv<Index, Value> rps(whole, 0, MPI_COMM_WORLD);
  Graph<OpBase> program;
  {
    auto product = std::make_shared<SpMV<Index, Value>>(rps, MPI_COMM_WORLD);
    program.start_then(product);
    program.then_finish(product);
  }
  Platform platform = Platform::make_n_streams(2, MPI_COMM_WORLD);

  // a schedule: the reference's solver over the first schedules it enumerates
  EmpiricalBenchmarker timer;
  tenzing::dfs::Opts searchOpts;
  searchOpts.maxSeqs = maxSeqs;
  searchOpts.benchOpts.nIters = 10;
  const tenzing::dfs::Result result = tenzing::dfs::explore(program, platform, timer, searchOpts);
  if (result.simResults.empty()) {
    if (0 == rank) std::printf("REF_E2E none\n");
    MPI_Finalize();
    return 0;
  }
  size_t best = 0;
  for (size_t i = 1; i < result.simResults.size(); ++i)
    if (result.simResults[i].benchResult.pct10 < result.simResults[best].benchResult.pct10) best = i;
  Sequence<BoundOp> seq = result.simResults[best].seq;

  // events for that schedule (what the solver does before every benchmark)
  CudaEventPool pool;
  ResourceMap rmap;
  for (const auto &op : seq)
    if (HasEvent *he = dynamic_cast<HasEvent *>(op.get()))
      for (Event e : he->get_events())
        if (!rmap.contains(e)) rmap.insert(e, pool.new_event());
  platform.resource_map() = rmap;

  // pinned host copies of this rank's x and y blocks
  const size_t nx = size_t(rps.lx().size()), ny = size_t(rps.ly().size());
  Value *hx = nullptr, *hy = nullptr;
  CUDA_RUNTIME(cudaHostAlloc(reinterpret_cast<void **>(&hx), std::max<size_t>(nx, 1) * sizeof(Value), cudaHostAllocDefault));
  CUDA_RUNTIME(cudaHostAlloc(reinterpret_cast<void **>(&hy), std::max<size_t>(ny, 1) * sizeof(Value), cudaHostAllocDefault));
  for (size_t i = 0; i < nx; ++i) hx[i] = Value(1) + Value(i % 7);
  cudaStream_t io;
  CUDA_RUNTIME(cudaStreamCreate(&io));
  Value *dx = const_cast<Value *>(rps.lx().data());
  const Value *dy = rps.ly().data();

  std::vector<double> t;
  for (int s = 0; s < warmup + steps; ++s) {
    CUDA_RUNTIME(cudaDeviceSynchronize());
    MPI_Barrier(MPI_COMM_WORLD);
    const double t0 = MPI_Wtime();
    CUDA_RUNTIME(cudaMemcpyAsync(dx, hx, nx * sizeof(Value), cudaMemcpyHostToDevice, io));
    CUDA_RUNTIME(cudaStreamSynchronize(io));
    for (auto &op : seq) op->run(platform);
    CUDA_RUNTIME(cudaMemcpyAsync(hy, dy, ny * sizeof(Value), cudaMemcpyDeviceToHost, io));
    CUDA_RUNTIME(cudaStreamSynchronize(io));
    double dt = MPI_Wtime() - t0;
    MPI_Allreduce(MPI_IN_PLACE, &dt, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);
    if (s >= warmup) t.push_back(dt);
  }
  std::sort(t.begin(), t.end());
  if (0 == rank && !t.empty())
    std::printf("REF_E2E median_us=%.3f p10_us=%.3f p90_us=%.3f h2d_bytes=%zu d2h_bytes=%zu schedule_pct10_us=%.3f\n",
                t[t.size() / 2] * 1e6, t[t.size() / 10] * 1e6, t[t.size() * 9 / 10] * 1e6, nx * sizeof(Value),
                ny * sizeof(Value), result.simResults[best].benchResult.pct10 * 1e6);
  std::fflush(stdout);
  MPI_Finalize();
  return 0;
}
