// Shared boilerplate of the native drivers (reference: tenzing-dfs/examples/spmv.cu:41-80,
// tenzing-mcts/examples/spmv_run_strategy.cuh:28-75): init, communicator from the launcher's
// environment (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT, as set by torchrun or tzrun),
// one GPU per rank, reproducibility header on rank 0.
#pragma once

#include <cstdlib>
#include <iostream>
#include <memory>
#include <string>

#include "tenzing/tenzing.hpp"

namespace tenzing {
namespace driver {

struct Context {
  std::shared_ptr<Communicator> comm;
  int rank = 0, size = 1, device = -1;
  bool haveGpu = false;
};

inline Context start(int &argc, char **&argv) {
  tenzing::init(argc, argv);
  Context c;
  c.comm = SocketComm::from_env();
  if (!c.comm) c.comm = std::make_shared<SelfComm>();
  c.rank = c.comm->rank();
  c.size = c.comm->size();
  log::set_rank(c.rank);
  if (c.rank == 0) reproduce::dump_with_cli(argc, argv);
  int devcount = 0;
  if (cudaGetDeviceCount(&devcount) == cudaSuccess && devcount > 0) {
    const char *lr = std::getenv("LOCAL_RANK");
    c.device = (lr ? std::atoi(lr) : c.rank) % devcount;
    CUDA_RUNTIME(cudaSetDevice(c.device));
    numa::bind_to_local_memory(c.device);
    c.haveGpu = true;
  } else {
    (void)cudaGetLastError();
  }
  return c;
}

// --resume <csv> / --cache / --seed, shared by the search drivers: wraps the driver's
// benchmarker in a CachingBenchmarker (see benchmarker.hpp).  With --resume and the same
// --seed (and the same --cache setting) a search retraces the interrupted run from its CSV and
// continues measuring where that run stopped; the file must be visible to every rank.
// --host-exec [--host-streams]: no GPU, but run the ops for real on the CPU (host-memory heap,
// tenzing/host_exec.hpp); with --host-streams every stream is a worker thread.
struct HostExec {
  bool on = false, streams = false;
  template <typename Parser> void add_options(Parser &p) {
    p.add_flag(on, "--host-exec", "", "execute the ops on the CPU (host memory) instead of a GPU");
    p.add_flag(streams, "--host-streams", "", "with --host-exec: one worker thread per stream");
  }
  // call before the workload allocates: host execution needs the host-memory heap
  void apply(Context &ctx) const {
    if (on || streams) ctx.haveGpu = false;
  }
  void apply(Platform &plat) const {
    if (streams) plat.enable_host_streams();
    else if (on) plat.set_host_execution(true);
  }
};

struct Resume {
  std::string csv;  // results of an earlier run of the same command ("" = none)
  bool cache = false;
  int seed = 0;
  template <typename Parser> void add_options(Parser &p) {
    p.add_option(csv, "--resume", "", "results CSV of an interrupted run to continue from");
    p.add_flag(cache, "--cache", "", "never measure the same schedule twice");
    p.add_option(seed, "--seed", "", "seed of the search");
  }
  bool active() const { return cache || !csv.empty(); }
  // the benchmarker the solver should use; `holder` keeps the wrapper alive
  Benchmarker &wrap(Benchmarker &inner, const Graph<OpBase> &g,
                    std::unique_ptr<CachingBenchmarker> &holder) const {
    if (!active()) return inner;
    holder.reset(new CachingBenchmarker(inner, cache));
    if (!csv.empty()) {
      CsvBenchmarker prior(csv, g);
      holder->preload(prior);
      TENZING_INFO("resume: " << prior.rows().size() << " rows, " << holder->size()
                              << " distinct schedules from " << csv);
    }
    return *holder;
  }
};

} // namespace driver
} // namespace tenzing
