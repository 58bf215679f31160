// Shared boilerplate of the native drivers (reference: tenzing-dfs/examples/spmv.cu:41-80,
// tenzing-mcts/examples/spmv_run_strategy.cuh:28-75): init, communicator from the launcher's
// environment (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT, as set by torchrun or tzrun),
// one GPU per rank, reproducibility header on rank 0.
#pragma once

#include <cstdlib>
#include <iostream>
#include <memory>

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


} // namespace driver
} // namespace tenzing
