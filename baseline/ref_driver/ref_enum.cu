// Search-engine throughput of the reference on the CPU (baseline only; no GPU needed).
//
// BASELINE.json config 1 is "noop-graph DFS schedule enumeration on CPU".  The reference's test
// (test/test_noop_graph.cpp) has one op; this program times the same public entry point,
// tenzing::dfs::get_all_sequences, on noop graphs large enough to time: K independent chains of
// L NoOps between Start and Finish, (K*L)! / (L!)^K schedules.  No streams, no CUDA call.
//
//   ref-enum K L        prints:  REF_ENUM chains=K len=L sequences=<n> seconds=<t>
#include "tenzing/graph.hpp"
#include "tenzing/init.hpp"
#include "tenzing/operation.hpp"
#include "tenzing/platform.hpp"

#include "tenzing/dfs/dfs.hpp"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <string>

int main(int argc, char **argv) {
  const int K = argc > 1 ? std::atoi(argv[1]) : 3;
  const int L = argc > 2 ? std::atoi(argv[2]) : 2;
  tenzing::init(argc, argv);
  MPI_Init(&argc, &argv);
  {
    Platform plat(MPI_COMM_WORLD);
    Graph<OpBase> g;
    for (int k = 0; k < K; ++k) {
      std::shared_ptr<NoOp> prev;
      for (int l = 0; l < L; ++l) {
        auto op = std::make_shared<NoOp>("op" + std::to_string(k) + "_" + std::to_string(l));
        if (prev) g.then(prev, op);
        else g.start_then(op);
        prev = op;
      }
      g.then_finish(prev);
    }
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<Sequence<BoundOp>> seqs = tenzing::dfs::get_all_sequences(g, plat);
    const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("REF_ENUM chains=%d len=%d sequences=%zu seconds=%.6f\n", K, L, seqs.size(), dt);
  }
  MPI_Finalize();
  return 0;
}
