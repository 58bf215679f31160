// Body shared by the halo-exchange drivers (reference:
// tenzing-mcts/examples/halo_run_strategy.hpp:21-163).
#pragma once

#include "tenzing/argparse.hpp"
#include "tenzing/cuda/device_barrier.hpp"
#include "tenzing/driver.hpp"
#include "tenzing/halo/halo_exchange.hpp"

namespace halo_driver {

using namespace tenzing;

inline int main_impl(int argc, char **argv, const char *defaultStrategy) {
  driver::Context ctx = driver::start(argc, argv);
  int mctsIters = 300, benchIters = 50, n = 512, nQ = 3, nGhost = 3, streams = 2, neighbors = 26;
  bool noExpandRollout = false, qxyz = false;
  std::string strategy = defaultStrategy, style = "choice";
  argparse::Parser p("3-D halo exchange schedule search");
  p.add_option(mctsIters, "--mcts-iters", "-i");
  p.add_option(benchIters, "--benchmark-iters", "-b");
  p.add_option(n, "--n", "-n", "interior cells per dimension per rank (reference: 512)");
  p.add_option(nQ, "--nq", "", "quantities per cell (reference: 3)");
  p.add_option(nGhost, "--ghost", "", "ghost radius (reference: 3)");
  p.add_option(neighbors, "--neighbors", "", "6, 18 or 26");
  p.add_option(streams, "--streams", "");
  p.add_option(strategy, "--strategy", "-s");
  p.add_option(style, "--style", "", "choice|reference|staged|direct|fused_all|fused_staged");
  p.add_flag(noExpandRollout, "--no-expand-rollout");
  p.add_flag(qxyz, "--qxyz", "", "QXYZ storage order (default XYZQ as in the reference example)");
  driver::Resume resume;
  resume.add_options(p);
  driver::HostExec hostExec;
  hostExec.add_options(p);
  p.no_unrecognized();
  if (!p.parse(argc, argv)) {
    std::cerr << p.help();
    return 1;
  }
  hostExec.apply(ctx);
  try {
    HaloExchange::Args args;
    args.nQ = nQ, args.nX = args.nY = args.nZ = n, args.nGhost = nGhost;
    args.pitch = 128; // reference: halo_run_strategy.hpp:46
    args.neighbors = neighbors;
    args.storageOrder = qxyz ? HaloExchange::StorageOrder::QXYZ : HaloExchange::StorageOrder::XYZQ;
    args.set_default_topology(ctx.size);
    const auto mem = ctx.haveGpu ? SymmetricHeap::Memory::Device : SymmetricHeap::Memory::Host;
    auto dom = std::make_shared<HaloExchange::Domain>(ctx.comm, args, mem);
    Graph<OpBase> g;
    if (style == "choice") {
      OpPtr op = HaloExchange::make_choice(dom);
      g.start_then(op);
      g.then_finish(op);
    } else {
      HaloExchange::Style st = HaloExchange::Style::Staged;
      for (auto s : {HaloExchange::Style::Reference, HaloExchange::Style::Staged,
                     HaloExchange::Style::Direct, HaloExchange::Style::FusedAll,
                     HaloExchange::Style::FusedStaged, HaloExchange::Style::FusedGroups, HaloExchange::Style::FusedMixed, HaloExchange::Style::FusedPipelined, HaloExchange::Style::FusedPull})
        if (style == HaloExchange::style_name(s)) st = s;
      HaloExchange::add_to_graph(g, dom, {g.start()}, {g.finish()}, st);
    }
    Platform plat = Platform::make_n_streams(streams, ctx.comm,
                                             ctx.haveGpu ? Platform::Device::Cuda : Platform::Device::Null);
    if (ctx.haveGpu && ctx.size > 1 && std::getenv("TENZING_DEVICE_BARRIER")) plat.set_device_barrier(std::make_shared<DeviceBarrier>(ctx.comm));
    hostExec.apply(plat);
    DeviceBenchmarker device;
    std::unique_ptr<CachingBenchmarker> cached;
    Benchmarker &bench = resume.wrap(device, g, cached);
    mcts::Opts opts;
    opts.seed = uint64_t(resume.seed);
    opts.nIters = size_t(mctsIters);
    opts.expandRollout = !noExpandRollout;
    opts.benchOpts.nIters = size_t(benchIters);
    auto strat = mcts::make_strategy(strategy);
    mcts::Result r = mcts::explore(g, plat, bench, *strat, opts);
    if (ctx.rank == 0) r.dump_csv();
  } catch (const std::exception &e) {
    STDERR("fatal: " << e.what());
    return 2;
  }
  return 0;
}

} // namespace halo_driver
