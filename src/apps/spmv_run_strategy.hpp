// Body shared by the SpMV drivers (reference: tenzing-mcts/examples/spmv_run_strategy.cuh:28-133,
// tenzing-dfs/examples/spmv.cu:41-123).
#pragma once

#include "tenzing/argparse.hpp"
#include "tenzing/cuda/device_barrier.hpp"
#include "tenzing/driver.hpp"
#include "tenzing/spmv/row_part_spmv.hpp"

namespace spmv_driver {

using namespace tenzing;

struct Config {
  int mctsIters = 300;      // reference: spmv_run_strategy.cuh:44
  int benchIters = 50;      // :45
  int m = 150000;           // :47
  int nnzPerRow = 10;
  int streams = 2;
  int maxSeqs = 15000;      // tenzing-dfs/examples/spmv.cu:116
  bool noExpandRollout = false;
  bool random = false;      // bw = m instead of m/size
  bool fp32 = false;
  std::string style = "choice";
  std::string strategy;
  driver::Resume resume;
  driver::HostExec hostExec;
};

inline bool parse(Config &c, int &argc, char **argv, bool mcts) {
  argparse::Parser p("distributed SpMV schedule search");
  if (mcts) p.add_option(c.mctsIters, "--mcts-iters", "-i", "MCTS iterations");
  if (mcts) p.add_flag(c.noExpandRollout, "--no-expand-rollout", "", "do not keep rollout nodes");
  if (mcts) p.add_option(c.strategy, "--strategy", "-s", "FastMin|Coverage|Random|...");
  if (!mcts) p.add_option(c.maxSeqs, "--max-seqs", "", "stop enumerating after this many sequences");
  p.add_option(c.benchIters, "--benchmark-iters", "-b", "measurements per sequence");
  p.add_option(c.m, "--matrix-m", "-m", "rows of the matrix");
  p.add_option(c.nnzPerRow, "--nnz-per-row", "", "non-zeros per row");
  p.add_option(c.streams, "--streams", "", "CUDA streams per rank");
  p.add_option(c.style, "--style", "", "choice|reference|fused|fused_separate");
  p.add_flag(c.random, "--random", "", "uniformly random columns instead of a band");
  p.add_flag(c.fp32, "--fp32", "", "fp32 values (the reference's type) instead of fp64");
  c.resume.add_options(p);
  c.hostExec.add_options(p);
  p.no_unrecognized();
  if (!p.parse(argc, argv)) {
    std::cerr << p.help();
    return false;
  }
  return true;
}

template <typename Scalar>
Graph<OpBase> make_graph(const std::shared_ptr<RowPartSpmv<Scalar>> &spmv, const std::string &style) {
  using Style = typename RowPartSpmv<Scalar>::Style;
  Graph<OpBase> g;
  OpPtr op;
  if (style == "choice") op = make_spmv_choice<Scalar>(spmv);
  else if (style == "reference") op = std::make_shared<SpMV<Scalar>>("spmv", spmv, Style::Reference);
  else if (style == "fused") op = std::make_shared<SpMV<Scalar>>("spmv", spmv, Style::Fused);
  else if (style == "fused_separate") op = std::make_shared<SpMV<Scalar>>("spmv", spmv, Style::FusedSeparate);
  else if (style == "fused_all") op = std::make_shared<SpMV<Scalar>>("spmv", spmv, Style::FusedAll);
  else if (style == "fused_pull") op = std::make_shared<SpMV<Scalar>>("spmv", spmv, Style::FusedPull);
  else if (style == "choice+") op = make_spmv_choice<Scalar>(spmv, true); // with experimental styles
  else THROW_RUNTIME("unknown style " << style);
  g.start_then(op);
  g.then_finish(op);
  return g;
}

template <typename Scalar> int run(const Config &c, driver::Context &ctx, bool mcts) {
  c.hostExec.apply(ctx);
  const auto mem = ctx.haveGpu ? SymmetricHeap::Memory::Device : SymmetricHeap::Memory::Host;
  const int64_t bw = c.random ? c.m : std::max(1, c.m / ctx.size); // reference: spmv.cu:87-88
  auto spmv = make_band_spmv<Scalar>(ctx.comm, c.m, bw, c.nnzPerRow, 0, mem);
  Graph<OpBase> g = make_graph<Scalar>(spmv, c.style);
  Platform plat = Platform::make_n_streams(c.streams, ctx.comm,
                                           ctx.haveGpu ? Platform::Device::Cuda : Platform::Device::Null);
  if (ctx.haveGpu && ctx.size > 1 && std::getenv("TENZING_DEVICE_BARRIER")) plat.set_device_barrier(std::make_shared<DeviceBarrier>(ctx.comm));
  c.hostExec.apply(plat);
  DeviceBenchmarker device;
  std::unique_ptr<CachingBenchmarker> cached;
  Benchmarker &bench = c.resume.wrap(device, g, cached);
  Benchmark::Opts bo;
  bo.nIters = size_t(c.benchIters);
  if (mcts) {
    mcts::Opts opts;
    opts.seed = uint64_t(c.resume.seed);
    opts.nIters = size_t(c.mctsIters);
    opts.expandRollout = !c.noExpandRollout;
    opts.benchOpts = bo;
    auto strategy = mcts::make_strategy(c.strategy.empty() ? "FastMin" : c.strategy);
    mcts::Result r = mcts::explore(g, plat, bench, *strategy, opts);
    if (ctx.rank == 0) r.dump_csv();
  } else {
    dfs::Opts opts;
    opts.maxSeqs = c.maxSeqs;
    opts.benchOpts = bo;
    dfs::Result r = dfs::explore(g, plat, bench, opts);
    if (ctx.rank == 0) r.dump_csv();
  }
  return 0;
}

inline int main_impl(int argc, char **argv, bool mcts, const char *defaultStrategy) {
  driver::Context ctx = driver::start(argc, argv);
  Config c;
  c.strategy = defaultStrategy ? defaultStrategy : "";
  if (!parse(c, argc, argv, mcts)) return 1;
  try {
    return c.fp32 ? run<float>(c, ctx, mcts) : run<double>(c, ctx, mcts);
  } catch (const std::exception &e) {
    STDERR("fatal: " << e.what());
    return 2;
  }
}

} // namespace spmv_driver
