// Native CPU test runner (reference: test/test_noop_graph.cpp, the in-source doctest cases of
// src/graph.cpp:422-501, src/operation.cpp:87-101, src/sequence.cpp:169-176, and
// test/test_gpu_graph.cu run here on a Null platform).  Exit code = number of failed checks.
#include <cstdio>
#include <thread>

#include "tenzing/tenzing.hpp"

using namespace tenzing;

static int g_failed = 0, g_checks = 0;
#define CHECK(cond)                                                                                \
  do {                                                                                             \
    ++g_checks;                                                                                    \
    if (!(cond)) {                                                                                 \
      ++g_failed;                                                                                  \
      std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);                       \
    }                                                                                              \
  } while (0)

static void test_graph() {
  Graph<OpBase> g;
  CHECK(g.vertex_size() == 2);
  auto a = std::make_shared<NoOp>("a");
  auto b = std::make_shared<NoOp>("b");
  g.start_then(a);
  g.then(a, b);
  g.then_finish(b);
  CHECK(g.vertex_size() == 4);
  CHECK(g.succs(a).size() == 1 && g.succs(a)[0]->name() == "b");
  Graph<OpBase> c = g.clone();
  CHECK(c.vertex_size() == 4 && c.find(a) != a && c.contains(a));
  Graph<OpBase> r = g.clone_but_replace(std::make_shared<NoOp>("z"), a);
  CHECK(!r.contains(a) && r.find_name("z") && g.contains(a));
  Graph<OpBase> inner;
  auto i1 = std::make_shared<NoOp>("i1");
  auto i2 = std::make_shared<NoOp>("i2");
  inner.start_then(i1);
  inner.then(i1, i2);
  inner.then_finish(i2);
  auto comp = std::make_shared<GraphCompoundOp>("comp", inner);
  Graph<OpBase> outer;
  outer.start_then(comp);
  outer.then_finish(comp);
  CHECK(outer.clone_but_expand(comp, comp->graph()).vertex_size() == inner.vertex_size());
  bool threw = false;
  try {
    g.then(g.start(), a);
  } catch (const std::exception &) {
    threw = true;
  }
  CHECK(threw);
  CHECK(NoOp("x").eq(NoOp("x")) && !NoOp("x").eq(NoOp("y")));
  Sequence<BoundOp> s;
  CHECK(s.size() == 0 && s.new_unique_event().id() == 0);
}

static void test_noop_state() {
  Graph<OpBase> g;
  auto op1 = std::make_shared<NoOp>("op1");
  g.start_then(op1);
  g.then_finish(op1);
  Platform plat = Platform::make_null(1);
  SDP::State st(g);
  CHECK(st.sequence().size() == 1);
  auto ds = st.get_decisions(plat);
  CHECK(ds.size() == 1 && ds[0]->kind() == Decision::Kind::ExecuteOp);
  for (auto &d : ds) CHECK(st.apply(*d).sequence().size() == 2);
  auto seqs = dfs::get_all_sequences(g, plat);
  CHECK(seqs.size() == 1 && seqs[0].size() == 3);
}

static void test_gpu_graph_on_null_platform() {
  auto k1 = std::make_shared<LambdaGpuOp>("k1");
  auto k2 = std::make_shared<LambdaGpuOp>("k2");
  auto k3 = std::make_shared<LambdaGpuOp>("k3");
  Graph<OpBase> g;
  g.start_then(k1);
  g.then(k1, k2);
  g.then(k1, k3);
  g.then_finish(k2);
  g.then_finish(k3);
  Platform plat = Platform::make_null(2);
  SDP::State st(g);
  auto ds = st.get_decisions(plat);
  CHECK(ds.size() == 2);
  SDP::State s0 = st.apply(*ds[0]), s1 = st.apply(*ds[1]);
  CHECK(bool(SDP::get_equivalence(s0, s1)));
  CHECK(bool(get_equivalence(s0.graph(), s1.graph())));
  CHECK(s0.canonical_key() == s1.canonical_key());
  CHECK(s0.graph().contains(std::make_shared<BoundGpuOp>(k1, Stream(0))));
  auto seqs = dfs::get_all_sequences(g, plat);
  CHECK(seqs.size() > 2);
  for (auto &s : seqs) {
    CHECK(s.back()->name() == "Finish");
    Sequence<BoundOp> copy = s;
    Schedule::remove_redundant_syncs(copy);
    CHECK(copy.size() <= s.size());
    // round trip through the broadcast format
    Json j = to_json(s, g);
    Sequence<BoundOp> back;
    from_json(Json::parse(j.dump()), g, back);
    CHECK(canonical_key(back) == canonical_key(s));
  }
}

static void test_mcts_and_loopback() {
  // fast/slow toy (reference: tenzing-mcts/src_mcts_test/mcts.cpp): optimum is known
  Graph<OpBase> g;
  for (const char *n : {"f0", "s0", "f1", "s1"}) {
    auto o = std::make_shared<NoOp>(n);
    g.start_then(o);
    g.then_finish(o);
  }
  auto cost = [](const Sequence<BoundOp> &seq) {
    auto pos = [&](const std::string &n) {
      for (size_t i = 0; i < seq.size(); ++i)
        if (seq[i]->name() == n) return i;
      return size_t(0);
    };
    return (pos("f0") < pos("s0") ? 1.0 : 2.0) + (pos("f1") < pos("s1") ? 1.0 : 2.0);
  };
  for (const auto &name : mcts::strategy_names()) {
    Platform plat = Platform::make_null(1);
    ModelBenchmarker bench(cost);
    auto strat = mcts::make_strategy(name);
    mcts::Opts opts;
    opts.nIters = 60;
    mcts::Result r = mcts::explore(g, plat, bench, *strat, opts);
    CHECK(!r.simResults.empty() && r.simResults.size() <= 60); // stops early once the 24-leaf tree is fully visited
    CHECK(r.simResults[size_t(r.best())].benchResult.pct10 == 2.0);
  }
  // world of two ranks (threads): rank 0 searches, both benchmark in lock step
  std::vector<size_t> counts(2, 0);
  LoopbackWorld::run(2, [&](int rank, std::shared_ptr<Communicator> comm) {
    Platform plat = Platform::make_null(1, comm);
    ModelBenchmarker bench(cost);
    dfs::Result r = dfs::explore(g, plat, bench);
    counts[size_t(rank)] = r.simResults.size();
  });
  CHECK(counts[0] == 24 && counts[1] == 24);
}

static void test_numeric() {
  CHECK(prime_factors(12) == (std::vector<int>{3, 2, 2}));
  CHECK(round_up(13, 8) == 16);
  CHECK(med(std::vector<double>{1, 3, 2, 10}) == 2.5);
  std::vector<double> drift;
  for (int i = 0; i < 100; ++i) drift.push_back(i);
  CHECK(randomness::runs_test(drift));
}

int main() {
  setenv("TENZING_ACK_NOTICE", "1", 0);
  test_graph();
  test_noop_state();
  test_gpu_graph_on_null_platform();
  test_mcts_and_loopback();
  test_numeric();
  std::printf("%d checks, %d failed\n", g_checks, g_failed);
  return g_failed;
}
