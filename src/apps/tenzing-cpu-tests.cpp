// Native CPU test runner (reference: test/test_noop_graph.cpp, the in-source doctest cases of
// src/graph.cpp:422-501, src/operation.cpp:87-101, src/sequence.cpp:169-176, and
// test/test_gpu_graph.cu run here on a Null platform; CI runs `tenzing-cpu -tc="*[cpu]*"`).
//
//   tenzing-cpu-tests                 run every case
//   tenzing-cpu-tests -tc='*graph*'   run the cases whose name matches the glob (* and ?)
//   tenzing-cpu-tests -ltc            list the case names
// Exit code = number of failed checks (0 = success).
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <chrono>
#include <functional>
#include <set>
#include <thread>

#include "tenzing/halo/halo_exchange.hpp"
#include "tenzing/spmv/row_part_spmv.hpp"
#include "tenzing/tenzing.hpp"

using namespace tenzing;

static std::atomic<int> g_failed{0}, g_checks{0}; // CHECK is also used inside rank threads
#define CHECK(cond)                                                                                \
  do {                                                                                             \
    ++g_checks;                                                                                    \
    if (!(cond)) {                                                                                 \
      ++g_failed;                                                                                  \
      std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);                       \
    }                                                                                              \
  } while (0)
#define CHECK_THROWS(stmt)                                                                         \
  do {                                                                                             \
    bool tz_threw_ = false;                                                                        \
    try {                                                                                          \
      stmt;                                                                                        \
    } catch (const std::exception &) {                                                             \
      tz_threw_ = true;                                                                            \
    }                                                                                              \
    CHECK(tz_threw_);                                                                              \
  } while (0)

struct Case {
  const char *name;
  std::function<void()> fn;
};
static std::vector<Case> &cases() {
  static std::vector<Case> c;
  return c;
}
struct Registrar {
  Registrar(const char *name, std::function<void()> fn) { cases().push_back({name, std::move(fn)}); }
};
#define TZ_CAT2(a, b) a##b
#define TZ_CAT(a, b) TZ_CAT2(a, b)
#define TEST_CASE(name)                                                                            \
  static void TZ_CAT(tz_case_, __LINE__)();                                                        \
  static Registrar TZ_CAT(tz_reg_, __LINE__)(name, TZ_CAT(tz_case_, __LINE__));                    \
  static void TZ_CAT(tz_case_, __LINE__)()

// glob with * and ?
static bool glob_match(const char *p, const char *s) {
  if (!*p) return !*s;
  if (*p == '*') {
    for (const char *t = s;; ++t) {
      if (glob_match(p + 1, t)) return true;
      if (!*t) return false;
    }
  }
  if (!*s) return false;
  return (*p == '?' || *p == *s) && glob_match(p + 1, s + 1);
}

// ---------------------------------------------------------------------------------------------
// substrate
// ---------------------------------------------------------------------------------------------

TEST_CASE("numeric helpers [cpu]") {
  CHECK(prime_factors(12) == (std::vector<int>{3, 2, 2}));
  CHECK(prime_factors(1).empty());
  CHECK(round_up(13, 8) == 16 && round_up(16, 8) == 16);
  CHECK(med(std::vector<double>{1, 3, 2, 10}) == 2.5);
  CHECK(avg(std::vector<double>{1, 2, 3}) == 2.0);
  CHECK(std::fabs(corr(std::vector<double>{1, 2, 3}, std::vector<double>{2, 4, 6}) - 1.0) < 1e-12);
  CHECK(std::fabs(corr(std::vector<double>{1, 2, 3}, std::vector<double>{3, 2, 1}) + 1.0) < 1e-12);
  std::vector<double> drift;
  for (int i = 0; i < 100; ++i) drift.push_back(i);
  CHECK(randomness::runs_test(drift));                          // monotone series: rejected
  CHECK(randomness::runs_test(std::vector<double>{1, 2, 3})); // too short: rejected
}

TEST_CASE("bijection, equivalence, stream and event handles [cpu]") {
  Bijection<Stream> b;
  CHECK(b.empty());
  CHECK(b.check_or_insert(Stream(0), Stream(1)));
  CHECK(b.check_or_insert(Stream(0), Stream(1)));  // same pair again
  CHECK(!b.check_or_insert(Stream(0), Stream(2))); // 0 is taken
  CHECK(!b.check_or_insert(Stream(3), Stream(1))); // 1 is taken: one-to-one in both directions
  CHECK(b.check_or_insert(Stream(1), Stream(0)));
  CHECK(b.str() == "{0->1,1->0}");
  Equivalence e;
  CHECK(bool(e) && !bool(Equivalence::falsy()));
  CHECK(e.check_or_insert(Event(4), Event(0)) && !e.check_or_insert(Event(4), Event(1)));
  CHECK(Stream(1) < Stream(2) && Stream(2) == Stream(2) && Event(0) != Event(1));
  Json js;
  to_json(js, Stream(3));
  CHECK(js.is_int() && js.as_int() == 3); // bare ints on the wire
  Event back;
  from_json(Json::parse("7"), back);
  CHECK(back == Event(7));
  ScalarValue<double> sv;
  ArrayValue<float> av;
  av.count = 10;
  CHECK(sv.bytes() == 8 && av.bytes() == 40);
}

TEST_CASE("cast helpers [cpu]") {
  OpPtr op = std::make_shared<NoOp>("n");
  CHECK(isa<NoOp>(op) && isa<BoundOp>(op) && !isa<GpuOp>(op));
  CHECK(TENZING_MUST_CAST(NoOp, op)->name() == "n");
  CHECK_THROWS(TENZING_MUST_CAST(GpuOp, op));
}

TEST_CASE("json value: compact, sorted keys, round trip [cpu]") {
  Json j = Json::object();
  j["name"] = "x";
  j["b"] = 1;
  j["a"] = Json::array();
  j["a"].push_back(1.5);
  j["a"].push_back(true);
  j["a"].push_back(Json());
  const std::string text = j.dump();
  CHECK(text == "{\"a\":[1.5,true,null],\"b\":1,\"name\":\"x\"}");
  CHECK(Json::parse(text).dump() == text);
  CHECK(Json::parse(" { \"k\" : \"q\\\"uote\\n\" } ").at("k").as_string() == "q\"uote\n");
  CHECK_THROWS(Json::parse("{\"a\":"));
  CHECK_THROWS(Json::parse("[1,2] trailing"));
}

TEST_CASE("reproduce header round trip [cpu]") {
  const std::string line = reproduce::header_json({"prog", "--flag", "3"});
  reproduce::Header h = reproduce::parse_header(line);
  CHECK(h.valid && h.same_build() && h.args.size() == 3 && h.args[1] == "--flag");
  CHECK(h.version() == reproduce::version_string());
  CHECK(!reproduce::parse_header("0|1|2").valid);
  CHECK(!reproduce::parse_header("{\"dfs__Opts\":{\"maxSeqs\":3}}").valid);
  reproduce::Header other = reproduce::make_header({});
  other.hash = "somethingelse";
  CHECK(!other.same_build());
}

TEST_CASE("numa: topology queries never fail hard [cpu]") {
  // no GPU here: the PCI lookup finds nothing and binding is declined, not fatal
  CHECK(numa::node_of_pci_device("ffff:ff:1f.7") == -1);
  CHECK(numa::cpus_of_node(1 << 20).empty());
  CHECK(!numa::bind_to_node(1 << 20));
  const std::vector<int> node0 = numa::cpus_of_node(0);
  for (int c : node0) CHECK(c >= 0);
}

TEST_CASE("counters [cpu]") {
  counters::ZeroInit<double> z;
  CHECK(double(z) == 0.0);
  z += 1.5;
  z += 1;
  CHECK(double(z) == 2.5);
  const double t0 = counters::now();
  CHECK(counters::now() >= t0);
  counters::reset();
  TENZING_COUNTER_OP(dfs, STATES_VISITED, += 2);
#ifndef TENZING_DISABLE_COUNTERS
  CHECK(uint64_t(counters::dfs().STATES_VISITED) == 2);
#endif
  counters::reset();
  CHECK(uint64_t(counters::dfs().STATES_VISITED) == 0);
}

// ---------------------------------------------------------------------------------------------
// SDP core
// ---------------------------------------------------------------------------------------------

TEST_CASE("graph construction and rewrites [cpu]") {
  Graph<OpBase> g;
  CHECK(g.vertex_size() == 2);
  auto a = std::make_shared<NoOp>("a");
  auto b = std::make_shared<NoOp>("b");
  g.start_then(a);
  g.then(a, b);
  g.then_finish(b);
  CHECK(g.vertex_size() == 4 && g.edge_size() == 3 && g.validate().empty());
  CHECK(g.succs(a).size() == 1 && g.succs(a)[0]->name() == "b");
  CHECK(g.preds_find_name("b").first && g.preds_find_name("b").second[0]->name() == "a");
  CHECK(!g.succs_find_name("nope").first);
  Graph<OpBase> c = g.clone();
  CHECK(c.vertex_size() == 4 && c.find(a) != a && c.contains(a));
  Graph<OpBase> r = g.clone_but_replace(std::make_shared<NoOp>("z"), a);
  CHECK(!r.contains(a) && r.find_name("z") && g.contains(a) && r.validate().empty());
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
  Graph<OpBase> expanded = outer.clone_but_expand(comp, comp->graph());
  CHECK(expanded.vertex_size() == inner.vertex_size() && expanded.validate().empty());
  CHECK_THROWS(g.then(g.start(), a));
  CHECK(NoOp("x").eq(NoOp("x")) && !NoOp("x").eq(NoOp("y")));
  CHECK(Start().eq(Start()) && Finish().eq(Finish()) && !Start().eq(Finish()));
  // a dangling vertex and a cycle are reported
  Graph<OpBase> bad = g;
  bad.then(a, std::make_shared<NoOp>("dangling"));
  CHECK(!bad.validate().empty());
  Graph<OpBase> cyc = g;
  cyc.then(b, a);
  CHECK(cyc.validate() == "the graph has a cycle");
  // erase keeps the adjacency mirrored
  Graph<OpBase> e = g;
  e.then_finish(a);
  e.erase(b);
  CHECK(e.vertex_size() == 3 && e.validate().empty());
  Sequence<BoundOp> s;
  CHECK(s.size() == 0 && s.new_unique_event().id() == 0);
}

TEST_CASE("graphviz text [cpu]") {
  Graph<OpBase> g;
  auto k = std::make_shared<LambdaGpuOp>("kern");
  g.start_then(k);
  g.then_finish(k);
  Graph<OpBase> bound = g.clone_but_replace(std::make_shared<BoundGpuOp>(k, Stream(1)), k);
  const std::string dot = bound.graphviz();
  CHECK(dot.rfind("digraph", 0) == 0);
  CHECK(dot.find("kern") != std::string::npos && dot.find("->") != std::string::npos);
}

TEST_CASE("noop graph: decisions and the single sequence [cpu]") {
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

TEST_CASE("gpu graph on a null platform: stream symmetry, cleaning, broadcast format [cpu]") {
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
  for (auto &d : ds) CHECK(d->kind() == Decision::Kind::AssignOpStream);
  SDP::State s0 = st.apply(*ds[0]), s1 = st.apply(*ds[1]);
  CHECK(bool(SDP::get_equivalence(s0, s1)));
  CHECK(bool(get_equivalence(s0.graph(), s1.graph())));
  CHECK(s0.canonical_key() == s1.canonical_key());
  CHECK(s0.graph().contains(std::make_shared<BoundGpuOp>(k1, Stream(0))));
  CHECK(st.frontier(plat).size() == 1); // the two assignments are the same up to renaming
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

TEST_CASE("event synchronizer: the three cross-op rules [cpu]") {
  auto a = std::make_shared<LambdaGpuOp>("a");
  auto b = std::make_shared<LambdaGpuOp>("b");
  auto c = std::make_shared<NoOp>("c");
  Graph<OpBase> g;
  g.start_then(a);
  g.then(a, b);
  g.then(a, c);
  g.then_finish(b);
  g.then_finish(c);
  auto a0 = std::make_shared<BoundGpuOp>(a, Stream(0));
  auto b0 = std::make_shared<BoundGpuOp>(b, Stream(0));
  auto b1 = std::make_shared<BoundGpuOp>(b, Stream(1));
  Graph<OpBase> gb = g.clone_but_replace(a0, a);
  Sequence<BoundOp> path;
  path.push_back(std::make_shared<Start>());
  path.push_back(a0);
  // same stream: ordered by the stream itself
  CHECK(EventSynchronizer::is_synced(b0, gb.clone_but_replace(b0, b), path));
  // other stream: record on a's stream, then wait on b's
  Graph<OpBase> g1 = gb.clone_but_replace(b1, b);
  CHECK(!EventSynchronizer::is_synced(b1, g1, path));
  auto s1 = EventSynchronizer::make_syncs(b1, g1, path);
  CHECK(s1.size() == 1 && isa<CudaEventRecord>(s1[0]));
  path.push_back(s1[0]);
  auto s2 = EventSynchronizer::make_syncs(b1, g1, path);
  CHECK(s2.size() == 1 && isa<CudaStreamWaitEvent>(s2[0]));
  path.push_back(s2[0]);
  CHECK(EventSynchronizer::is_synced(b1, g1, path));
  // host op after a GPU op: the same record can serve, then the host must block on it
  CHECK(!EventSynchronizer::is_synced(c, g1, path));
  auto s3 = EventSynchronizer::make_syncs(c, g1, path);
  CHECK(s3.size() == 1);
  path.push_back(s3[0]);
  if (!isa<CudaEventSync>(s3[0])) {
    auto s4 = EventSynchronizer::make_syncs(c, g1, path);
    CHECK(s4.size() == 1 && isa<CudaEventSync>(s4[0]));
    path.push_back(s4[0]);
  }
  CHECK(EventSynchronizer::is_synced(c, g1, path));
}

TEST_CASE("schedule cleaning removes an unconsumed record [cpu]") {
  auto k = std::make_shared<LambdaGpuOp>("k");
  Sequence<BoundOp> seq;
  seq.push_back(std::make_shared<Start>());
  seq.push_back(std::make_shared<BoundGpuOp>(k, Stream(0)));
  seq.push_back(std::make_shared<CudaEventRecord>(Event(0), Stream(0)));
  seq.push_back(std::make_shared<Finish>());
  const int removed = Schedule::remove_redundant_syncs(seq);
  CHECK(removed == 1 && seq.size() == 3);
}

TEST_CASE("make_schedules enumerates the linear extensions [cpu]") {
  Graph<OpBase> g;
  for (const char *n : {"p", "q", "r"}) {
    auto o = std::make_shared<NoOp>(n);
    g.start_then(o);
    g.then_finish(o);
  }
  CHECK(count_schedules(g) == 6);
  CHECK(make_schedules(g).size() == 6);
  auto rnd = make_schedules_random(g, 5, 3);
  CHECK(rnd.size() == 5);
  for (auto &s : rnd) CHECK(s.order.size() == 5);
}

// ---------------------------------------------------------------------------------------------
// solvers and benchmarkers
// ---------------------------------------------------------------------------------------------

namespace {
// fast/slow toy (reference: tenzing-mcts/src_mcts_test/mcts.cpp): the optimum is known
Graph<OpBase> fast_slow_graph() {
  Graph<OpBase> g;
  for (const char *n : {"f0", "s0", "f1", "s1"}) {
    auto o = std::make_shared<NoOp>(n);
    g.start_then(o);
    g.then_finish(o);
  }
  return g;
}
double fast_slow_cost(const Sequence<BoundOp> &seq) {
  auto pos = [&](const std::string &n) {
    for (size_t i = 0; i < seq.size(); ++i)
      if (seq[i]->name() == n) return i;
    return size_t(0);
  };
  return (pos("f0") < pos("s0") ? 1.0 : 2.0) + (pos("f1") < pos("s1") ? 1.0 : 2.0);
}
} // namespace

TEST_CASE("mcts: every strategy finds the optimum of the toy problem [cpu]") {
  Graph<OpBase> g = fast_slow_graph();
  for (const auto &name : mcts::strategy_names()) {
    Platform plat = Platform::make_null(1);
    ModelBenchmarker bench(fast_slow_cost);
    auto strat = mcts::make_strategy(name);
    mcts::Opts opts;
    opts.nIters = 60;
    mcts::Result r = mcts::explore(g, plat, bench, *strat, opts);
    // stops early once the 24-leaf tree is fully visited
    CHECK(!r.simResults.empty() && r.simResults.size() <= 60);
    CHECK(r.simResults[size_t(r.best())].benchResult.pct10 == 2.0);
  }
}

TEST_CASE("dfs in a world of two ranks (threads) runs in lock step [cpu]") {
  Graph<OpBase> g = fast_slow_graph();
  std::vector<size_t> counts(2, 0);
  LoopbackWorld::run(2, [&](int rank, std::shared_ptr<Communicator> comm) {
    Platform plat = Platform::make_null(1, comm);
    ModelBenchmarker bench(fast_slow_cost);
    dfs::Result r = dfs::explore(g, plat, bench);
    counts[size_t(rank)] = r.simResults.size();
  });
  CHECK(counts[0] == 24 && counts[1] == 24);
}

TEST_CASE("dfs csv replays through CsvBenchmarker [cpu]") {
  Graph<OpBase> g = fast_slow_graph();
  Platform plat = Platform::make_null(1);
  ModelBenchmarker bench(fast_slow_cost);
  dfs::Result r = dfs::explore(g, plat, bench);
  const std::string text = reproduce::header_json({"native-test"}) + "\n" + r.csv();
  CsvBenchmarker replay = CsvBenchmarker::from_text(text, g);
  CHECK(replay.rows().size() == r.simResults.size());
  CHECK(replay.producer().valid && replay.producer().args[0] == "native-test");
  for (auto &sr : r.simResults) {
    Sequence<BoundOp> seq = sr.seq;
    CHECK(replay.benchmark(seq, plat).pct10 == sr.benchResult.pct10);
  }
  Sequence<BoundOp> unknown;
  unknown.push_back(std::make_shared<Start>());
  unknown.push_back(std::make_shared<Finish>());
  CHECK_THROWS(replay.benchmark(unknown, plat));
}

TEST_CASE("timeline simulator: overlap needs two streams [cpu]") {
  auto k1 = std::make_shared<LambdaGpuOp>("k1");
  auto k2 = std::make_shared<LambdaGpuOp>("k2");
  SimBenchmarker::Model m;
  m.gpuSeconds["k1"] = 100e-6;
  m.gpuSeconds["k2"] = 100e-6;
  m.launchCost = 1e-6;
  SimBenchmarker sim(m);
  auto seq_on = [&](int s2) {
    Sequence<BoundOp> s;
    s.push_back(std::make_shared<Start>());
    s.push_back(std::make_shared<BoundGpuOp>(k1, Stream(0)));
    s.push_back(std::make_shared<BoundGpuOp>(k2, Stream(s2)));
    s.push_back(std::make_shared<Finish>());
    return s;
  };
  // spans: (start, end) of every op on its own timeline; [2] is k2
  const double serial = sim.timeline(seq_on(0))[2].end, overlapped = sim.timeline(seq_on(1))[2].end;
  CHECK(serial > 199e-6 && overlapped < 110e-6);
  // the host does not wait for a GPU op unless the sequence says so
  CHECK(sim.simulate(seq_on(1)) < 10e-6);
}

// ---------------------------------------------------------------------------------------------
// workloads (host memory, Null platform: structure and set-up only)
// ---------------------------------------------------------------------------------------------

TEST_CASE("partition, coo/csr, local/remote split [cpu]") {
  CHECK(get_partition(10, 0, 3).lb == 0 && get_partition(10, 0, 3).ub == 4);
  CHECK(get_partition(10, 2, 3).lb == 7 && get_partition(10, 2, 3).ub == 10);
  for (int64_t i = 0; i < 10; ++i) {
    const int o = get_owner(10, i, 3);
    CHECK(get_partition(10, o, 3).lb <= i && i < get_partition(10, o, 3).ub);
  }
  CooMat<int, double> coo(3, 6);
  coo.push_back(2, 5, 1.0);
  coo.push_back(0, 1, 2.0);
  coo.push_back(0, 1, 9.0); // duplicate coordinate
  coo.push_back(1, 3, 3.0);
  coo.push_back(1, 0, 4.0);
  coo.sort_and_dedup();
  CHECK(coo.nnz() == 4);
  CsrMat<int, double> m(coo);
  CHECK(m.num_rows() == 3 && m.nnz() == 4);
  CHECK(m.retain_rows(1, 3).nnz() == 3);
  // columns [2, 4) local
  SplitMat<int, double> s = split_local_remote(m, 2, 4);
  CHECK(s.local.nnz() == 1 && s.remote.nnz() == 3);
  CHECK(s.globals == (std::vector<int>{0, 1, 5}));
  std::vector<double> x{1, 2, 3, 4, 5, 6};
  std::vector<double> xl{x[2], x[3]}, xr{x[0], x[1], x[5]};
  std::vector<double> y = m.multiply(x), yl = s.local.multiply(xl), yr = s.remote.multiply(xr);
  for (size_t i = 0; i < y.size(); ++i) CHECK(y[i] == yl[i] + yr[i]);
  auto parts = part_by_rows(m, 2);
  CHECK(parts.size() == 2 && parts[0].num_rows() + parts[1].num_rows() == 3);
  // the generator gives the same rows whichever block they are generated in
  auto whole = random_band_rows<int, double>(100, 10, 5, 0, 100, 42);
  auto block = random_band_rows<int, double>(100, 10, 5, 30, 60, 42);
  CHECK(whole.retain_rows(30, 60).colInd_ == block.colInd_);
}

TEST_CASE("spmv set-up on two ranks and the shape of its graphs [cpu]") {
  std::vector<int> vertices(2, 0);
  std::vector<size_t> seqCount(2, 0);
  LoopbackWorld::run(2, [&](int rank, std::shared_ptr<Communicator> comm) {
    auto spmv = make_band_spmv<double>(comm, 200, 100, 6, 1, SymmetricHeap::Memory::Host);
    CHECK(spmv->size() == 2 && spmv->row_range().extent() == 100);
    CHECK(spmv->lA().rows == 100 && spmv->num_remote_cols() > 0);
    CHECK(spmv->peers().size() == 1 && spmv->peers()[0] == 1 - rank);
    SpMV<double> ref("spmv", spmv, RowPartSpmv<double>::Style::Reference);
    vertices[size_t(rank)] = int(ref.graph().vertex_size());
    CHECK(ref.graph().validate().empty());
    Graph<OpBase> g;
    auto fused = std::make_shared<SpMV<double>>("spmv", spmv, RowPartSpmv<double>::Style::Fused);
    g.start_then(fused);
    g.then_finish(fused);
    Platform plat = Platform::make_null(2, comm);
    auto seqs = dfs::get_all_sequences(g, plat);
    seqCount[size_t(rank)] = seqs.size();
    for (auto &s : seqs) CHECK(FlagSynchronizer::is_deadlock_free(s));
  });
  CHECK(vertices[0] == 10 && vertices[1] == 10); // the reference's 8 ops + Start + Finish
  CHECK(seqCount[0] > 0 && seqCount[0] == seqCount[1]);
}

TEST_CASE("halo geometry and graphs of every style [cpu]") {
  using namespace HaloExchange;
  CHECK(directions(6).size() == 6 && directions(18).size() == 18 && directions(26).size() == 26);
  CHECK(dir_to_tag(Dim3<int>(-1, -1, -1)) == 0 && dir_to_tag(Dim3<int>(1, 1, 1)) == 26);
  CHECK(rank_grid(8) == Dim3<int>(2, 2, 2) && rank_grid(12).flatten() == 12);
  CHECK(-Dim3<int>(1, -2, 3) == Dim3<int>(-1, 2, -3)); // the reference returns the copy un-negated
  Args args;
  args.nX = 8;
  args.nY = 6;
  args.nZ = 4;
  args.nQ = 2;
  args.nGhost = 2;
  args.set_default_topology(1);
  auto dom = std::make_shared<Domain>(std::make_shared<SelfComm>(), args, SymmetricHeap::Memory::Host);
  CHECK(dom->extent() == Dim3<int>(12, 10, 8));
  CHECK(dom->region_extent(Dim3<int>(1, 0, 0)) == Dim3<int>(2, 6, 4));
  CHECK(dom->region_extent(Dim3<int>(1, 1, 1)) == Dim3<int>(2, 2, 2));
  CHECK(dom->neighbor(Dim3<int>(1, 0, 0)) == 0); // periodic world of one
  size_t bytes = 0;
  for (const auto &d : dom->dirs()) bytes += dom->region_elems(d) * 8;
  CHECK(bytes == dom->bytes_per_exchange());
  CHECK((dom->index(dom->args().nGhost, 0, 0, 0) * 8) % 16 == 0); // first interior cell aligned
  Platform plat = Platform::make_null(2);
  for (Style st : {Style::Reference, Style::Staged, Style::Direct, Style::FusedAll, Style::FusedStaged,
                   Style::FusedGroups, Style::FusedMixed, Style::FusedPipelined, Style::FusedPull}) {
    Graph<OpBase> g;
    add_to_graph(g, dom, {g.start()}, {g.finish()}, st);
    CHECK(g.validate().empty());
    CHECK(g.vertex_size() > 2);
  }
  // the small styles can be enumerated completely; every schedule publishes before it waits
  Graph<OpBase> g;
  add_to_graph(g, dom, {g.start()}, {g.finish()}, Style::FusedMixed);
  auto seqs = dfs::get_all_sequences(g, plat);
  CHECK(!seqs.empty());
  for (auto &s : seqs) CHECK(FlagSynchronizer::is_deadlock_free(s));
}

// ---------------------------------------------------------------------------------------------
// host execution: the workloads really run, ranks are threads (this is what TSan checks)
// ---------------------------------------------------------------------------------------------

namespace {

// a complete sequence: always the k-th decision offered (k varies the schedule)
Sequence<BoundOp> walk(const Graph<OpBase> &g, const Platform &plat, size_t k) {
  SDP::State st(g);
  for (;;) {
    auto ds = st.get_decisions(plat);
    if (ds.empty()) break;
    st = st.apply(*ds[k % ds.size()]);
  }
  Sequence<BoundOp> seq = st.sequence();
  Schedule::remove_redundant_syncs(seq);
  return seq;
}

double cell_value(const Dim3<int> &global, const Dim3<int> &G, int q, int nQ) {
  auto wrap = [](int v, int n) { return ((v % n) + n) % n; };
  const int64_t cell = (int64_t(wrap(global.z, G.z)) * G.y + wrap(global.y, G.y)) * G.x + wrap(global.x, G.x);
  return double(cell * nQ + q);
}

// interior <- global id, ghosts <- -1
void halo_fill(const HaloExchange::Domain &dom) {
  const auto &a = dom.args();
  const Dim3<int> c = a.rankToCoord(dom.rank());
  const Dim3<int> G(a.rankGrid.x * a.nX, a.rankGrid.y * a.nY, a.rankGrid.z * a.nZ);
  double *grid = static_cast<double *>(dom.grid());
  const Dim3<int> e = dom.extent();
  const int g = a.nGhost;
  for (int q = 0; q < a.nQ; ++q)
    for (int z = 0; z < e.z; ++z)
      for (int y = 0; y < e.y; ++y)
        for (int x = 0; x < e.x; ++x) {
          const bool interior = x >= g && x < g + a.nX && y >= g && y < g + a.nY && z >= g && z < g + a.nZ;
          const Dim3<int> global(c.x * a.nX + x - g, c.y * a.nY + y - g, c.z * a.nZ + z - g);
          grid[dom.index(x, y, z, q)] = interior ? cell_value(global, G, q, a.nQ) : -1.0;
        }
}

// every cell (ghosts included) holds the periodic global field
int halo_mismatches(const HaloExchange::Domain &dom) {
  const auto &a = dom.args();
  const Dim3<int> c = a.rankToCoord(dom.rank());
  const Dim3<int> G(a.rankGrid.x * a.nX, a.rankGrid.y * a.nY, a.rankGrid.z * a.nZ);
  const double *grid = static_cast<const double *>(dom.grid());
  const Dim3<int> e = dom.extent();
  const int g = a.nGhost;
  int bad = 0;
  for (int q = 0; q < a.nQ; ++q)
    for (int z = 0; z < e.z; ++z)
      for (int y = 0; y < e.y; ++y)
        for (int x = 0; x < e.x; ++x) {
          const Dim3<int> global(c.x * a.nX + x - g, c.y * a.nY + y - g, c.z * a.nZ + z - g);
          if (grid[dom.index(x, y, z, q)] != cell_value(global, G, q, a.nQ)) ++bad;
        }
  return bad;
}

} // namespace

TEST_CASE("halo exchange executes on the host: 8 thread ranks, every style, three epochs [cpu]") {
  using namespace HaloExchange;
  std::atomic<int> bad{0}, runs{0};
  LoopbackWorld::run(8, [&](int, std::shared_ptr<Communicator> comm) {
    Args args;
    args.nX = 6, args.nY = 5, args.nZ = 4, args.nQ = 2, args.nGhost = 2;
    args.set_default_topology(comm->size());
    auto dom = std::make_shared<Domain>(comm, args, SymmetricHeap::Memory::Host);
    Platform plat = Platform::make_null(2, comm);
    plat.set_host_execution(true);
    for (Style st : {Style::Reference, Style::Staged, Style::Direct, Style::FusedAll, Style::FusedStaged,
                     Style::FusedGroups, Style::FusedMixed, Style::FusedPipelined, Style::FusedPull}) {
      Graph<OpBase> g;
      add_to_graph(g, dom, {g.start()}, {g.finish()}, st);
      for (size_t k = 0; k < 2; ++k) {
        Sequence<BoundOp> seq = walk(g, plat, k);
        halo_fill(*dom);
        comm->barrier();
        for (int epoch = 0; epoch < 3; ++epoch) run_sequence(seq, plat);
        comm->barrier();
        bad += halo_mismatches(*dom);
        ++runs;
        comm->barrier();
      }
    }
  });
  CHECK(bad.load() == 0);
  CHECK(runs.load() == 8 * 9 * 2);
}

TEST_CASE("distributed spmv executes on the host: 4 thread ranks, every style, inline and with stream threads [cpu]") {
  std::atomic<int> bad{0};
  LoopbackWorld::run(4, [&](int rank, std::shared_ptr<Communicator> comm) {
    using R = RowPartSpmv<double>;
    const int64_t n = 203;
    for (int64_t bw : {int64_t(60), n}) { // band and all-to-all
      auto spmv = make_band_spmv<double>(comm, n, bw, 6, 9, SymmetricHeap::Memory::Host);
      Platform plat = Platform::make_null(2, comm);
      if (bw == n) plat.enable_host_streams(); // the all-to-all case also with stream threads
      else plat.set_host_execution(true);
      // x = global column id + 1; the oracle multiplies this rank's rows on the host
      const Range rr = spmv->row_range();
      for (int64_t i = 0; i < rr.extent(); ++i) spmv->lx()[i] = double(rr.lb + i + 1);
      std::vector<double> x(static_cast<size_t>(n));
      for (int64_t i = 0; i < n; ++i) x[size_t(i)] = double(i + 1);
      const CsrMat<int, double> mine = random_band_rows<int, double>(n, bw, 6, rr.lb, rr.ub, 9);
      const std::vector<double> want = mine.multiply(x);
      for (R::Style st : {R::Style::Reference, R::Style::Fused, R::Style::FusedSeparate, R::Style::FusedAll,
                          R::Style::FusedPull}) {
        Graph<OpBase> g;
        auto op = std::make_shared<SpMV<double>>("spmv", spmv, st);
        g.start_then(op);
        g.then_finish(op);
        for (size_t k = 0; k < 2; ++k) {
          Sequence<BoundOp> seq = walk(g, plat, k);
          comm->barrier();
          for (int epoch = 0; epoch < 3; ++epoch) run_sequence(seq, plat);
          plat.synchronize();
          comm->barrier();
          for (int64_t i = 0; i < rr.extent(); ++i)
            if (std::fabs(spmv->ly()[i] - want[size_t(i)]) > 1e-9 * std::fabs(want[size_t(i)]) + 1e-12) ++bad;
          comm->barrier();
        }
      }
      (void)rank;
    }
  });
  CHECK(bad.load() == 0);
}

TEST_CASE("streams as threads: halo schedules run concurrently and correctly [cpu]") {
  using namespace HaloExchange;
  std::atomic<int> bad{0};
  std::atomic<uint64_t> tasks{0};
  LoopbackWorld::run(4, [&](int, std::shared_ptr<Communicator> comm) {
    Args args;
    args.nX = 6, args.nY = 5, args.nZ = 4, args.nQ = 2, args.nGhost = 2;
    args.set_default_topology(comm->size());
    auto dom = std::make_shared<Domain>(comm, args, SymmetricHeap::Memory::Host);
    Platform plat = Platform::make_null(2, comm);
    plat.enable_host_streams();
    for (Style st : {Style::Reference, Style::Staged, Style::Direct, Style::FusedAll, Style::FusedStaged,
                     Style::FusedGroups, Style::FusedMixed, Style::FusedPipelined, Style::FusedPull}) {
      Graph<OpBase> g;
      add_to_graph(g, dom, {g.start()}, {g.finish()}, st);
      for (size_t k = 0; k < 3; ++k) {
        Sequence<BoundOp> seq = walk(g, plat, k); // k = 1, 2 put ops on the second stream
        halo_fill(*dom);
        comm->barrier();
        for (int epoch = 0; epoch < 3; ++epoch) run_sequence(seq, plat);
        plat.synchronize();
        comm->barrier();
        bad += halo_mismatches(*dom);
        comm->barrier();
      }
    }
    tasks += plat.host_streams()->tasks_done();
  });
  CHECK(bad.load() == 0);
  CHECK(tasks.load() > 0);
}

TEST_CASE("host streams: event semantics [cpu]") {
  HostStreams hs(2);
  std::atomic<int> order{0};
  int a = -1, b = -1, c = -1;
  hs.enqueue(0, [&] {
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
    a = order++;
  });
  hs.record(7, 0);
  hs.stream_wait(1, 7); // stream 1 continues only after stream 0's first task
  hs.enqueue(1, [&] { b = order++; });
  hs.stream_wait(1, 99); // never recorded: a no-op, as in CUDA
  hs.enqueue(1, [&] { c = order++; });
  hs.event_sync(7);
  CHECK(a == 0);
  hs.sync_all();
  CHECK(a == 0 && b == 1 && c == 2);
  // an exception in a task reaches the host at the next synchronisation, nothing hangs
  hs.enqueue(0, [] { throw std::runtime_error("kernel failed"); });
  hs.enqueue(0, [&] { a = 100; }); // skipped: the stream has failed
  hs.record(8, 0);
  bool threw = false;
  try {
    hs.event_sync(8);
  } catch (const std::exception &e) {
    threw = std::string(e.what()) == "kernel failed";
  }
  CHECK(threw && a == 0);
  hs.enqueue(0, [&] { a = 5; }); // usable again
  hs.sync_all();
  CHECK(a == 5);
}

// Not tagged [cpu]: scripts/sanitize.sh runs it under ThreadSanitizer and EXPECTS a report.
// Two ops on different streams write the same buffer and the graph has no edge between them
// (the reference's SpMV has exactly this defect between yl and yr): every
// schedule the solver builds is "correct" by its own rules, and only execution can tell.
TEST_CASE("racy graph: two unordered ops write one buffer [tsan-must-report]") {
  Platform plat = Platform::make_null(2);
  plat.enable_host_streams();
  std::vector<double> buf(4096, 0.0);
  struct Writer : GpuOp {
    std::string n;
    std::vector<double> *b;
    double v;
    Writer(std::string name, std::vector<double> *buf, double val) : n(std::move(name)), b(buf), v(val) {}
    std::string name() const override { return n; }
    CLONE_DEF(Writer)
    void run(cudaStream_t) override {
      for (auto &x : *b) x = v;
    }
  };
  auto w0 = std::make_shared<Writer>("w0", &buf, 1.0);
  auto w1 = std::make_shared<Writer>("w1", &buf, 2.0);
  Graph<OpBase> g;
  g.start_then(w0);
  g.start_then(w1); // no edge w0 -> w1
  g.then_finish(w0);
  g.then_finish(w1);
  Sequence<BoundOp> seq = walk(g, plat, 1); // binds the two ops to different streams
  std::set<int> streams;
  for (const auto &op : seq)
    if (auto b = std::dynamic_pointer_cast<BoundGpuOp>(op)) streams.insert(b->stream().id());
  CHECK(streams.size() == 2);
  for (int i = 0; i < 20; ++i) run_sequence(seq, plat);
  plat.synchronize();
  CHECK(buf[0] == 1.0 || buf[0] == 2.0);
}

TEST_CASE("a flag nobody publishes makes the wait throw, and the world is released [cpu]") {
  host::set_wait_timeout_seconds(0.3);
  bool threw = false;
  try {
    LoopbackWorld::run(2, [&](int rank, std::shared_ptr<Communicator> comm) {
      uint64_t flag = 0;
      if (rank == 0) host::flags_wait(&flag, 1, 1); // never published
      comm->barrier();
    });
  } catch (const std::exception &e) {
    threw = std::string(e.what()).find("did not reach epoch") != std::string::npos;
  }
  host::set_wait_timeout_seconds(20.0);
  CHECK(threw);
}

// ---------------------------------------------------------------------------------------------

int main(int argc, char **argv) {
  setenv("TENZING_ACK_NOTICE", "1", 0);
  const char *filter = "*";
  bool list = false;
  for (int i = 1; i < argc; ++i) {
    if (std::strncmp(argv[i], "-tc=", 4) == 0) filter = argv[i] + 4;
    else if (std::strncmp(argv[i], "--test-case=", 12) == 0) filter = argv[i] + 12;
    else if (std::strcmp(argv[i], "-ltc") == 0 || std::strcmp(argv[i], "--list-test-cases") == 0) list = true;
    else {
      std::fprintf(stderr, "usage: %s [-tc=<glob>] [-ltc]\n", argv[0]);
      return 2;
    }
  }
  int ran = 0;
  for (const Case &c : cases()) {
    if (!glob_match(filter, c.name)) continue;
    if (list) {
      std::printf("%s\n", c.name);
      continue;
    }
    const int before = g_failed.load();
    try {
      c.fn();
    } catch (const std::exception &e) {
      ++g_failed;
      std::fprintf(stderr, "FAILED %s: uncaught exception: %s\n", c.name, e.what());
    }
    ++ran;
    std::printf("[%s] %s\n", g_failed.load() == before ? " ok " : "FAIL", c.name);
  }
  if (list) return 0;
  std::printf("%d cases, %d checks, %d failed\n", ran, g_checks.load(), g_failed.load());
  return g_failed.load();
}
