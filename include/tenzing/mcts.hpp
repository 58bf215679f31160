// Monte-Carlo tree search over the decision process, with pluggable exploit scores.
//
// Reference: tenzing-mcts/include/tenzing/mcts/{mcts,mcts_node,mcts_strategy*}.hpp,
// tenzing-mcts/src/mcts.cpp.  What is kept: UCT with c = sqrt(2), -inf explore for
// fully-visited children, random tie-break, "leaf = has an unplayed child",
// optional persistence of rollout nodes, pct10 as the value signal, the nine
// strategies and their formulae, the tree dump schedule and the CSV rows.
// What differs:
//  * nodes own their children through unique_ptr and carry a real parent link
//    (the reference never assigns parent_: SURVEY D1/D2);
//  * strategies are objects behind one interface `select(child)` /
//    `backprop(node, result)` (six of the reference's nine use a stale 3-argument
//    signature and do not compile: D3), chosen at run time by name; the
//    template spelling `explore<FastMin>(...)` is kept as a thin wrapper;
//  * every node keeps the same small statistics block (sorted times, min, max)
//    that all strategies read;
//  * randomness comes from a seeded engine in the search context, not rand().
#pragma once

#include <cstdint>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <random>
#include <string>
#include <vector>

#include "tenzing/benchmarker.hpp"
#include "tenzing/decision.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/platform.hpp"
#include "tenzing/sequence.hpp"
#include "tenzing/state.hpp"

namespace tenzing {
namespace mcts {

struct Opts {
  size_t nIters = 0;          // searches to do; 0 = until the tree is fully visited
  bool dumpTree = false;      // write tree .dot files at iterations 0-9, 10..40 step 10, 50, 75
  std::string dumpTreePrefix = "mcts_";
  bool expandRollout = true;  // keep the nodes a rollout passes through
  uint64_t seed = 0;
  bool removeRedundantSyncs = true;
  size_t maxTreeNodes = 0;    // stop (Stop::large_tree) once the tree holds this many nodes; 0 = no limit
  Benchmark::Opts benchOpts;
};

struct SimResult {
  Sequence<BoundOp> path;
  Benchmark::Result benchResult;
};

struct Result {
  std::vector<SimResult> simResults;
  // header-less rows idx|pct01|pct10|pct50|pct90|pct99|stddev|<op json>|...
  void dump_csv(std::ostream &os) const;
  void dump_csv() const { dump_csv(std::cout); }
  std::string csv() const;
  int best() const; // fastest by pct10, -1 if empty
};

struct Stop {
  enum class Reason { full_tree, large_tree, iterations };
  Reason reason_ = Reason::full_tree;
  bool value_ = false;
  Stop() = default;
  Stop(bool b, Reason r) : reason_(r), value_(b) {}
  explicit operator bool() const { return value_; }
  const char *c_str() const;
  void bcast(int root, Communicator &comm);
};

// statistics every strategy may use
struct NodeStats {
  std::vector<double> times; // sorted ascending, one pct10 per rollout through the node
  double tMin = std::numeric_limits<double>::infinity();
  double tMax = -std::numeric_limits<double>::infinity();
  void add(double t);
};

class Strategy;

class Node {
public:
  struct RolloutResult {
    Sequence<BoundOp> sequence;
    Node *backpropStart = nullptr;
  };

  Node *parent_ = nullptr;
  std::vector<std::unique_ptr<Node>> children_;
  BoundOpPtr op_; // null for a "non-op decision" (graph rewrite) node
  bool expanded_ = false;
  bool fullyVisited_ = false;
  size_t n_ = 0; // rollouts through this node
  // For op nodes the graph of the parent; for graph-rewrite nodes (stream assignment,
  // expansion, choice) the revised graph, built from `rewrite_` the first time it is needed:
  // a node may have a hundred rewrite children of which a search visits a few, and every
  // revised graph is a copy of the whole program graph.
  mutable std::shared_ptr<const Graph<OpBase>> graph_;
  DecisionPtr rewrite_;
  const std::shared_ptr<const Graph<OpBase>> &graph() const;
  NodeStats state_;

  Node(std::shared_ptr<const Graph<OpBase>> graph, BoundOpPtr op, Node *parent = nullptr)
      : parent_(parent), op_(std::move(op)), graph_(std::move(graph)) {}
  // graph-rewrite child: the graph is derived from the parent's on demand
  Node(DecisionPtr rewrite, Node *parent) : parent_(parent), rewrite_(std::move(rewrite)) {}
  // root: sequence = [Start]
  explicit Node(const Graph<OpBase> &graph);
  Node(const Node &) = delete;
  Node &operator=(const Node &) = delete;

  size_t size() const;
  size_t unvisited_size() const;
  size_t fully_visited_size() const;

  bool is_terminal() const;
  bool is_leaf() const;

  // UCT descent to a leaf or terminal node
  Node &select(Strategy &strategy, std::mt19937_64 &rng);
  // create children (once) and return the first unplayed one (this if terminal)
  Node &expand(const Platform &plat);
  // random playout to a terminal state
  RolloutResult get_rollout(const Platform &plat, std::mt19937_64 &rng, bool expand = true);
  void backprop(Strategy &strategy, const Benchmark::Result &br);

  const Node &root() const;
  Node &root();
  // ops on the path root..this
  Sequence<BoundOp> get_sequence() const;

  std::string graphviz_label(const Strategy *strategy = nullptr) const;
  std::string graphviz_name() const;
  std::string desc() const;

  // `at`: the state of this node (its graph, the ops root..this) if the caller has it; a rollout
  // walks down with one state and derives each level from the one above instead of rebuilding
  // it from the root at every level
  void ensure_children(const Platform &plat, const SDP::State *at = nullptr);
};

// exploit score in [0,1] (or +inf to force a child) plus per-rollout bookkeeping
class Strategy {
public:
  virtual ~Strategy() = default;
  virtual std::string name() const = 0;
  virtual double select(const Node &child) = 0;
  virtual void backprop(Node &node, const Benchmark::Result &br) = 0;
  virtual std::string graphviz_label_line(const Node &node) const;
  void seed(uint64_t s) { rng_.seed(s); }

protected:
  std::mt19937_64 rng_{0};
};

#define TENZING_MCTS_STRATEGY(NAME)                                                                \
  class NAME : public Strategy {                                                                   \
  public:                                                                                          \
    using Context = NAME;                                                                          \
    using State = NodeStats;                                                                       \
    std::string name() const override { return #NAME; }                                            \
    double select(const Node &child) override;                                                     \
    void backprop(Node &node, const Benchmark::Result &br) override;

TENZING_MCTS_STRATEGY(FastMin)
};
TENZING_MCTS_STRATEGY(Coverage)
  double minT = std::numeric_limits<double>::infinity();
  double maxT = -std::numeric_limits<double>::infinity();
};
TENZING_MCTS_STRATEGY(Random)
  std::map<const Node *, size_t> selected;
};
TENZING_MCTS_STRATEGY(AvgTime)
};
TENZING_MCTS_STRATEGY(AntiCorrelation)
};
TENZING_MCTS_STRATEGY(NormalizedAntiCorrelation)
  const Node *root = nullptr;
};
TENZING_MCTS_STRATEGY(NormRootCorr)
  const Node *root = nullptr;
};
TENZING_MCTS_STRATEGY(BalanceHistogram)
  const Node *root = nullptr;
};
TENZING_MCTS_STRATEGY(Unvisited)
};
#undef TENZING_MCTS_STRATEGY

// "FastMin", "Coverage", "Random", "AvgTime", "AntiCorrelation",
// "NormalizedAntiCorrelation", "NormRootCorr", "BalanceHistogram", "Unvisited"
std::unique_ptr<Strategy> make_strategy(const std::string &name);
std::vector<std::string> strategy_names();

// 10-bin histogram over [tMin, tMax] shared by the correlation strategies
std::vector<uint64_t> histogram(const std::vector<double> &v, double tMin, double tMax,
                                int nBins = 10);

void dump_graphviz(const std::string &path, const Node &root, const Strategy *strategy = nullptr);
std::string graphviz(const Node &root, const Strategy *strategy = nullptr);

// the search; rank 0 owns the tree, all ranks benchmark in lock step
Result explore(const Graph<OpBase> &g, Platform &plat, Benchmarker &benchmarker,
               Strategy &strategy, const Opts &opts = Opts());

template <typename StrategyT, typename BenchmarkerT>
Result explore(const Graph<OpBase> &g, Platform &plat, BenchmarkerT &benchmarker,
               const Opts &opts = Opts()) {
  StrategyT strategy;
  return explore(g, plat, static_cast<Benchmarker &>(benchmarker), strategy, opts);
}

// legacy frontier (reference: tenzing-mcts/src/mcts_node.cpp:37-149): ready ops
// in all platform variations, replaced by their missing syncs where unsynced
std::vector<BoundOpPtr> get_frontier(const Platform &plat, const Graph<OpBase> &g,
                                     const std::vector<BoundOpPtr> &completed);

} // namespace mcts
} // namespace tenzing
