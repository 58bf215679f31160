#include "tenzing/mcts.hpp"

#include <algorithm>
#include <cmath>
#include <fstream>
#include <functional>
#include <sstream>

#include "tenzing/communicator.hpp"
#include "tenzing/counters.hpp"
#include "tenzing/numeric.hpp"
#include "tenzing/schedule.hpp"
#include "tenzing/trap.hpp"

namespace tenzing {
namespace mcts {

// ---------------------------------------------------------------- results
void Result::dump_csv(std::ostream &os) const {
  for (size_t i = 0; i < simResults.size(); ++i)
    os << csv_row(i, simResults[i].benchResult, simResults[i].path) << "\n";
  os.flush();
}
std::string Result::csv() const {
  std::ostringstream ss;
  dump_csv(ss);
  return ss.str();
}
int Result::best() const {
  int b = -1;
  for (size_t i = 0; i < simResults.size(); ++i)
    if (b < 0 || simResults[i].benchResult.pct10 < simResults[b].benchResult.pct10) b = int(i);
  return b;
}

const char *Stop::c_str() const {
  switch (reason_) {
  case Reason::full_tree: return "full_tree";
  case Reason::large_tree: return "large_tree";
  case Reason::iterations: return "iterations";
  }
  return "?";
}
void Stop::bcast(int root, Communicator &comm) {
  int64_t v = (value_ ? 1 : 0) | (int64_t(reason_) << 1);
  v = comm.bcast_int(v, root);
  value_ = v & 1;
  reason_ = Reason(v >> 1);
}

void NodeStats::add(double t) {
  times.insert(std::upper_bound(times.begin(), times.end(), t), t);
  tMin = std::min(tMin, t);
  tMax = std::max(tMax, t);
}

// ---------------------------------------------------------------- node
Node::Node(const Graph<OpBase> &graph)
    : graph_(std::make_shared<const Graph<OpBase>>(graph)) {
  op_ = std::dynamic_pointer_cast<BoundOp>(graph_->start());
}

size_t Node::size() const {
  size_t n = 1;
  for (const auto &c : children_) n += c->size();
  return n;
}
size_t Node::unvisited_size() const {
  size_t n = (n_ == 0);
  for (const auto &c : children_) n += c->unvisited_size();
  return n;
}
size_t Node::fully_visited_size() const {
  size_t n = fullyVisited_;
  for (const auto &c : children_) n += c->fully_visited_size();
  return n;
}

const std::shared_ptr<const Graph<OpBase>> &Node::graph() const {
  if (!graph_) {
    if (!rewrite_ || !parent_) THROW_RUNTIME("mcts node without a graph");
    const SDP::State before(parent_->graph(), parent_->get_sequence());
    graph_ = before.apply(*rewrite_).graph_ptr();
  }
  return graph_;
}

bool Node::is_terminal() const { return op_ && op_->uid() == graph()->finish()->uid(); }

bool Node::is_leaf() const {
  if (children_.empty()) return true;
  for (const auto &c : children_)
    if (c->n_ == 0) return true;
  return false;
}

const Node &Node::root() const { return parent_ ? parent_->root() : *this; }
Node &Node::root() { return parent_ ? parent_->root() : *this; }

Sequence<BoundOp> Node::get_sequence() const {
  std::vector<BoundOpPtr> rev;
  for (const Node *cur = this; cur; cur = cur->parent_)
    if (cur->op_) rev.push_back(cur->op_);
  Sequence<BoundOp> seq;
  for (auto it = rev.rbegin(); it != rev.rend(); ++it) seq.push_back(*it);
  return seq;
}

std::string Node::graphviz_name() const { return op_ ? op_->name() : "non-op decision"; }
std::string Node::desc() const { return op_ ? op_->desc() : "non-op decision"; }

std::string Node::graphviz_label(const Strategy *strategy) const {
  std::ostringstream ss;
  auto hs = op_ ? std::dynamic_pointer_cast<HasStream>(op_) : nullptr;
  auto he = op_ ? std::dynamic_pointer_cast<HasEvent>(op_) : nullptr;
  if (hs) {
    ss << "s ";
    const auto v = hs->get_streams();
    for (size_t i = 0; i < v.size(); ++i) ss << (i ? "," : "") << v[i].str();
  }
  if (he) {
    if (hs) ss << " ";
    ss << "e ";
    const auto v = he->get_events();
    for (size_t i = 0; i < v.size(); ++i) ss << (i ? "," : "") << v[i].str();
  }
  if (hs || he) ss << "\n";
  if (strategy) ss << strategy->graphviz_label_line(*this);
  std::string s = ss.str();
  while (!s.empty() && s.back() == '\n') s.pop_back();
  return s;
}

void Node::ensure_children(const Platform &plat, const SDP::State *at) {
  if (expanded_) return;
  expanded_ = true;
  if (is_terminal()) return;
  std::unique_ptr<SDP::State> own;
  if (!at) {
    own.reset(new SDP::State(graph(), get_sequence()));
    at = own.get();
  }
  for (const DecisionPtr &d : at->get_decisions(plat)) {
    if (d->kind() == Decision::Kind::ExecuteOp) {
      const auto &e = static_cast<const ExecuteOp &>(*d);
      children_.emplace_back(new Node(graph(), e.op, this));
    } else {
      // graph rewrite: no op; the revised graph is built when the child is first visited
      children_.emplace_back(new Node(d, this));
    }
  }
}

Node &Node::select(Strategy &strategy, std::mt19937_64 &rng) {
  if (is_leaf() || is_terminal()) return *this;
  const double inf = std::numeric_limits<double>::infinity();
  const double c = std::sqrt(2.0);
  std::vector<double> ucts;
  ucts.reserve(children_.size());
  for (const auto &child : children_) {
    const double exploit = strategy.select(*child);
    const double explore =
        child->fullyVisited_ ? -inf : c * std::sqrt(std::log(double(n_)) / double(child->n_));
    double u = exploit + explore;
    if (std::isnan(u)) u = -inf; // +inf exploit on a fully-visited child
    ucts.push_back(u);
  }
  const double best = *std::max_element(ucts.begin(), ucts.end());
  std::vector<size_t> ties;
  for (size_t i = 0; i < ucts.size(); ++i)
    if (ucts[i] == best) ties.push_back(i);
  const size_t pick = ties[std::uniform_int_distribution<size_t>(0, ties.size() - 1)(rng)];
  return children_[pick]->select(strategy, rng);
}

Node &Node::expand(const Platform &plat) {
  ensure_children(plat);
  if (children_.empty()) return *this;
  for (auto &c : children_)
    if (c->n_ == 0) return *c;
  THROW_RUNTIME("expand called on a node whose children have all been played: " << desc());
}

Node::RolloutResult Node::get_rollout(const Platform &plat, std::mt19937_64 &rng, bool expand) {
  RolloutResult res;
  if (parent_) res.sequence = parent_->get_sequence();
  if (expand) {
    if (op_) res.sequence.push_back(op_);
    SDP::State state(graph(), res.sequence); // of `cur`, derived level by level
    Node *cur = this;
    while (cur) {
      res.backpropStart = cur;
      cur->ensure_children(plat, &state);
      if (cur->children_.empty()) break;
      cur = cur->children_[std::uniform_int_distribution<size_t>(0, cur->children_.size() - 1)(rng)]
                .get();
      if (cur->op_) {
        state = state.apply(ExecuteOp(cur->op_));
      } else {
        state = state.apply(*cur->rewrite_);
        if (!cur->graph_) cur->graph_ = state.graph_ptr(); // the revised graph, built once
      }
    }
    res.sequence = state.sequence();
  } else {
    // play out on states only; nothing below this node is kept
    if (op_) res.sequence.push_back(op_);
    SDP::State state(graph(), res.sequence);
    while (!state.is_terminal()) {
      const auto decisions = state.get_decisions(plat);
      if (decisions.empty()) THROW_RUNTIME("rollout reached a dead end at " << state.sequence().desc());
      state = state.apply(
          *decisions[std::uniform_int_distribution<size_t>(0, decisions.size() - 1)(rng)]);
    }
    res.sequence = state.sequence();
    res.backpropStart = this;
  }
  return res;
}

void Node::backprop(Strategy &strategy, const Benchmark::Result &br) {
  for (Node *cur = this; cur; cur = cur->parent_) {
    ++cur->n_;
    if (cur->children_.empty()) {
      if (cur->expanded_) cur->fullyVisited_ = true;
    } else {
      bool all = true;
      for (const auto &c : cur->children_) all = all && c->fullyVisited_;
      if (all) cur->fullyVisited_ = true;
    }
    strategy.backprop(*cur, br);
  }
}

// ---------------------------------------------------------------- strategies
std::string Strategy::graphviz_label_line(const Node &node) const {
  if (node.state_.times.empty()) return "";
  std::ostringstream ss;
  ss << std::scientific;
  ss.precision(2);
  ss << node.state_.tMin << " - " << node.state_.tMax;
  return ss.str();
}

std::vector<uint64_t> histogram(const std::vector<double> &v, double tMin, double tMax, int nBins) {
  std::vector<uint64_t> h(size_t(nBins), 0);
  const double span = tMax - tMin;
  for (double e : v) {
    size_t i = span > 0 ? size_t(std::max(0.0, (e - tMin) / span * nBins)) : 0;
    if (i >= size_t(nBins)) i = size_t(nBins) - 1;
    ++h[i];
  }
  return h;
}

static double clamp01(double v) { return std::min(1.0, std::max(0.0, v)); }

// -- FastMin: how close the child's best time is to the best time seen at the root
double FastMin::select(const Node &child) {
  const Node &root = child.root();
  if (&child == &root) return 1;
  if (root.n_ < 2 || root.state_.tMax == root.state_.tMin) return 1;
  if (child.n_ < 1) return select(*child.parent_);
  return clamp01(1.0 - (child.state_.tMin - root.state_.tMin) / (root.state_.tMax - root.state_.tMin));
}
void FastMin::backprop(Node &node, const Benchmark::Result &br) { node.state_.add(br.pct10); }

// -- Coverage: how much of the parent's [fastest, slowest] range the child spans
double Coverage::select(const Node &child) {
  if (!child.parent_) return 1;
  const auto &pt = child.parent_->state_.times;
  const auto &ct = child.state_.times;
  if (pt.size() < 2 || ct.empty()) return 1;
  const double pMin = pt.front(), pMax = pt.back();
  if (pMin == pMax) return 1;
  if (ct.size() < 2) return clamp01(std::max(ct[0] - pMin, pMax - ct[0]) / (pMax - pMin));
  return clamp01((ct.back() - ct.front()) / (pMax - pMin));
}
void Coverage::backprop(Node &node, const Benchmark::Result &br) {
  node.state_.add(br.pct10);
  if (!node.parent_) {
    minT = node.state_.times.front();
    maxT = node.state_.times.back();
  }
}

// -- Random: one uniformly chosen child per parent per traversal
double Random::select(const Node &child) {
  const Node *parent = child.parent_;
  if (!parent) return 0;
  auto it = selected.find(parent);
  if (it == selected.end()) {
    const size_t pick =
        std::uniform_int_distribution<size_t>(0, parent->children_.size() - 1)(rng_);
    it = selected.emplace(parent, pick).first;
  }
  return parent->children_[it->second].get() == &child ? std::numeric_limits<double>::infinity()
                                                       : 0.0;
}
void Random::backprop(Node &node, const Benchmark::Result &br) {
  node.state_.add(br.pct10);
  if (!node.parent_) selected.clear();
}

// -- AvgTime: mean closeness-to-best of all the child's rollouts
double AvgTime::select(const Node &child) {
  const Node &root = child.root();
  if (child.n_ < 1 || root.n_ < 2 || root.state_.tMax == root.state_.tMin) return 0;
  double acc = 0;
  for (double t : child.state_.times)
    acc += clamp01(1.0 - (t - root.state_.tMin) / (root.state_.tMax - root.state_.tMin));
  return acc / double(child.state_.times.size());
}
void AvgTime::backprop(Node &node, const Benchmark::Result &br) { node.state_.add(br.pct10); }

// -- AntiCorrelation: prefer children whose time histogram differs from the parent's
double AntiCorrelation::select(const Node &child) {
  if (!child.parent_) return 0;
  const auto &pt = child.parent_->state_.times;
  const auto &ct = child.state_.times;
  if (pt.size() < 2 || ct.size() < 2) return 0;
  const double tMin = std::min(pt.front(), ct.front()), tMax = std::max(pt.back(), ct.back());
  const double c = corr(histogram(pt, tMin, tMax), histogram(ct, tMin, tMax));
  return clamp01((1.0 - c) / 2.0);
}
void AntiCorrelation::backprop(Node &node, const Benchmark::Result &br) {
  node.state_.add(br.pct10);
}

namespace {
// (1 - corr) or (1 + corr) of a node's histogram against the root's, over the root's range
double root_corr_score(const Node &root, const Node &n, bool anti) {
  const auto &rt = root.state_.times;
  const double tMin = rt.front(), tMax = rt.back();
  const double c = corr(histogram(rt, tMin, tMax), histogram(n.state_.times, tMin, tMax));
  return anti ? 1.0 - c : 1.0 + c; // [0, 2]
}
double sibling_normalised(const Node *root, const Node &child, bool anti) {
  const Node *parent = child.parent_;
  if (!parent || !root) return 0;
  if (parent->state_.times.size() < 2 || child.state_.times.size() < 2) return 0;
  if (root->state_.times.size() < 2) return 0;
  double best = 0;
  for (const auto &sib : parent->children_)
    if (!sib->state_.times.empty()) best = std::max(best, root_corr_score(*root, *sib, anti));
  if (best <= 0) return 0;
  return root_corr_score(*root, child, anti) / best;
}
} // namespace

// -- NormalizedAntiCorrelation: anti-correlation with the ROOT histogram, / best sibling
double NormalizedAntiCorrelation::select(const Node &child) {
  return sibling_normalised(root, child, true);
}
void NormalizedAntiCorrelation::backprop(Node &node, const Benchmark::Result &br) {
  node.state_.add(br.pct10);
  if (!node.parent_) root = &node;
}

// -- NormRootCorr: positive correlation with the root histogram, / best sibling
double NormRootCorr::select(const Node &child) { return sibling_normalised(root, child, false); }
void NormRootCorr::backprop(Node &node, const Benchmark::Result &br) {
  node.state_.add(br.pct10);
  if (!node.parent_) root = &node;
}

// -- BalanceHistogram: favour children that can feed the emptiest root histogram bin
double BalanceHistogram::select(const Node &child) {
  if (!child.parent_ || !root) return 0;
  if (child.parent_->state_.times.empty() || child.state_.times.empty()) return 0;
  const auto &rt = root->state_.times;
  const auto rHist = histogram(rt, rt.front(), rt.back());
  const auto cHist = histogram(child.state_.times, rt.front(), rt.back());
  int smallest = -1, largest = -1;
  for (size_t i = 0; i < rHist.size(); ++i) {
    if (rHist[i] > 0 && cHist[i] > 0 && (smallest < 0 || rHist[i] < rHist[smallest]))
      smallest = int(i);
    if (rHist[i] > 0 && (largest < 0 || rHist[i] > rHist[largest])) largest = int(i);
  }
  if (smallest < 0 || largest < 0) return 0;
  return 1.0 - double(rHist[smallest]) / double(rHist[largest]);
}
void BalanceHistogram::backprop(Node &node, const Benchmark::Result &br) {
  node.state_.add(br.pct10);
  if (!node.parent_) root = &node;
}

// -- Unvisited: force children that have never been measured
double Unvisited::select(const Node &child) {
  return child.state_.times.empty() ? std::numeric_limits<double>::infinity() : 0.0;
}
void Unvisited::backprop(Node &node, const Benchmark::Result &br) { node.state_.add(br.pct10); }

std::vector<std::string> strategy_names() {
  return {"FastMin",       "Coverage",        "Random",
          "AvgTime",       "AntiCorrelation", "NormalizedAntiCorrelation",
          "NormRootCorr",  "BalanceHistogram", "Unvisited"};
}

std::unique_ptr<Strategy> make_strategy(const std::string &name) {
#define TZ_CASE(N)                                                                                 \
  if (name == #N) return std::unique_ptr<Strategy>(new N())
  TZ_CASE(FastMin);
  TZ_CASE(Coverage);
  TZ_CASE(Random);
  TZ_CASE(AvgTime);
  TZ_CASE(AntiCorrelation);
  TZ_CASE(NormalizedAntiCorrelation);
  TZ_CASE(NormRootCorr);
  TZ_CASE(BalanceHistogram);
  TZ_CASE(Unvisited);
#undef TZ_CASE
  THROW_RUNTIME("unknown MCTS strategy '" << name << "'");
}

// ---------------------------------------------------------------- graphviz
std::string graphviz(const Node &root, const Strategy *strategy) {
  std::ostringstream os;
  auto escape = [](std::string s) {
    std::string o;
    for (char c : s) {
      if (c == '\n') o += "\\n";
      else if (c == '"') o += "\\\"";
      else o.push_back(c);
    }
    return o;
  };
  // children of fully-visited or once-played nodes, and never-played nodes, are hidden
  auto hides_children = [](const Node &n) { return n.fullyVisited_ || n.n_ == 1; };
  std::function<void(const Node &)> nodes = [&](const Node &n) {
    os << "node_" << &n << " [label=\""
       << escape(n.graphviz_name() + "\nrollouts: " + std::to_string(n.n_) + "\n" +
                 n.graphviz_label(strategy))
       << "\"";
    if (n.fullyVisited_) os << " fillcolor=green style=filled";
    else if (n.n_ == 1) os << " fillcolor=gray style=filled";
    os << "];\n";
    if (hides_children(n)) return;
    for (const auto &c : n.children_)
      if (c->n_ > 0) nodes(*c);
  };
  std::function<void(const Node &)> edges = [&](const Node &n) {
    if (hides_children(n)) return;
    for (const auto &c : n.children_)
      if (c->n_ > 0) os << "node_" << &n << " -> node_" << c.get() << ";\n";
    for (const auto &c : n.children_)
      if (c->n_ > 0) edges(*c);
  };
  os << "digraph D {\n";
  nodes(root);
  edges(root);
  os << "}\n";
  return os.str();
}

void dump_graphviz(const std::string &path, const Node &root, const Strategy *strategy) {
  std::ofstream os(path);
  if (!os) THROW_RUNTIME("cannot open " << path);
  os << graphviz(root, strategy);
}

// ---------------------------------------------------------------- search
Result explore(const Graph<OpBase> &g, Platform &plat, Benchmarker &benchmarker,
               Strategy &strategy, const Opts &opts) {
  Communicator &comm = plat.comm();
  const int rank = comm.rank();
  Result result;
  std::unique_ptr<Node> root;
  std::mt19937_64 rng(opts.seed);
  strategy.seed(opts.seed ^ 0x9e3779b97f4a7c15ULL);
  if (rank == 0) {
    root.reset(new Node(g));
    register_handler([&result](int) { result.dump_csv(); });
  }
  auto &ctr = counters::mcts();
  (void)ctr;

  for (size_t iter = 0; opts.nIters == 0 || iter < opts.nIters; ++iter) {
    Stop stop;
    if (rank == 0 && root->fullyVisited_) stop = Stop(true, Stop::Reason::full_tree);
    if (rank == 0 && !stop && opts.maxTreeNodes && root->size() >= opts.maxTreeNodes)
      stop = Stop(true, Stop::Reason::large_tree);
    stop.bcast(0, comm);
    if (stop) {
      TENZING_INFO("mcts: stop (" << stop.c_str() << ") after " << iter << " iterations");
      break;
    }

    Sequence<BoundOp> order;
    Node *endpoint = nullptr;
    if (rank == 0) {
      double t = counters::now();
      Node &selected = root->select(strategy, rng);
      TENZING_COUNTER_OP(mcts, SELECT_TIME, += counters::now() - t);
      t = counters::now();
      Node &child = selected.expand(plat);
      TENZING_COUNTER_OP(mcts, EXPAND_TIME, += counters::now() - t);
      t = counters::now();
      Node::RolloutResult rr = child.get_rollout(plat, rng, opts.expandRollout);
      TENZING_COUNTER_OP(mcts, ROLLOUT_TIME, += counters::now() - t);
      order = std::move(rr.sequence);
      endpoint = rr.backpropStart;
      t = counters::now();
      if (opts.removeRedundantSyncs) Schedule::remove_redundant_syncs(order);
      TENZING_COUNTER_OP(mcts, REDUNDANT_SYNC_TIME, += counters::now() - t);
    }
    order = bcast(order, g, comm, 0);

    double t = counters::now();
    plat.provision_events(order);
    TENZING_COUNTER_OP(mcts, RMAP_TIME, += counters::now() - t);

    comm.barrier();
    t = counters::now();
    const Benchmark::Result br = benchmarker.benchmark(order, plat, opts.benchOpts);
    TENZING_COUNTER_OP(mcts, BENCHMARK_TIME, += counters::now() - t);
    comm.barrier();

    {
      // every rank keeps the (identical, max-reduced) results; only rank 0 owns the tree
      SimResult sr;
      sr.path = order;
      sr.benchResult = br;
      result.simResults.push_back(std::move(sr));
    }
    if (rank == 0) {
      t = counters::now();
      endpoint->backprop(strategy, br);
      TENZING_COUNTER_OP(mcts, BACKPROP_TIME, += counters::now() - t);
      TENZING_COUNTER_OP(mcts, ITERS, += 1);
      if (opts.dumpTree &&
          (iter < 10 || (iter < 50 && iter % 10 == 0) || (iter >= 50 && iter < 100 && iter % 25 == 0))) {
        dump_graphviz(opts.dumpTreePrefix + std::to_string(iter) + ".dot", *root, &strategy);
      }
    }
  }
  if (rank == 0) unregister_handler();
  return result;
}

// ---------------------------------------------------------------- legacy frontier
std::vector<BoundOpPtr> get_frontier(const Platform &plat, const Graph<OpBase> &g,
                                     const std::vector<BoundOpPtr> &completed) {
  Sequence<BoundOp> path(completed);
  std::vector<BoundOpPtr> out;
  for (const OpPtr &op : g.frontier(completed)) {
    if (!std::dynamic_pointer_cast<GpuOp>(op) && !std::dynamic_pointer_cast<BoundOp>(op)) continue;
    for (const BoundOpPtr &v : make_platform_variations(plat, op)) {
      if (EventSynchronizer::is_synced(v, g, path)) {
        out.push_back(v);
      } else {
        for (const BoundOpPtr &s : EventSynchronizer::make_syncs(v, g, path)) out.push_back(s);
      }
    }
  }
  keep_uniques(out);
  return out;
}

} // namespace mcts
} // namespace tenzing
