// The program as a DAG of operations, always bracketed by Start and Finish.
//
// Reference: include/tenzing/graph.hpp:19-580, src/graph.cpp:13-420.  Same
// vocabulary (start_then / then / then_finish / clone_but_replace /
// clone_but_expand / frontier / dump_graphviz ...), different representation:
// vertices are slots of a dense op table, edges are integer adjacency lists
// (short ones stored inline) in a block that copies share until an edge
// changes, and sorted flat arrays map an op's interned uid to its slot.
// Rewrites share the op objects, so the solvers can afford a graph per
// decision: binding one op copies the op table and the two indexes.
#pragma once

#include <algorithm>
#include <fstream>
#include <memory>
#include <sstream>
#include <unordered_map>
#include <vector>

#include "tenzing/cuda/ops_cuda.hpp"
#include "tenzing/operation.hpp"

namespace tenzing {

namespace detail {
// The solvers make a graph per rewrite decision (a stream assignment copies the whole program
// graph), so a copy must cost a few block copies, not two allocations per vertex plus a node
// per index entry: adjacency lists keep their first entries inline and the indexes are sorted
// flat arrays.

// a short list of vertex numbers; up to N live inside the object
template <int N> class SmallIds {
  int n_ = 0;
  int inl_[N];
  std::vector<int> more_; // entries N.. when the list is longer

public:
  int size() const { return n_; }
  bool empty() const { return n_ == 0; }
  int operator[](int i) const { return i < N ? inl_[i] : more_[size_t(i - N)]; }
  void push_back(int x) {
    if (n_ < N) inl_[n_] = x;
    else more_.push_back(x);
    ++n_;
  }
  void clear() {
    n_ = 0;
    more_.clear();
  }
  bool contains(int x) const {
    for (int i = 0; i < n_; ++i)
      if ((*this)[i] == x) return true;
    return false;
  }
  void remove(int x) {
    int w = 0;
    const int n = n_;
    std::vector<int> keep;
    for (int i = 0; i < n; ++i) {
      const int v = (*this)[i];
      if (v == x) continue;
      if (w < N) inl_[w] = v;
      else keep.push_back(v);
      ++w;
    }
    n_ = w;
    more_.swap(keep);
  }
  class const_iterator {
    const SmallIds *s_;
    int i_;

  public:
    using iterator_category = std::forward_iterator_tag;
    using value_type = int;
    using difference_type = std::ptrdiff_t;
    using pointer = const int *;
    using reference = int;
    const_iterator(const SmallIds *s, int i) : s_(s), i_(i) {}
    int operator*() const { return (*s_)[i_]; }
    const_iterator &operator++() {
      ++i_;
      return *this;
    }
    bool operator!=(const const_iterator &o) const { return i_ != o.i_; }
    bool operator==(const const_iterator &o) const { return i_ == o.i_; }
  };
  const_iterator begin() const { return const_iterator(this, 0); }
  const_iterator end() const { return const_iterator(this, n_); }
};

// uid -> vertex number, sorted by uid
class FlatIndex {
  std::vector<std::pair<uint32_t, int>> v_;
  std::vector<std::pair<uint32_t, int>>::const_iterator lower(uint32_t k) const {
    return std::lower_bound(v_.begin(), v_.end(), k,
                            [](const std::pair<uint32_t, int> &x, uint32_t key) { return x.first < key; });
  }

public:
  int find(uint32_t k) const { // -1 if absent
    auto it = lower(k);
    return it == v_.end() || it->first != k ? -1 : it->second;
  }
  bool count(uint32_t k) const { return find(k) >= 0; }
  // keeps an existing entry (like unordered_map::emplace)
  void emplace(uint32_t k, int id) {
    auto it = lower(k);
    if (it != v_.end() && it->first == k) return;
    v_.insert(it, std::make_pair(k, id));
  }
  void erase(uint32_t k) {
    auto it = lower(k);
    if (it != v_.end() && it->first == k) v_.erase(it);
  }
  void clear() { v_.clear(); }
};
} // namespace detail

template <typename T> class Graph {
public:
  using op_t = std::shared_ptr<T>;
  using vid_t = int;
  using ids_t = detail::SmallIds<6>;

private:
  // A vertex is a slot number.  The op table belongs to the graph; the edges live in a block
  // that a graph shares with its copies until one of them changes an edge: binding an op to a
  // stream or choosing an alternative (the bulk of the search's rewrites) copies the op table
  // and the indexes, and no adjacency list.
  struct Adj {
    ids_t succs, preds;
    bool alive = true;
  };
  std::vector<op_t> ops_;
  std::shared_ptr<std::vector<Adj>> adj_ = std::make_shared<std::vector<Adj>>();
  const std::vector<Adj> &adj() const { return *adj_; }
  std::vector<Adj> &adj_mut() {
    if (adj_.use_count() > 1) adj_ = std::make_shared<std::vector<Adj>>(*adj_);
    return *adj_;
  }
  detail::FlatIndex index_;  // op uid -> vertex
  detail::FlatIndex uindex_; // op unbound uid -> vertex
  vid_t start_ = -1, finish_ = -1;

  static void add_unique(ids_t &v, vid_t x) {
    if (!v.contains(x)) v.push_back(x);
  }
  static void remove_one(ids_t &v, vid_t x) { v.remove(x); }
  vid_t add_vertex(const op_t &op) {
    const vid_t have = index_.find(op->uid());
    if (have >= 0) return have;
    vid_t id = static_cast<vid_t>(ops_.size());
    ops_.push_back(op);
    adj_mut().push_back(Adj{});
    index_.emplace(op->uid(), id);
    uindex_.emplace(op->unbound_uid(), id);
    return id;
  }
  void add_edge(vid_t a, vid_t b) {
    if (adj()[a].succs.contains(b) && adj()[b].preds.contains(a)) return;
    auto &m = adj_mut();
    add_unique(m[a].succs, b);
    add_unique(m[b].preds, a);
  }
  std::vector<op_t> ops_of(const ids_t &ids) const {
    std::vector<op_t> r;
    r.reserve(size_t(ids.size()));
    for (vid_t i : ids) r.push_back(ops_[i]);
    return r;
  }
  vid_t must_find(const OpBase &op) const {
    vid_t i = find_vertex(op);
    if (i < 0) THROW_RUNTIME("op " << op.desc() << " is not in the graph");
    return i;
  }

public:
  Graph() {
    start_ = add_vertex(std::make_shared<Start>());
    finish_ = add_vertex(std::make_shared<Finish>());
  }

  // ---- construction ----
  const op_t &start() const { return ops_[start_]; }
  const op_t &finish() const { return ops_[finish_]; }

  // Start -> op
  const op_t &start_then(const op_t &op) {
    vid_t b = add_vertex(op);
    add_edge(start_, b);
    return ops_[b];
  }
  // a -> b; neither may be Start or Finish (use start_then / then_finish)
  const op_t &then(const op_t &a, const op_t &b) {
    if (a->uid() == ops_[start_]->uid() || a->uid() == ops_[finish_]->uid() ||
        b->uid() == ops_[start_]->uid() || b->uid() == ops_[finish_]->uid()) {
      THROW_RUNTIME("Graph::then does not take Start/Finish; use start_then/then_finish");
    }
    vid_t ia = add_vertex(a), ib = add_vertex(b);
    add_edge(ia, ib);
    return ops_[ib];
  }
  // op -> Finish
  void then_finish(const op_t &op) {
    vid_t a = add_vertex(op);
    add_edge(a, finish_);
  }
  // generic edge that accepts Start/Finish endpoints (used by graph builders)
  void add_edge(const op_t &a, const op_t &b) {
    vid_t ia = add_vertex(a), ib = add_vertex(b);
    if (ia == finish_ || ib == start_) THROW_RUNTIME("edge out of Finish / into Start");
    add_edge(ia, ib);
  }

  // ---- queries ----
  size_t vertex_size() const {
    size_t n = 0;
    for (const auto &a : adj()) n += a.alive;
    return n;
  }
  size_t edge_size() const {
    size_t n = 0;
    for (const auto &a : adj())
      if (a.alive) n += a.succs.size();
    return n;
  }
  vid_t find_vertex(const OpBase &op) const { return index_.find(op.uid()); }
  // vertex whose unbound identity equals op's unbound identity
  vid_t find_vertex_unbound(const OpBase &op) const { return uindex_.find(op.unbound_uid()); }
  // by vertex number, for callers that walk the graph without copying op lists
  const op_t &op_at(vid_t i) const { return ops_[size_t(i)]; }
  const ids_t &pred_ids(vid_t i) const { return adj()[size_t(i)].preds; }
  const ids_t &succ_ids(vid_t i) const { return adj()[size_t(i)].succs; }
  size_t vertex_slots() const { return ops_.size(); } // vertex numbers are < this (dead ones included)
  bool contains(const op_t &op) const { return op && find_vertex(*op) >= 0; }
  bool contains_unbound(const op_t &op) const { return op && find_vertex_unbound(*op) >= 0; }
  // the graph's own pointer for an op equal to `op` (null if absent)
  op_t find(const op_t &op) const {
    vid_t i = find_vertex(*op);
    return i < 0 ? nullptr : ops_[i];
  }
  op_t find_unbound(const op_t &op) const {
    vid_t i = find_vertex_unbound(*op);
    return i < 0 ? nullptr : ops_[i];
  }
  op_t find_name(const std::string &name) const {
    for (size_t i = 0; i < ops_.size(); ++i)
      if (adj()[i].alive && ops_[i]->name() == name) return ops_[i];
    return nullptr;
  }

  std::vector<op_t> vertices() const {
    std::vector<op_t> r;
    for (size_t i = 0; i < ops_.size(); ++i)
      if (adj()[i].alive) r.push_back(ops_[i]);
    return r;
  }
  template <typename U> std::vector<std::shared_ptr<U>> nodes_cast() const {
    std::vector<std::shared_ptr<U>> r;
    for (size_t i = 0; i < ops_.size(); ++i)
      if (adj()[i].alive)
        if (auto p = std::dynamic_pointer_cast<U>(ops_[i])) r.push_back(p);
    return r;
  }
  std::vector<op_t> start_vertices() const { return ops_of(adj()[start_].succs); }
  std::vector<op_t> finish_vertices() const { return ops_of(adj()[finish_].preds); }

  std::vector<op_t> succs(const op_t &op) const { return ops_of(adj()[must_find(*op)].succs); }
  std::vector<op_t> preds(const op_t &op) const { return ops_of(adj()[must_find(*op)].preds); }
  std::vector<op_t> succs_vec(const op_t &op) const { return succs(op); }
  std::vector<op_t> preds_vec(const op_t &op) const { return preds(op); }
  // (found, list) - `found` false when the op is not a vertex
  std::pair<bool, std::vector<op_t>> preds_find(const op_t &op) const {
    vid_t i = find_vertex(*op);
    return i < 0 ? std::make_pair(false, std::vector<op_t>{})
                 : std::make_pair(true, ops_of(adj()[i].preds));
  }
  std::pair<bool, std::vector<op_t>> succs_find(const op_t &op) const {
    vid_t i = find_vertex(*op);
    return i < 0 ? std::make_pair(false, std::vector<op_t>{})
                 : std::make_pair(true, ops_of(adj()[i].succs));
  }
  std::pair<bool, std::vector<op_t>> preds_find_or_find_unbound(const op_t &op) const {
    vid_t i = find_vertex_unbound(*op);
    return i < 0 ? std::make_pair(false, std::vector<op_t>{})
                 : std::make_pair(true, ops_of(adj()[i].preds));
  }
  std::pair<bool, std::vector<op_t>> succs_find_or_find_unbound(const op_t &op) const {
    vid_t i = find_vertex_unbound(*op);
    return i < 0 ? std::make_pair(false, std::vector<op_t>{})
                 : std::make_pair(true, ops_of(adj()[i].succs));
  }
  std::pair<bool, std::vector<op_t>> succs_find_name(const std::string &name) const {
    for (size_t i = 0; i < ops_.size(); ++i)
      if (adj()[i].alive && ops_[i]->name() == name) return {true, ops_of(adj()[i].succs)};
    return {false, {}};
  }
  std::pair<bool, std::vector<op_t>> preds_find_name(const std::string &name) const {
    for (size_t i = 0; i < ops_.size(); ++i)
      if (adj()[i].alive && ops_[i]->name() == name) return {true, ops_of(adj()[i].preds)};
    return {false, {}};
  }

  // Structural invariants every well-formed program graph satisfies: adjacency lists mirror
  // each other and name live vertices only, the indexes point at the right vertices, nothing
  // enters Start or leaves Finish, the graph is acyclic, and every live vertex lies on a path
  // Start -> ... -> Finish.  Returns an empty string if all hold, else the first violation.
  // Builds with TENZING_RUNTIME_SANITY_CHECKS run it after every rewrite (reference:
  // CMakeLists.txt always defines that macro; here it is an option, off by default).
  std::string validate() const {
    const vid_t n = static_cast<vid_t>(ops_.size());
    std::ostringstream err;
    auto live = [&](vid_t i) { return i >= 0 && i < n && adj()[i].alive; };
    for (vid_t i = 0; i < n; ++i) {
      if (!adj()[i].alive) continue;
      for (vid_t s : adj()[i].succs) {
        if (!live(s)) { err << ops_[i]->desc() << " has a dead successor"; return err.str(); }
        if (!adj()[s].preds.contains(i)) {
          err << "edge " << ops_[i]->desc() << " -> " << ops_[s]->desc() << " has no mirror";
          return err.str();
        }
      }
      for (vid_t p : adj()[i].preds) {
        if (!live(p)) { err << ops_[i]->desc() << " has a dead predecessor"; return err.str(); }
        if (!adj()[p].succs.contains(i)) {
          err << "edge " << ops_[p]->desc() << " -> " << ops_[i]->desc() << " has no mirror";
          return err.str();
        }
      }
      if (index_.find(ops_[i]->uid()) != i) {
        err << ops_[i]->desc() << " is not indexed";
        return err.str();
      }
    }
    if (!live(start_) || !live(finish_)) return "Start or Finish missing";
    if (!adj()[start_].preds.empty()) return "an edge enters Start";
    if (!adj()[finish_].succs.empty()) return "an edge leaves Finish";
    // Kahn: all live vertices must come out, and only Start may have no predecessor
    std::vector<int> indeg(n, 0);
    std::vector<vid_t> ready;
    size_t nLive = 0, seen = 0;
    for (vid_t i = 0; i < n; ++i) {
      if (!adj()[i].alive) continue;
      ++nLive;
      indeg[i] = int(adj()[i].preds.size());
      if (indeg[i] == 0) {
        if (i != start_) { err << ops_[i]->desc() << " is not reachable from Start"; return err.str(); }
        ready.push_back(i);
      }
      if (adj()[i].succs.empty() && i != finish_) {
        err << ops_[i]->desc() << " does not lead to Finish";
        return err.str();
      }
    }
    while (!ready.empty()) {
      vid_t i = ready.back();
      ready.pop_back();
      ++seen;
      for (vid_t s : adj()[i].succs)
        if (--indeg[s] == 0) ready.push_back(s);
    }
    if (seen != nLive) return "the graph has a cycle";
    return "";
  }
  void sanity_check(const char *where) const {
#ifdef TENZING_RUNTIME_SANITY_CHECKS
    const std::string e = validate();
    if (!e.empty()) THROW_RUNTIME("graph invariant broken after " << where << ": " << e);
#else
    (void)where;
#endif
  }

  // ---- rewrites ----
  // deep copy: every op is cloned
  Graph clone() const {
    Graph g(*this);
    g.index_.clear();
    g.uindex_.clear();
    for (vid_t i = 0; i < static_cast<vid_t>(g.ops_.size()); ++i) {
      if (!g.adj()[i].alive) continue;
      g.ops_[i] = std::dynamic_pointer_cast<T>(ops_[i]->clone());
      g.index_.emplace(g.ops_[i]->uid(), i);
      g.uindex_.emplace(g.ops_[i]->unbound_uid(), i);
    }
    return g;
  }
  // in place: vertex `src` now holds `dst` (same edges)
  void replace(const op_t &src, const op_t &dst) {
    vid_t i = must_find(*src);
    if (dst->uid() != src->uid() && index_.count(dst->uid()))
      THROW_RUNTIME("replacement " << dst->desc() << " is already in the graph");
    index_.erase(ops_[i]->uid());
    uindex_.erase(ops_[i]->unbound_uid());
    ops_[i] = dst;
    index_.emplace(dst->uid(), i);
    uindex_.emplace(dst->unbound_uid(), i);
  }
  // copy with `src` replaced by `dst` (ops shared, not cloned)
  Graph clone_but_replace(const op_t &dst, const op_t &src) const {
    Graph g(*this);
    g.replace(src, dst);
    g.sanity_check("clone_but_replace");
    return g;
  }
  // copy with vertex `op` replaced by the interior of `sub`:
  //   preds(op) -> sub.start_vertices ; sub.finish_vertices -> succs(op)
  Graph clone_but_expand(const op_t &op, const Graph &sub) const {
    Graph g(*this);
    const vid_t me = g.must_find(*op);
    const ids_t myPreds = g.adj()[me].preds, mySuccs = g.adj()[me].succs;
    g.erase_vertex(me);
    std::vector<vid_t> map(sub.ops_.size(), -1);
    for (vid_t i = 0; i < static_cast<vid_t>(sub.ops_.size()); ++i) {
      if (!sub.adj()[i].alive || i == sub.start_ || i == sub.finish_) continue;
      if (g.index_.count(sub.ops_[i]->uid()))
        THROW_RUNTIME("expanding " << op->name() << ": inner op " << sub.ops_[i]->name()
                                   << " already exists in the outer graph");
      map[i] = g.add_vertex(sub.ops_[i]);
    }
    for (vid_t i = 0; i < static_cast<vid_t>(sub.ops_.size()); ++i) {
      if (map[i] < 0) continue;
      for (vid_t s : sub.adj()[i].succs) {
        if (s == sub.finish_) {
          for (vid_t o : mySuccs) g.add_edge(map[i], o);
        } else {
          g.add_edge(map[i], map[s]);
        }
      }
      for (vid_t p : sub.adj()[i].preds) {
        if (p == sub.start_)
          for (vid_t o : myPreds) g.add_edge(o, map[i]);
      }
    }
    // an empty compound (Start -> Finish) collapses to direct edges
    for (vid_t s : sub.adj()[sub.start_].succs)
      if (s == sub.finish_)
        for (vid_t p : myPreds)
          for (vid_t q : mySuccs) g.add_edge(p, q);
    g.sanity_check("clone_but_expand");
    return g;
  }
  // remove a vertex and its edges
  void erase(const op_t &op) { erase_vertex(must_find(*op)); }
  void erase_edge_only(const op_t &a, const op_t &b) {
    vid_t ia = must_find(*a), ib = must_find(*b);
    auto &m = adj_mut();
    remove_one(m[ia].succs, ib);
    remove_one(m[ib].preds, ia);
  }

  // ops whose predecessors are all in `done` and that are not in `done`
  // themselves; comparison ignores stream bindings
  template <typename U> std::vector<op_t> frontier(const std::vector<std::shared_ptr<U>> &done) const {
    std::vector<char> isDone(ops_.size(), 0);
    for (const auto &d : done) {
      vid_t i = find_vertex_unbound(*d);
      if (i >= 0) isDone[i] = 1;
    }
    std::vector<op_t> r;
    for (vid_t i = 0; i < static_cast<vid_t>(ops_.size()); ++i) {
      if (!adj()[i].alive || isDone[i]) continue;
      bool ready = true;
      for (vid_t p : adj()[i].preds) {
        if (!isDone[p]) {
          ready = false;
          break;
        }
      }
      if (ready) r.push_back(ops_[i]);
    }
    return r;
  }

  // the same from a per-vertex "done" table (see mark_done): no lookups, what a search that
  // extends one path step by step uses
  std::vector<op_t> frontier_of_mask(const std::vector<char> &isDone) const {
    std::vector<op_t> r;
    for (vid_t i = 0; i < static_cast<vid_t>(ops_.size()); ++i) {
      if (!adj()[i].alive || (size_t(i) < isDone.size() && isDone[size_t(i)])) continue;
      bool ready = true;
      for (vid_t p : adj()[i].preds) {
        if (!(size_t(p) < isDone.size() && isDone[size_t(p)])) {
          ready = false;
          break;
        }
      }
      if (ready) r.push_back(ops_[i]);
    }
    return r;
  }
  std::vector<vid_t> frontier_ids_of_mask(const std::vector<char> &isDone) const {
    std::vector<vid_t> r;
    for (vid_t i = 0; i < static_cast<vid_t>(ops_.size()); ++i) {
      if (!adj()[i].alive || (size_t(i) < isDone.size() && isDone[size_t(i)])) continue;
      bool ready = true;
      for (vid_t p : adj()[i].preds) {
        if (!(size_t(p) < isDone.size() && isDone[size_t(p)])) {
          ready = false;
          break;
        }
      }
      if (ready) r.push_back(i);
    }
    return r;
  }
  // record in `isDone` (indexed by vertex, grown as needed) that `op` has been executed;
  // ops that are not vertices (synchronisation inserted by the search) leave it unchanged.
  // Vertex numbers survive clone_but_replace / clone_but_expand, so a table filled against a
  // graph stays valid for the graphs rewritten from it.
  void mark_done(std::vector<char> &isDone, const OpBase &op) const {
    const vid_t i = find_vertex_unbound(op);
    if (i < 0) return;
    if (isDone.size() < ops_.size()) isDone.resize(ops_.size(), 0);
    isDone[size_t(i)] = 1;
  }

  // ---- output ----
  std::string str() const {
    std::ostringstream ss;
    for (size_t i = 0; i < ops_.size(); ++i) {
      if (!adj()[i].alive) continue;
      ss << ops_[i]->desc() << " ->";
      for (vid_t s : adj()[i].succs) ss << " " << ops_[size_t(s)]->desc();
      ss << "\n";
    }
    return ss.str();
  }
  void dump() const { std::cerr << str(); }
  std::string graphviz() const {
    std::ostringstream ss;
    ss << "digraph D {\n";
    for (vid_t i = 0; i < static_cast<vid_t>(ops_.size()); ++i) {
      if (!adj()[i].alive) continue;
      ss << "  op_" << i << " [label=\"" << ops_[i]->name();
      if (auto hs = std::dynamic_pointer_cast<HasStream>(ops_[i])) {
        ss << "\\nstream";
        for (const Stream &s : hs->get_streams()) ss << " " << s.str();
      }
      if (auto he = std::dynamic_pointer_cast<HasEvent>(ops_[i])) {
        ss << "\\nevent";
        for (const Event &e : he->get_events()) ss << " " << e.str();
      }
      ss << "\"];\n";
    }
    for (vid_t i = 0; i < static_cast<vid_t>(ops_.size()); ++i) {
      if (!adj()[i].alive) continue;
      for (vid_t s : adj()[i].succs) ss << "  op_" << i << " -> op_" << s << ";\n";
    }
    ss << "}\n";
    return ss.str();
  }
  void dump_graphviz(const std::string &path) const {
    std::ofstream os(path);
    if (!os) THROW_RUNTIME("cannot open " << path);
    os << graphviz();
  }

private:
  void erase_vertex(vid_t i) {
    if (i == start_ || i == finish_) THROW_RUNTIME("cannot erase Start/Finish");
    auto &m = adj_mut();
    for (vid_t p : m[i].preds) remove_one(m[p].succs, i);
    for (vid_t s : m[i].succs) remove_one(m[s].preds, i);
    m[i].preds.clear();
    m[i].succs.clear();
    m[i].alive = false;
    index_.erase(ops_[i]->uid());
    uindex_.erase(ops_[i]->unbound_uid());
  }
};

// ---- stream-assignment enumerators and equivalence (src/core/graph.cpp) ----
// reference: src/graph.cpp:42-90,94-121,155-234,236-346,348-420

// every assignment of the graph's GpuOps to `streams` (S^k graphs)
std::vector<Graph<OpBase>> use_streams(const Graph<OpBase> &g, const std::vector<Stream> &streams);
// canonical assignments only: the i-th GpuOp (in vertex order) may use streams
// 0..min(i, S-1) and may open at most one new stream, which removes
// assignments that differ only by renaming streams
std::vector<Graph<OpBase>> use_streams2(const Graph<OpBase> &g, const std::vector<Stream> &streams);
// bind GpuOps by name -> stream
Graph<OpBase> apply_assignment(const Graph<OpBase> &g,
                               const std::vector<std::pair<std::string, Stream>> &assignment);
bool is_equivalent_stream_mapping(const Graph<OpBase> &a, const Graph<OpBase> &b);
// same vertices (by name), same edges, streams/events consistent under one bijection
Equivalence get_equivalence(const Graph<OpBase> &a, const Graph<OpBase> &b);
// make every cross-stream / gpu->cpu edge explicit by inserting CER/CSWE/CES
// vertices (the reference declares this and never defines it)
Graph<OpBase> insert_synchronization(const Graph<OpBase> &g);

} // namespace tenzing
