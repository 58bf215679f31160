#include "tenzing/state.hpp"

#include <algorithm>
#include <map>
#include <unordered_set>

namespace tenzing {
namespace SDP {

State::State(const Graph<OpBase> &graph)
    : graph_(std::make_shared<const Graph<OpBase>>(graph)) {
  sequence_.push_back(std::dynamic_pointer_cast<BoundOp>(graph_->start()));
}

State::State(const Graph<OpBase> &graph, const Sequence<BoundOp> &sequence)
    : graph_(std::make_shared<const Graph<OpBase>>(graph)), sequence_(sequence) {}

std::vector<BoundOpPtr> State::get_syncs_before_op(const BoundOpPtr &op) const {
  if (Synchronizer::is_synced(op, *graph_, sequence_)) return {};
  return Synchronizer::make_syncs(op, *graph_, sequence_, true);
}

std::vector<DecisionPtr> State::get_decisions(const Platform &plat, bool quiet) const {
  std::vector<DecisionPtr> decisions;
  std::unordered_set<uint32_t> offeredSyncs; // one ExecuteOp per distinct sync op
  const PathIndex idx(sequence_);            // one pass over the path serves every frontier op
  for (const OpPtr &op : graph_->frontier(sequence_.vector())) {
    if (auto bop = std::dynamic_pointer_cast<BoundOp>(op)) {
      const std::vector<BoundOpPtr> syncs =
          Synchronizer::is_synced(bop, *graph_, sequence_, idx)
              ? std::vector<BoundOpPtr>{}
              : Synchronizer::make_syncs(bop, *graph_, sequence_, idx, true);
      if (syncs.empty()) {
        decisions.push_back(std::make_shared<ExecuteOp>(bop));
      } else {
        for (const BoundOpPtr &s : syncs)
          if (offeredSyncs.insert(s->uid()).second)
            decisions.push_back(std::make_shared<ExecuteOp>(s));
      }
    } else if (auto gop = std::dynamic_pointer_cast<GpuOp>(op)) {
      for (const Stream &s : plat.streams_)
        decisions.push_back(std::make_shared<AssignOpStream>(gop, s));
    } else if (auto comp = std::dynamic_pointer_cast<CompoundOp>(op)) {
      decisions.push_back(std::make_shared<ExpandOp>(comp));
    } else if (auto choice = std::dynamic_pointer_cast<ChoiceOp>(op)) {
      for (const OpPtr &alt : choice->choices())
        decisions.push_back(std::make_shared<ChooseOp>(choice, alt));
    } else {
      THROW_RUNTIME("op " << op->desc() << " is not Bound/Gpu/Compound/Choice");
    }
  }
  if (!quiet)
    for (const auto &d : decisions) TENZING_INFO("decision: " << d->desc());
  return decisions;
}

State State::apply(const Decision &d) const {
  switch (d.kind()) {
  case Decision::Kind::ExecuteOp: {
    State r(*this);
    r.keyCache_.clear();
    r.sequence_.push_back(static_cast<const ExecuteOp &>(d).op);
    return r;
  }
  case Decision::Kind::ExpandOp: {
    const auto &e = static_cast<const ExpandOp &>(d);
    return State(graph_->clone_but_expand(e.op, e.op->graph()), sequence_);
  }
  case Decision::Kind::AssignOpStream: {
    const auto &a = static_cast<const AssignOpStream &>(d);
    return State(graph_->clone_but_replace(std::make_shared<BoundGpuOp>(a.op, a.stream), a.op),
                 sequence_);
  }
  case Decision::Kind::ChooseOp: {
    const auto &c = static_cast<const ChooseOp &>(d);
    return State(graph_->clone_but_replace(c.replacement, c.orig), sequence_);
  }
  }
  THROW_RUNTIME("failed to apply decision, unexpected Decision type");
}

std::vector<State> State::frontier(const Platform &plat, bool quiet) const {
  std::vector<State> result;
  std::unordered_set<std::string> seen;
  for (const auto &d : get_decisions(plat, quiet)) {
    State s = apply(*d);
    if (seen.insert(s.canonical_key()).second) result.push_back(std::move(s));
  }
  return result;
}

bool State::is_terminal() const {
  return !sequence_.empty() && sequence_.back()->uid() == graph_->finish()->uid();
}

// Sequence first, then the graph's vertices in a binding-independent order, with ONE
// renumbering of streams and events (first use) across both, so sequence and graph agree on
// the renaming.  Ops are named by their interned unbound identity (kind + name, a 32-bit id
// that every op caches) instead of by their name strings: the key is compared within one
// process only (DFS de-duplication), and building it is the inner loop of the enumeration.
std::string State::canonical_key() const {
  if (!keyCache_.empty()) return keyCache_;
  std::vector<std::pair<int, int>> sMap, eMap; // id -> first-use rank; a handful of entries
  auto relabel = [](std::vector<std::pair<int, int>> &m, int id) {
    for (const auto &kv : m)
      if (kv.first == id) return kv.second;
    const int n = static_cast<int>(m.size());
    m.emplace_back(id, n);
    return n;
  };
  std::string key;
  auto put = [&key](char tag, uint32_t v) {
    key.push_back(tag);
    key.append(reinterpret_cast<const char *>(&v), sizeof(v));
  };
  auto emit = [&](const OpBase &op) {
    if (const auto *sync = dynamic_cast<const SyncOpBase *>(&op)) {
      // synchronisation ops are the same op whatever they are called: kind, not name
      put('k', intern_key(sync->kind()));
    } else {
      put('o', op.unbound_uid());
    }
    if (const auto *hs = dynamic_cast<const HasStream *>(&op))
      for (const Stream &st : hs->get_streams()) put('s', uint32_t(relabel(sMap, st.id())));
    if (const auto *he = dynamic_cast<const HasEvent *>(&op))
      for (const Event &e : he->get_events()) put('e', uint32_t(relabel(eMap, e.id())));
  };
  key.reserve(16 * (sequence_.size() + 4 * graph_->vertex_size()));
  for (const auto &op : sequence_) emit(*op);
  key.push_back('#');
  std::vector<OpPtr> verts = graph_->vertices();
  std::sort(verts.begin(), verts.end(),
            [](const OpPtr &x, const OpPtr &y) { return x->unbound_uid() < y->unbound_uid(); });
  std::vector<uint32_t> succIds;
  for (const auto &v : verts) {
    emit(*v);
    key.push_back('>');
    succIds.clear();
    for (const auto &sc : graph_->succs(v)) succIds.push_back(sc->unbound_uid());
    std::sort(succIds.begin(), succIds.end());
    for (uint32_t id : succIds) put(',', id);
    key.push_back('|');
  }
  keyCache_ = key;
  return key;
}

Equivalence get_equivalence(const State &a, const State &b) {
  Equivalence seqEq = tenzing::get_equivalence(a.sequence(), b.sequence());
  if (!seqEq) return Equivalence::falsy();
  Equivalence graphEq = tenzing::get_equivalence(a.graph(), b.graph());
  if (!graphEq) return Equivalence::falsy();
  // one renaming must serve both
  if (!seqEq.merge(graphEq)) return Equivalence::falsy();
  return seqEq;
}

} // namespace SDP
} // namespace tenzing
