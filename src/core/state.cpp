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

// sequence first, then graph vertices in name order, one renumbering of streams
// and events across both (so sequence and graph agree on the renaming)
std::string State::canonical_key() const {
  if (!keyCache_.empty()) return keyCache_;
  std::map<int, int> sMap, eMap;
  auto relabel = [](std::map<int, int> &m, int id) {
    auto it = m.find(id);
    if (it != m.end()) return it->second;
    const int n = static_cast<int>(m.size());
    m.emplace(id, n);
    return n;
  };
  auto emit = [&](const OpBase &op, std::string &key) {
    const bool sync = dynamic_cast<const SyncOpBase *>(&op) != nullptr;
    key += sync ? op.kind() : op.name();
    if (const auto *hs = dynamic_cast<const HasStream *>(&op))
      for (const Stream &s : hs->get_streams()) key += "/s" + std::to_string(relabel(sMap, s.id()));
    if (const auto *he = dynamic_cast<const HasEvent *>(&op))
      for (const Event &e : he->get_events()) key += "/e" + std::to_string(relabel(eMap, e.id()));
  };
  std::string key;
  for (const auto &op : sequence_) {
    emit(*op, key);
    key += "|";
  }
  key += "#";
  std::vector<OpPtr> verts = graph_->vertices();
  std::sort(verts.begin(), verts.end(),
            [](const OpPtr &a, const OpPtr &b) { return a->name() < b->name(); });
  for (const auto &v : verts) {
    emit(*v, key);
    key += ">";
    std::vector<std::string> succNames;
    for (const auto &s : graph_->succs(v)) succNames.push_back(s->name());
    std::sort(succNames.begin(), succNames.end());
    for (const auto &n : succNames) key += n + ",";
    key += "|";
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
