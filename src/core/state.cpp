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
  std::vector<uint32_t> offeredSyncs;        // one ExecuteOp per distinct sync op
  pathIndex_.extend(sequence_);              // one index serves every frontier op; only new ops cost
  const PathIndex &idx = pathIndex_;
  for (; doneCounted_ < sequence_.size(); ++doneCounted_) graph_->mark_done(done_, *sequence_[doneCounted_]);
  if (synced_.size() < graph_->vertex_slots()) synced_.resize(graph_->vertex_slots(), 0);
  for (const int v : graph_->frontier_ids_of_mask(done_)) {
    const OpPtr &op = graph_->op_at(v);
    const uint16_t t = op->traits();
    if (t & op_trait::Bound) {
      const BoundOpPtr bop = std::static_pointer_cast<BoundOp>(op);
      // A frontier op that was synchronised with its predecessors stays so while the path only
      // grows (its predecessors are all executed, and records and waits are never taken back),
      // so the answer is kept per vertex and inherited by the states derived from this one.
      std::vector<BoundOpPtr> syncs;
      if (!synced_[size_t(v)]) {
        syncs = Synchronizer::make_syncs(bop, *graph_, sequence_, idx, true); // empty iff synced
        if (syncs.empty()) synced_[size_t(v)] = 1;
      }
      if (syncs.empty()) {
        decisions.push_back(std::make_shared<ExecuteOp>(bop));
      } else {
        for (const BoundOpPtr &s : syncs) {
          if (std::find(offeredSyncs.begin(), offeredSyncs.end(), s->uid()) != offeredSyncs.end()) continue;
          offeredSyncs.push_back(s->uid());
          decisions.push_back(std::make_shared<ExecuteOp>(s));
        }
      }
    } else if (t & op_trait::Gpu) {
      const auto gop = std::static_pointer_cast<GpuOp>(op);
      for (const Stream &s : plat.streams_)
        decisions.push_back(std::make_shared<AssignOpStream>(gop, s));
    } else if (t & op_trait::Compound) {
      decisions.push_back(std::make_shared<ExpandOp>(std::static_pointer_cast<CompoundOp>(op)));
    } else if (t & op_trait::Choice) {
      const auto choice = std::static_pointer_cast<ChoiceOp>(op);
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

State State::with_graph(Graph<OpBase> &&g) const {
  State r(std::make_shared<const Graph<OpBase>>(std::move(g)), sequence_);
  r.seqDigest_ = seqDigest_;
  r.seqDigested_ = seqDigested_;
  r.done_ = done_; // vertex numbers survive the rewrite
  r.doneCounted_ = doneCounted_;
  r.pathIndex_ = pathIndex_;
  r.synced_ = synced_;
  return r;
}

State State::apply(const Decision &d) const {
  switch (d.kind()) {
  case Decision::Kind::ExecuteOp: {
    State r(*this); // shares the graph and its prepared tokens, carries the sequence digest on
    r.keyCache_.clear();
    r.sequence_.push_back(static_cast<const ExecuteOp &>(d).op);
    return r;
  }
  case Decision::Kind::ExpandOp: {
    const auto &e = static_cast<const ExpandOp &>(d);
    return with_graph(graph_->clone_but_expand(e.op, e.op->graph()));
  }
  case Decision::Kind::AssignOpStream: {
    const auto &a = static_cast<const AssignOpStream &>(d);
    return with_graph(graph_->clone_but_replace(std::make_shared<BoundGpuOp>(a.op, a.stream), a.op));
  }
  case Decision::Kind::ChooseOp: {
    const auto &c = static_cast<const ChooseOp &>(d);
    return with_graph(graph_->clone_but_replace(c.replacement, c.orig));
  }
  }
  THROW_RUNTIME("failed to apply decision, unexpected Decision type");
}

std::vector<State> State::frontier(const Platform &plat, bool quiet) const {
  std::vector<State> result;
  std::vector<StateKey> seen; // a handful of siblings
  canonical_hash();           // so that every child inherits the digest of the common prefix
  for (const auto &d : get_decisions(plat, quiet)) {
    State s = apply(*d);
    const StateKey k = s.canonical_hash();
    if (std::find(seen.begin(), seen.end(), k) == seen.end()) {
      seen.push_back(k);
      result.push_back(std::move(s));
    }
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

// ---- the digest form of the same key ----
// Token for token what canonical_key() writes (sequence ops, '#', then the graph's vertices in
// unbound-uid order with their sorted successor lists, one renumbering across both), folded into
// 128 bits instead of a string.  Two things make it cheap: a state derived by ExecuteOp digests
// only the op it added, and the graph's tokens are laid out once per graph object.
void detail::KeyDigest::op(const OpBase &o) {
  const uint16_t t = o.traits();
  if (t & op_trait::Sync)
    token('k', intern_key(o.kind()));
  else
    token('o', o.unbound_uid());
  if (t & op_trait::BoundGpu) {
    stream(static_cast<const BoundGpuOp &>(o).stream().id());
    return;
  }
  if (t & op_trait::WithStream)
    for (const Stream &st : dynamic_cast<const HasStream &>(o).get_streams()) stream(st.id());
  if (t & op_trait::WithEvent)
    for (const Event &e : dynamic_cast<const HasEvent &>(o).get_events()) event(e.id());
}

static std::shared_ptr<const detail::GraphTokens> make_graph_tokens(const Graph<OpBase> &g) {
  auto gt = std::make_shared<detail::GraphTokens>();
  auto put = [&gt](char tag, uint32_t v) { gt->tokens.push_back((uint64_t(uint8_t(tag)) << 32) | v); };
  std::vector<OpPtr> verts = g.vertices();
  std::sort(verts.begin(), verts.end(),
            [](const OpPtr &x, const OpPtr &y) { return x->unbound_uid() < y->unbound_uid(); });
  std::vector<uint32_t> succIds;
  for (const auto &v : verts) {
    if (const auto *sync = dynamic_cast<const SyncOpBase *>(v.get()))
      put('k', intern_key(sync->kind()));
    else
      put('o', v->unbound_uid());
    if (const auto *hs = dynamic_cast<const HasStream *>(v.get()))
      for (const Stream &st : hs->get_streams()) {
        put('s', uint32_t(st.id()));
        gt->hasResources = true;
      }
    if (const auto *he = dynamic_cast<const HasEvent *>(v.get()))
      for (const Event &e : he->get_events()) {
        put('e', uint32_t(e.id()));
        gt->hasResources = true;
      }
    put('>', 0);
    succIds.clear();
    for (const auto &sc : g.succs(v)) succIds.push_back(sc->unbound_uid());
    std::sort(succIds.begin(), succIds.end());
    for (uint32_t id : succIds) put(',', id);
    put('|', 0);
  }
  if (!gt->hasResources) {
    detail::KeyDigest d;
    for (uint64_t t : gt->tokens) d.mix(t);
    gt->plain = StateKey{d.a, d.b};
  }
  return gt;
}

StateKey State::sequence_hash() const {
  for (; seqDigested_ < sequence_.size(); ++seqDigested_) seqDigest_.op(*sequence_[seqDigested_]);
  return StateKey{seqDigest_.a, seqDigest_.b};
}

StateKey State::canonical_hash() const {
  sequence_hash();
  if (!graphTokens_) graphTokens_ = make_graph_tokens(*graph_);
  detail::KeyDigest d;
  d.a = seqDigest_.a;
  d.b = seqDigest_.b;
  d.mix('#');
  if (!graphTokens_->hasResources) {
    d.mix(graphTokens_->plain.a);
    d.mix(graphTokens_->plain.b);
  } else {
    // the graph's streams and events continue the sequence's renumbering
    d.streams = seqDigest_.streams;
    d.events = seqDigest_.events;
    for (uint64_t t : graphTokens_->tokens) {
      const char tag = char(t >> 32);
      if (tag == 's') d.stream(int(uint32_t(t)));
      else if (tag == 'e') d.event(int(uint32_t(t)));
      else d.mix(t);
    }
  }
  return StateKey{d.a, d.b};
}

StateKey sequence_hash(const Sequence<BoundOp> &seq) {
  detail::KeyDigest d;
  for (const auto &op : seq) d.op(*op);
  return StateKey{d.a, d.b};
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
