#include "tenzing/flag_synchronizer.hpp"

#include <unordered_set>

namespace tenzing {

namespace {
const HasFlags *flags_of(const OpBase &op) {
  if (const auto *b = dynamic_cast<const BoundGpuOp *>(&op))
    return dynamic_cast<const HasFlags *>(b->unbound().get());
  return dynamic_cast<const HasFlags *>(&op);
}
} // namespace

std::vector<FlagSynchronizer::Violation> FlagSynchronizer::check(const Sequence<BoundOp> &seq) {
  std::vector<Violation> out;
  std::unordered_set<uint64_t> published;
  for (size_t i = 0; i < seq.size(); ++i) {
    const HasFlags *f = flags_of(*seq[i]);
    if (!f) continue;
    if (f->publishes_before_waiting())
      for (uint64_t c : f->flags_published()) published.insert(c);
    for (uint64_t c : f->flags_awaited())
      if (!published.count(c)) out.push_back(Violation{i, seq[i]->name(), c});
    for (uint64_t c : f->flags_published()) published.insert(c);
  }
  return out;
}

bool FlagSynchronizer::is_ready(const BoundOp &op, const Sequence<BoundOp> &path) {
  const HasFlags *f = flags_of(op);
  if (!f) return true;
  std::unordered_set<uint64_t> published;
  for (const auto &p : path)
    if (const HasFlags *pf = flags_of(*p))
      for (uint64_t c : pf->flags_published()) published.insert(c);
  if (f->publishes_before_waiting())
    for (uint64_t c : f->flags_published()) published.insert(c);
  for (uint64_t c : f->flags_awaited())
    if (!published.count(c)) return false;
  return true;
}

int FlagSynchronizer::add_edges(Graph<OpBase> &g) {
  int added = 0;
  const auto verts = g.vertices();
  for (const auto &w : verts) {
    const HasFlags *wf = flags_of(*w);
    if (!wf) continue;
    const auto awaited = wf->flags_awaited();
    if (awaited.empty()) continue;
    const std::unordered_set<uint64_t> need(awaited.begin(), awaited.end());
    for (const auto &p : verts) {
      if (p->uid() == w->uid()) continue;
      const HasFlags *pf = flags_of(*p);
      if (!pf) continue;
      bool shares = false;
      for (uint64_t c : pf->flags_published())
        if (need.count(c)) shares = true;
      if (!shares) continue;
      bool have = false;
      for (const auto &s : g.succs(p))
        if (s->uid() == w->uid()) have = true;
      if (!have) {
        g.add_edge(p, w);
        ++added;
      }
    }
  }
  return added;
}

} // namespace tenzing
