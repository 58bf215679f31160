#include "tenzing/graph.hpp"

#include <map>

namespace tenzing {

namespace {
using G = Graph<OpBase>;

std::vector<GpuOpPtr> unbound_gpu_ops(const G &g) { return g.nodes_cast<GpuOp>(); }

void assign_rec(const G &g, const std::vector<GpuOpPtr> &ops, size_t i, int maxStream, int opened,
                bool canonical, const std::vector<Stream> &streams, std::vector<G> &out) {
  if (i == ops.size()) {
    out.push_back(g);
    return;
  }
  // canonical mode: op i may reuse any stream opened so far or open exactly the next one
  const int limit = canonical ? std::min<int>(opened + 1, int(streams.size())) : int(streams.size());
  (void)maxStream;
  for (int s = 0; s < limit; ++s) {
    G next = g.clone_but_replace(std::make_shared<BoundGpuOp>(ops[i], streams[s]), ops[i]);
    assign_rec(next, ops, i + 1, maxStream, std::max(opened, s + 1), canonical, streams, out);
  }
}
} // namespace

std::vector<G> use_streams(const G &g, const std::vector<Stream> &streams) {
  std::vector<G> out;
  assign_rec(g, unbound_gpu_ops(g), 0, 0, 0, false, streams, out);
  return out;
}

std::vector<G> use_streams2(const G &g, const std::vector<Stream> &streams) {
  std::vector<G> out;
  assign_rec(g, unbound_gpu_ops(g), 0, 0, 0, true, streams, out);
  return out;
}

G apply_assignment(const G &g, const std::vector<std::pair<std::string, Stream>> &assignment) {
  G out = g;
  for (const auto &kv : assignment) {
    OpPtr op = out.find_name(kv.first);
    if (!op) THROW_RUNTIME("apply_assignment: no op named " << kv.first);
    GpuOpPtr gop = std::dynamic_pointer_cast<GpuOp>(op);
    if (!gop) {
      if (auto b = std::dynamic_pointer_cast<BoundGpuOp>(op)) gop = b->unbound();
    }
    if (!gop) THROW_RUNTIME("apply_assignment: " << kv.first << " is not a GPU op");
    out.replace(op, std::make_shared<BoundGpuOp>(gop, kv.second));
  }
  return out;
}

Equivalence get_equivalence(const G &a, const G &b) {
  if (a.vertex_size() != b.vertex_size()) return Equivalence::falsy();
  std::map<std::string, OpPtr> byName;
  for (const auto &v : b.vertices()) byName[v->name()] = v;
  Equivalence eq;
  for (const auto &va : a.vertices()) {
    auto it = byName.find(va->name());
    if (it == byName.end()) return Equivalence::falsy();
    const OpPtr &vb = it->second;
    const auto *sa = dynamic_cast<const HasStream *>(va.get());
    const auto *sb = dynamic_cast<const HasStream *>(vb.get());
    if (bool(sa) != bool(sb)) return Equivalence::falsy();
    if (sa) {
      const auto xa = sa->get_streams(), xb = sb->get_streams();
      if (xa.size() != xb.size()) return Equivalence::falsy();
      for (size_t i = 0; i < xa.size(); ++i)
        if (!eq.check_or_insert(xa[i], xb[i])) return Equivalence::falsy();
    }
    const auto *ea = dynamic_cast<const HasEvent *>(va.get());
    const auto *eb = dynamic_cast<const HasEvent *>(vb.get());
    if (bool(ea) != bool(eb)) return Equivalence::falsy();
    if (ea) {
      const auto xa = ea->get_events(), xb = eb->get_events();
      if (xa.size() != xb.size()) return Equivalence::falsy();
      for (size_t i = 0; i < xa.size(); ++i)
        if (!eq.check_or_insert(xa[i], xb[i])) return Equivalence::falsy();
    }
    // same successor names
    std::vector<std::string> na, nb;
    for (const auto &s : a.succs(va)) na.push_back(s->name());
    for (const auto &s : b.succs(vb)) nb.push_back(s->name());
    std::sort(na.begin(), na.end());
    std::sort(nb.begin(), nb.end());
    if (na != nb) return Equivalence::falsy();
  }
  return eq;
}

bool is_equivalent_stream_mapping(const G &a, const G &b) { return bool(get_equivalence(a, b)); }

G insert_synchronization(const G &g) {
  G out = g;
  int nextEvent = 0;
  for (const auto &he : g.nodes_cast<HasEvent>())
    for (const Event &e : he->get_events()) nextEvent = std::max(nextEvent, e.id() + 1);
  for (const auto &u : g.vertices()) {
    auto ug = std::dynamic_pointer_cast<BoundGpuOp>(u);
    if (!ug) continue;
    for (const auto &v : g.succs(u)) {
      auto vg = std::dynamic_pointer_cast<BoundGpuOp>(v);
      const bool vCpu = bool(std::dynamic_pointer_cast<CpuOp>(v));
      if (vg && vg->stream() == ug->stream()) continue;
      if (!vg && !vCpu) continue; // unbound successor: cannot decide yet
      Event e(nextEvent++);
      auto cer = std::make_shared<CudaEventRecord>(e, ug->stream());
      cer->update_name({u}, {v});
      std::shared_ptr<OpBase> wait;
      if (vg) {
        auto w = std::make_shared<CudaStreamWaitEvent>(vg->stream(), e);
        w->update_name({u}, {v});
        wait = w;
      } else {
        auto w = std::make_shared<CudaEventSync>(e);
        w->update_name({u}, {v});
        wait = w;
      }
      out.erase_edge_only(u, v);
      out.add_edge(u, std::static_pointer_cast<OpBase>(cer));
      out.add_edge(std::static_pointer_cast<OpBase>(cer), wait);
      out.add_edge(wait, v);
    }
  }
  return out;
}

} // namespace tenzing
