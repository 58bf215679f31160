#include "tenzing/sequence.hpp"

#include <map>

#include "tenzing/communicator.hpp"
#include "tenzing/operation_serdes.hpp"

namespace tenzing {

std::string get_desc_delim(const Sequence<BoundOp> &seq, const std::string &delim) {
  return seq.desc(delim);
}

// streams/events of two ops must pair up one-to-one under `eq`
static bool resources_consistent(const OpBase &a, const OpBase &b, Equivalence &eq) {
  const auto *sa = dynamic_cast<const HasStream *>(&a);
  const auto *sb = dynamic_cast<const HasStream *>(&b);
  if (bool(sa) != bool(sb)) return false;
  if (sa) {
    const auto va = sa->get_streams(), vb = sb->get_streams();
    if (va.size() != vb.size()) return false;
    for (size_t i = 0; i < va.size(); ++i)
      if (!eq.check_or_insert(va[i], vb[i])) return false;
  }
  const auto *ea = dynamic_cast<const HasEvent *>(&a);
  const auto *eb = dynamic_cast<const HasEvent *>(&b);
  if (bool(ea) != bool(eb)) return false;
  if (ea) {
    const auto va = ea->get_events(), vb = eb->get_events();
    if (va.size() != vb.size()) return false;
    for (size_t i = 0; i < va.size(); ++i)
      if (!eq.check_or_insert(va[i], vb[i])) return false;
  }
  return true;
}

// sync ops are compared by kind (their names carry search context), other ops by name
static bool same_logical_op(const OpBase &a, const OpBase &b) {
  const bool sa = dynamic_cast<const SyncOpBase *>(&a) != nullptr;
  const bool sb = dynamic_cast<const SyncOpBase *>(&b) != nullptr;
  if (sa != sb) return false;
  if (sa) return a.kind() == b.kind();
  return a.name() == b.name();
}

Equivalence get_equivalence(const Sequence<BoundOp> &a, const Sequence<BoundOp> &b) {
  if (a.size() != b.size()) return Equivalence::falsy();
  Equivalence eq;
  for (size_t i = 0; i < a.size(); ++i) {
    if (!same_logical_op(*a[i], *b[i])) return Equivalence::falsy();
    if (!resources_consistent(*a[i], *b[i], eq)) return Equivalence::falsy();
  }
  return eq;
}

std::string canonical_key(const Sequence<BoundOp> &seq) {
  std::map<int, int> sMap, eMap;
  auto relabel = [](std::map<int, int> &m, int id) {
    auto it = m.find(id);
    if (it != m.end()) return it->second;
    const int n = static_cast<int>(m.size());
    m.emplace(id, n);
    return n;
  };
  std::string key;
  for (const auto &op : seq) {
    const bool sync = dynamic_cast<const SyncOpBase *>(op.get()) != nullptr;
    key += sync ? op->kind() : op->name();
    if (const auto *hs = dynamic_cast<const HasStream *>(op.get()))
      for (const Stream &s : hs->get_streams()) key += "/s" + std::to_string(relabel(sMap, s.id()));
    if (const auto *he = dynamic_cast<const HasEvent *>(op.get()))
      for (const Event &e : he->get_events()) key += "/e" + std::to_string(relabel(eMap, e.id()));
    key += "|";
  }
  return key;
}

Sequence<BoundOp> bcast(const Sequence<BoundOp> &seq, const Graph<OpBase> &g, Communicator &comm,
                        int root) {
  if (comm.size() == 1) return seq;
  std::string text;
  if (comm.rank() == root) text = to_json(seq, g).dump();
  text = comm.bcast_string(text, root);
  if (comm.rank() == root) return seq;
  Sequence<BoundOp> out;
  from_json(Json::parse(text), g, out);
  return out;
}

} // namespace tenzing
