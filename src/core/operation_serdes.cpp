#include "tenzing/operation_serdes.hpp"

#include <unordered_map>
#include <unordered_set>

namespace tenzing {

void to_json(Json &j, const std::shared_ptr<const OpBase> &op, const Graph<OpBase> &g) {
  j = op->json();
  j["in_graph"] = g.find_vertex_unbound(*op) >= 0;
}

void to_json(Json &j, const Sequence<BoundOp> &seq, const Graph<OpBase> &g) {
  j = Json::array();
  for (const auto &op : seq) {
    Json e;
    to_json(e, op, g);
    j.push_back(e);
  }
}

Json to_json(const Sequence<BoundOp> &seq, const Graph<OpBase> &g) {
  Json j;
  to_json(j, seq, g);
  return j;
}

std::vector<Json> to_json_ops(const Sequence<BoundOp> &seq) {
  std::vector<Json> r;
  for (const auto &op : seq) r.push_back(op->json());
  return r;
}

static OpPtr find_in_op(const OpPtr &op, const std::string &name);

static OpPtr find_in_graph(const Graph<OpBase> &g, const std::string &name) {
  for (const auto &v : g.vertices())
    if (auto r = find_in_op(v, name)) return r;
  return nullptr;
}

static OpPtr find_in_op(const OpPtr &op, const std::string &name) {
  if (op->name() == name) return op;
  if (auto c = std::dynamic_pointer_cast<CompoundOp>(op)) {
    return find_in_graph(c->graph(), name);
  }
  if (auto c = std::dynamic_pointer_cast<ChoiceOp>(op)) {
    for (const auto &alt : c->choices())
      if (auto r = find_in_op(alt, name)) return r;
  }
  return nullptr;
}

OpPtr find_op_by_name(const Graph<OpBase> &g, const std::string &name) {
  return find_in_graph(g, name);
}

namespace {

using NameMap = std::unordered_map<std::string, OpPtr>;

// Joint resolution of a whole sequence.  Alternatives of a ChoiceOp may contain ops with the
// same names (every SpMV style has a "yl" and a "yr", and they are different ops), so a name
// alone does not say which op is meant; the SET of names does: a complete sequence contains
// exactly the ops of the alternatives that were chosen.  For every ChoiceOp the alternative
// is taken whose ops are all named by the sequence (the largest such one; failing that, the
// one with the most names in the sequence), and names are then resolved inside it.
struct Match {
  size_t matched = 0, total = 0;
  bool complete() const { return matched == total; }
  bool better_than(const Match &o) const {
    if (complete() != o.complete()) return complete();
    return matched > o.matched;
  }
};

Match collect(const OpPtr &op, const std::unordered_set<std::string> &names, NameMap &out);

Match collect_graph(const Graph<OpBase> &g, const std::unordered_set<std::string> &names, NameMap &out) {
  Match m;
  for (const auto &v : g.vertices()) {
    if (v == g.start() || v == g.finish()) continue;
    const Match c = collect(v, names, out);
    m.matched += c.matched;
    m.total += c.total;
  }
  return m;
}

Match collect(const OpPtr &op, const std::unordered_set<std::string> &names, NameMap &out) {
  if (auto c = std::dynamic_pointer_cast<CompoundOp>(op)) return collect_graph(c->graph(), names, out);
  if (auto c = std::dynamic_pointer_cast<ChoiceOp>(op)) {
    Match best;
    NameMap bestMap;
    bool have = false;
    for (const auto &alt : c->choices()) {
      NameMap tmp;
      const Match m = collect(alt, names, tmp);
      if (!have || m.better_than(best)) {
        best = m;
        bestMap = std::move(tmp);
        have = true;
      }
    }
    for (auto &kv : bestMap) out.emplace(kv.first, kv.second);
    return best;
  }
  Match m;
  m.total = 1;
  if (names.count(op->name())) {
    m.matched = 1;
    out.emplace(op->name(), op); // first one wins, as find_op_by_name
  }
  return m;
}

void bind_found(const Json &j, const std::string &name, const OpPtr &found, BoundOpPtr &op) {
  if (auto b = std::dynamic_pointer_cast<BoundGpuOp>(found)) {
    // graph already holds a binding; honour the stream recorded in the JSON
    if (j.contains("stream") && int(j.at("stream").as_int()) != b->stream().id())
      op = std::make_shared<BoundGpuOp>(b->unbound(), Stream(int(j.at("stream").as_int())));
    else
      op = b;
    return;
  }
  if (auto b = std::dynamic_pointer_cast<BoundOp>(found)) {
    op = b;
    return;
  }
  if (auto gop = std::dynamic_pointer_cast<GpuOp>(found)) {
    if (!j.contains("stream")) THROW_RUNTIME("GPU op " << name << " has no stream in " << j.dump());
    op = std::make_shared<BoundGpuOp>(gop, Stream(int(j.at("stream").as_int())));
    return;
  }
  THROW_RUNTIME("op " << name << " is in the graph but is not executable");
}

void resolve(const Json &j, const Graph<OpBase> &g, const NameMap *names, BoundOpPtr &op) {
  const std::string name = j.contains("name") ? j.at("name").as_string() : std::string();
  // sync ops are self-describing and never live in a graph
  if (auto s = sync_op_from_json(j)) {
    op = s;
    return;
  }
  OpPtr found;
  if (names && !name.empty()) {
    auto it = names->find(name);
    if (it != names->end()) found = it->second;
  }
  if (!found && !name.empty()) found = find_op_by_name(g, name);
  if (found) return bind_found(j, name, found, op);
  if (j.contains("kind") && j.at("kind").is_string() && j.at("kind").as_string() == "NoOp") {
    op = std::make_shared<NoOp>(name);
    return;
  }
  THROW_RUNTIME("cannot resolve op from " << j.dump());
}

} // namespace

// a single op: first op of that name anywhere in the graph (as the reference does); ambiguous
// when alternatives of a ChoiceOp share names - resolve whole sequences where that matters
void from_json(const Json &j, const Graph<OpBase> &g, BoundOpPtr &op) { resolve(j, g, nullptr, op); }

void from_json(const Json &j, const Graph<OpBase> &g, Sequence<BoundOp> &seq) {
  seq.clear();
  std::unordered_set<std::string> wanted;
  for (const Json &e : j.as_array())
    if (e.is_object() && e.contains("name") && e.at("name").is_string())
      wanted.insert(e.at("name").as_string());
  NameMap names;
  collect_graph(g, wanted, names);
  for (const Json &e : j.as_array()) {
    BoundOpPtr op;
    resolve(e, g, &names, op);
    seq.push_back(op);
  }
}

} // namespace tenzing
