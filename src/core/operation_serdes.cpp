#include "tenzing/operation_serdes.hpp"

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

void from_json(const Json &j, const Graph<OpBase> &g, BoundOpPtr &op) {
  const std::string name = j.contains("name") ? j.at("name").as_string() : std::string();
  // sync ops are self-describing and never live in a graph
  if (auto s = sync_op_from_json(j)) {
    op = s;
    return;
  }
  OpPtr found = name.empty() ? nullptr : find_op_by_name(g, name);
  if (found) {
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
  if (j.contains("kind") && j.at("kind").is_string() && j.at("kind").as_string() == "NoOp") {
    op = std::make_shared<NoOp>(name);
    return;
  }
  THROW_RUNTIME("cannot resolve op from " << j.dump());
}

void from_json(const Json &j, const Graph<OpBase> &g, Sequence<BoundOp> &seq) {
  seq.clear();
  for (const Json &e : j.as_array()) {
    BoundOpPtr op;
    from_json(e, g, op);
    seq.push_back(op);
  }
}

} // namespace tenzing
