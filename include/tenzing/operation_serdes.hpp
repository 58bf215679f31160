// Op / sequence (de)serialisation against a graph.
// Reference: include/tenzing/operation_serdes.hpp:14-46, src/operation_serdes.cpp:14-76;
// format in SURVEY App. A.  Deserialisation resolves ops BY NAME in the rank's
// own graph (recursing into compound sub-graphs and choice alternatives),
// binds GpuOps to j["stream"], and falls back to constructing a sync op from
// j["kind"].
#pragma once

#include "tenzing/graph.hpp"
#include "tenzing/json.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {

void to_json(Json &j, const std::shared_ptr<const OpBase> &op, const Graph<OpBase> &g);
void to_json(Json &j, const Sequence<BoundOp> &seq, const Graph<OpBase> &g);
Json to_json(const Sequence<BoundOp> &seq, const Graph<OpBase> &g);
// plain op JSON (no "in_graph") for every op of the sequence - the CSV tail columns
std::vector<Json> to_json_ops(const Sequence<BoundOp> &seq);

void from_json(const Json &j, const Graph<OpBase> &g, BoundOpPtr &op);
void from_json(const Json &j, const Graph<OpBase> &g, Sequence<BoundOp> &seq);

// find an op named `name` anywhere in g (compound graphs / choice alternatives included)
OpPtr find_op_by_name(const Graph<OpBase> &g, const std::string &name);

} // namespace tenzing
