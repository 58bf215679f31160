// Concrete compound / choice ops built from values (reference declares only the
// abstract CompoundOp, include/tenzing/operation_compound.hpp:8-12; concrete
// ones such as SpMV live with the workloads).
#pragma once

#include "tenzing/graph.hpp"
#include "tenzing/operation.hpp"

namespace tenzing {

// a compound op that simply carries its sub-graph
class GraphCompoundOp : public CompoundOp {
  std::string name_;
  Graph<OpBase> graph_;

public:
  GraphCompoundOp(const std::string &name, const Graph<OpBase> &g) : name_(name), graph_(g) {}
  std::string name() const override { return name_; }
  Graph<OpBase> graph() const override { return graph_; }
  CLONE_DEF(GraphCompoundOp)
};

// a choice between explicitly listed alternatives
class ListChoiceOp : public ChoiceOp {
  std::string name_;
  std::vector<OpPtr> choices_;

public:
  ListChoiceOp(const std::string &name, const std::vector<OpPtr> &choices)
      : name_(name), choices_(choices) {}
  std::string name() const override { return name_; }
  std::vector<OpPtr> choices() const override { return choices_; }
  CLONE_DEF(ListChoiceOp)
};

} // namespace tenzing
