// The moves of the sequential decision process.
// Reference: include/tenzing/decision.hpp:13-63.
#pragma once

#include <memory>
#include <string>

#include "tenzing/cuda/ops_cuda.hpp"
#include "tenzing/operation.hpp"

namespace tenzing {

class Decision {
public:
  enum class Kind { ExecuteOp, ExpandOp, ChooseOp, AssignOpStream };
  virtual ~Decision() = default;
  virtual Kind kind() const = 0;
  virtual std::string desc() const = 0;
};
using DecisionPtr = std::shared_ptr<Decision>;

// append an executable op (a program op or a synchronisation op) to the sequence
class ExecuteOp final : public Decision {
public:
  BoundOpPtr op;
  explicit ExecuteOp(const BoundOpPtr &o) : op(o) {}
  Kind kind() const override { return Kind::ExecuteOp; }
  std::string desc() const override { return "Execute " + op->desc(); }
};

// replace a compound op by its sub-graph
class ExpandOp final : public Decision {
public:
  std::shared_ptr<CompoundOp> op;
  explicit ExpandOp(const std::shared_ptr<CompoundOp> &o) : op(o) {}
  Kind kind() const override { return Kind::ExpandOp; }
  std::string desc() const override { return "Expand " + op->desc(); }
};

// replace a choice op by one of its alternatives
class ChooseOp final : public Decision {
public:
  std::shared_ptr<ChoiceOp> orig;
  OpPtr replacement;
  ChooseOp(const std::shared_ptr<ChoiceOp> &o, const OpPtr &r) : orig(o), replacement(r) {}
  Kind kind() const override { return Kind::ChooseOp; }
  std::string desc() const override {
    return "Choose " + replacement->desc() + " for " + orig->desc();
  }
};

// bind a GPU op to a stream
class AssignOpStream final : public Decision {
public:
  GpuOpPtr op;
  Stream stream;
  AssignOpStream(const GpuOpPtr &o, Stream s) : op(o), stream(s) {}
  Kind kind() const override { return Kind::AssignOpStream; }
  std::string desc() const override {
    return "Assign " + op->desc() + " to stream " + stream.str();
  }
};

} // namespace tenzing
