// An ordered list of executable ops: a (partial) implementation of the program.
//
// Reference: include/tenzing/sequence.hpp:21-143, src/sequence.cpp:13-167.
#pragma once

#include <memory>
#include <string>
#include <vector>

#include "tenzing/cuda/ops_cuda.hpp"
#include "tenzing/graph.hpp"
#include "tenzing/operation.hpp"

namespace tenzing {

class Communicator;

template <typename T> class Sequence {
public:
  using value_type = std::shared_ptr<T>;
  using vector_t = std::vector<value_type>;
  using iterator = typename vector_t::iterator;
  using const_iterator = typename vector_t::const_iterator;

private:
  vector_t ops_;

public:
  Sequence() = default;
  Sequence(std::initializer_list<value_type> il) : ops_(il) {}
  explicit Sequence(const vector_t &v) : ops_(v) {}

  iterator begin() { return ops_.begin(); }
  iterator end() { return ops_.end(); }
  const_iterator begin() const { return ops_.begin(); }
  const_iterator end() const { return ops_.end(); }
  size_t size() const { return ops_.size(); }
  bool empty() const { return ops_.empty(); }
  void clear() { ops_.clear(); }
  void push_back(const value_type &op) { ops_.push_back(op); }
  void push_back(value_type &&op) { ops_.push_back(std::move(op)); }
  value_type &operator[](size_t i) { return ops_[i]; }
  const value_type &operator[](size_t i) const { return ops_[i]; }
  const value_type &back() const { return ops_.back(); }
  iterator erase(const_iterator it) { return ops_.erase(it); }
  iterator insert(const_iterator it, const value_type &op) { return ops_.insert(it, op); }
  vector_t &vector() { return ops_; }
  const vector_t &vector() const { return ops_; }

  bool contains(const OpBase &op) const {
    for (const auto &e : ops_)
      if (e->uid() == op.uid()) return true;
    return false;
  }
  // first element whose unbound identity equals op's unbound identity
  const_iterator find_unbound(const OpBase &op) const {
    const uint32_t u = op.unbound_uid();
    for (auto it = ops_.begin(); it != ops_.end(); ++it)
      if ((*it)->unbound_uid() == u) return it;
    return ops_.end();
  }
  const_iterator find_unbound(const std::shared_ptr<const OpBase> &op) const {
    return find_unbound(*op);
  }
  bool contains_unbound(const OpBase &op) const { return find_unbound(op) != ops_.end(); }
  bool contains_unbound(const std::shared_ptr<const OpBase> &op) const {
    return contains_unbound(*op);
  }

  // smallest event id no op of the sequence uses
  Event new_unique_event() const {
    std::vector<char> used;
    for (const auto &e : ops_) {
      if (auto he = std::dynamic_pointer_cast<HasEvent>(e)) {
        for (const Event &ev : he->get_events()) {
          if (ev.id() < 0) continue;
          if (size_t(ev.id()) >= used.size()) used.resize(ev.id() + 1, 0);
          used[ev.id()] = 1;
        }
      }
    }
    int id = 0;
    while (size_t(id) < used.size() && used[id]) ++id;
    return Event(id);
  }

  std::string desc(const std::string &delim = ", ") const {
    std::string s;
    for (size_t i = 0; i < ops_.size(); ++i) {
      if (i) s += delim;
      s += ops_[i]->desc();
    }
    return s;
  }
};

std::string get_desc_delim(const Sequence<BoundOp> &seq, const std::string &delim);

// position-wise: same names, streams and events consistent under one bijection
Equivalence get_equivalence(const Sequence<BoundOp> &a, const Sequence<BoundOp> &b);

// canonical text of a sequence: op names with streams/events renumbered in
// first-use order.  Two sequences are equivalent iff their keys are equal, so
// de-duplication is a hash lookup (the reference compares all pairs).
std::string canonical_key(const Sequence<BoundOp> &seq);

// rank `root` serialises the sequence to JSON and broadcasts it; the other
// ranks re-resolve every op by name in their own graph (reference:
// src/sequence.cpp:88-125 `mpi_bcast`).  Returns the sequence valid on this rank.
Sequence<BoundOp> bcast(const Sequence<BoundOp> &seq, const Graph<OpBase> &g, Communicator &comm,
                        int root = 0);
inline Sequence<BoundOp> mpi_bcast(const Sequence<BoundOp> &seq, const Graph<OpBase> &g,
                                   Communicator &comm) {
  return bcast(seq, g, comm, 0);
}

} // namespace tenzing
