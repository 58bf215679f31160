// Operation hierarchy of the sequential-decision-process model.
//
// Reference: include/tenzing/operation.hpp:35-173, operation_compound.hpp:8-12.
// Same public surface (OpBase{name,desc,json,clone,eq,lt,tag}, BoundOp::run,
// CpuOp, ChoiceOp::choices, CompoundOp::graph, Start, Finish, NoOp), different
// mechanics:
//  * identity is by value through one string `key()` ("<kind>:<name>", plus
//    stream/event fields where they matter).  Keys are interned to a 32-bit
//    `uid()` so graphs and states index ops with integers instead of walking
//    std::map<shared_ptr, ...> with dynamic_casts on every comparison;
//  * ops are treated as immutable once they are in a graph, so graph rewrites
//    share them instead of deep-cloning (the reference clones every op - and
//    a cuSPARSE handle per SpMV op - on every decision);
//  * `kind()` is a stable string (also written to JSON for sync ops), `tag()`
//    is its hash.
// As in the reference, two ops with the same key ARE the same op: give every
// op of a program a unique name.
#pragma once

#include <cstdint>
#include <functional>
#include <atomic>
#include <memory>
#include <set>
#include <string>
#include <vector>

#include "tenzing/json.hpp"
#include "tenzing/macro_at.hpp"
#include "tenzing/platform.hpp"

namespace tenzing {

class OpBase;
using OpPtr = std::shared_ptr<OpBase>;

// process-wide key interning
uint32_t intern_key(const std::string &key);
const std::string &interned_key(uint32_t uid);

#define CLONE_DEF(TYPE)                                                                            \
  std::shared_ptr<::tenzing::OpBase> clone() const override {                                      \
    return std::make_shared<TYPE>(*this);                                                          \
  }
// Kept for source compatibility with ops written against the reference: eq/lt
// are derived from key() here, so these expand to nothing.
#define EQ_DEF(TYPE)
#define LT_DEF(TYPE)

// What an op is, as bits.  The search asks these questions of every op of every state it
// visits; the answers depend on the dynamic type only, so each object works them out once
// (dynamic_casts, src/core/operation.cpp) and keeps them.
namespace op_trait {
enum : uint16_t {
  Bound = 1u << 0,           // BoundOp
  Cpu = 1u << 1,             // CpuOp
  BoundGpu = 1u << 2,        // exactly BoundGpuOp
  Gpu = 1u << 3,             // GpuOp (unbound)
  Compound = 1u << 4,        // CompoundOp
  Choice = 1u << 5,          // ChoiceOp
  Sync = 1u << 6,            // SyncOpBase
  WithStream = 1u << 7,      // HasStream
  WithEvent = 1u << 8,       // HasEvent
  EventRecord = 1u << 9,     // exactly CudaEventRecord
  EventSync = 1u << 10,      // exactly CudaEventSync
  StreamWaitEvent = 1u << 11, // exactly CudaStreamWaitEvent
  IsStart = 1u << 12,        // exactly Start
  Known = 1u << 15
};
} // namespace op_trait

class OpBase {
  // Caches filled on first use.  Relaxed atomics: two threads that race to fill one compute the
  // same value, and ops are shared between the threads of a multi-rank-in-one-process run.
  mutable std::atomic<uint32_t> uid_{0}; // 0 = not interned yet
  mutable std::atomic<uint32_t> unboundUid_{0};
  mutable std::atomic<uint16_t> traits_{0}; // 0 = not worked out yet
  uint16_t compute_traits() const;

public:
  OpBase() = default;
  // copies must re-intern (a derived copy-constructor may change key fields)
  OpBase(const OpBase &) {}
  OpBase &operator=(const OpBase &) {
    invalidate_identity();
    traits_.store(0, std::memory_order_relaxed);
    return *this;
  }
  virtual ~OpBase() = default;

  // op_trait bits of this object's dynamic type
  uint16_t traits() const {
    uint16_t t = traits_.load(std::memory_order_relaxed);
    if (!t) {
      t = compute_traits();
      traits_.store(t, std::memory_order_relaxed);
    }
    return t;
  }
  bool is(uint16_t trait) const { return (traits() & trait) != 0; }

  virtual std::string name() const = 0;
  virtual std::string desc() const { return name(); }
  virtual std::string kind() const { return "Op"; }
  virtual Json json() const; // {"name": name()}
  virtual std::shared_ptr<OpBase> clone() const = 0;

  // value identity
  virtual std::string key() const { return kind() + ":" + name(); }
  // identity with any stream binding removed (BoundGpuOp -> the GpuOp's key)
  virtual std::string unbound_key() const { return key(); }

  uint32_t uid() const {
    uint32_t u = uid_.load(std::memory_order_relaxed);
    if (!u) {
      u = intern_self();
      uid_.store(u, std::memory_order_relaxed);
    }
    return u;
  }
  uint32_t unbound_uid() const {
    uint32_t u = unboundUid_.load(std::memory_order_relaxed);
    if (!u) {
      u = intern_unbound_self();
      unboundUid_.store(u, std::memory_order_relaxed);
    }
    return u;
  }
  bool eq(const OpBase &rhs) const { return uid() == rhs.uid(); }
  bool eq(const std::shared_ptr<const OpBase> &rhs) const { return rhs && uid() == rhs->uid(); }
  bool lt(const OpBase &rhs) const { return key() < rhs.key(); }
  bool lt(const std::shared_ptr<const OpBase> &rhs) const { return key() < rhs->key(); }
  int tag() const { return static_cast<int>(std::hash<std::string>{}(kind())); }

  struct compare_lt {
    bool operator()(const std::shared_ptr<const OpBase> &a,
                    const std::shared_ptr<const OpBase> &b) const {
      return a->lt(*b);
    }
  };

protected:
  // call after changing a field that key() depends on
  void invalidate_identity() {
    uid_.store(0, std::memory_order_relaxed);
    unboundUid_.store(0, std::memory_order_relaxed);
  }
  // how uid() / unbound_uid() get their number; ops made by the thousand during a search
  // override these with something cheaper than building and hashing the key string, and must
  // return exactly intern_key(key()) / intern_key(unbound_key())
  virtual uint32_t intern_self() const { return intern_key(key()); }
  virtual uint32_t intern_unbound_self() const { return intern_key(unbound_key()); }
};

// not executable: several implementation alternatives for one logical operation
class ChoiceOp : public OpBase {
public:
  std::string kind() const override { return "ChoiceOp"; }
  virtual std::vector<OpPtr> choices() const = 0;
};

// executable on the platform as is
class BoundOp : public OpBase {
public:
  virtual void run(Platform &plat) = 0;
};
using BoundOpPtr = std::shared_ptr<BoundOp>;

// executes on the host thread
class CpuOp : public BoundOp {
public:
  std::string kind() const override { return "CpuOp"; }
};

class Start final : public CpuOp {
public:
  std::string name() const override { return "Start"; }
  std::string kind() const override { return "Start"; }
  std::string key() const override { return "Start"; }
  CLONE_DEF(Start)
  void run(Platform &) override {}
};

class Finish final : public CpuOp {
public:
  std::string name() const override { return "Finish"; }
  std::string kind() const override { return "Finish"; }
  std::string key() const override { return "Finish"; }
  CLONE_DEF(Finish)
  void run(Platform &) override {}
};

// does nothing; identity is its name
class NoOp : public CpuOp {
  std::string name_;

public:
  explicit NoOp(const std::string &name) : name_(name) {}
  std::string name() const override { return name_; }
  std::string kind() const override { return "NoOp"; }
  Json json() const override;
  CLONE_DEF(NoOp)
  void run(Platform &) override {}
};

// host op from a callable - handy for tests, toy cost models and Python ops
class LambdaCpuOp : public CpuOp {
  std::string name_;
  std::function<void(Platform &)> fn_;

public:
  LambdaCpuOp(const std::string &name, std::function<void(Platform &)> fn)
      : name_(name), fn_(std::move(fn)) {}
  std::string name() const override { return name_; }
  CLONE_DEF(LambdaCpuOp)
  void run(Platform &plat) override {
    if (fn_) fn_(plat);
  }
};

template <typename T> class Graph;

// an op that stands for a sub-graph the solver may expand in place
class CompoundOp : public OpBase {
public:
  std::string kind() const override { return "CompoundOp"; }
  virtual Graph<OpBase> graph() const = 0;
};

// ---- helpers (reference: src/operation.cpp:25-84) ----

// drop later duplicates (by eq)
void keep_uniques(std::vector<BoundOpPtr> &v);
// the ways `op` can be placed on the platform: a GpuOp becomes one BoundGpuOp
// per stream, a BoundOp is returned as is
std::vector<BoundOpPtr> make_platform_variations(const Platform &plat, const OpPtr &op);
// true iff the unbound version of e equals the unbound version of some element of v
bool unbound_contains(const std::vector<BoundOpPtr> &v, const OpPtr &e);
bool contains(const std::vector<OpPtr> &v, const OpPtr &e);

} // namespace tenzing
