#include "tenzing/schedule.hpp"

#include <limits>
#include <random>
#include <unordered_map>

namespace tenzing {

namespace {

using Seq = Sequence<BoundOp>;

// what an op of the sequence is, from its cached type bits: these scans run over every pair of
// positions for every redundancy removed, so no reference counting and no failing casts
template <typename T> const T *as(const BoundOpPtr &p);
template <> const CudaEventRecord *as<CudaEventRecord>(const BoundOpPtr &p) {
  return p->is(op_trait::EventRecord) ? static_cast<const CudaEventRecord *>(p.get()) : nullptr;
}
template <> const CudaEventSync *as<CudaEventSync>(const BoundOpPtr &p) {
  return p->is(op_trait::EventSync) ? static_cast<const CudaEventSync *>(p.get()) : nullptr;
}
template <> const CudaStreamWaitEvent *as<CudaStreamWaitEvent>(const BoundOpPtr &p) {
  return p->is(op_trait::StreamWaitEvent) ? static_cast<const CudaStreamWaitEvent *>(p.get()) : nullptr;
}
template <> const StreamSync *as<StreamSync>(const BoundOpPtr &p) {
  return p->is(op_trait::Sync) ? dynamic_cast<const StreamSync *>(p.get()) : nullptr;
}

bool gpu_op_on_stream(const BoundOpPtr &op, const Stream &s) {
  return op->is(op_trait::BoundGpu) && static_cast<const BoundGpuOp &>(*op).stream() == s;
}

// positions of the waits that consume event e after position `from`
struct Consumers {
  std::vector<size_t> hostSyncs, streamWaits;
};
Consumers consumers_of(const Seq &o, size_t from, const Event &e) {
  Consumers c;
  for (size_t i = from + 1; i < o.size(); ++i) {
    if (auto ces = as<CudaEventSync>(o[i])) {
      if (ces->event() == e) c.hostSyncs.push_back(i);
    } else if (auto w = as<CudaStreamWaitEvent>(o[i])) {
      if (w->event() == e) c.streamWaits.push_back(i);
    } // a legacy StreamWait records and consumes its own event
  }
  return c;
}

void erase_positions(Seq &o, std::vector<size_t> pos) {
  std::sort(pos.begin(), pos.end(), std::greater<size_t>());
  for (size_t p : pos) o.erase(o.begin() + p);
}

// one pass; returns number removed (0 = fix point)
int one_pass(Seq &o) {
  // R1: record nobody consumes
  for (size_t i = 0; i < o.size(); ++i) {
    if (auto cer = as<CudaEventRecord>(o[i])) {
      const Consumers c = consumers_of(o, i, cer->event());
      if (c.hostSyncs.empty() && c.streamWaits.empty()) {
        erase_positions(o, {i});
        return 1;
      }
    }
  }
  // R2: stream wait with no later GPU op in the waiting stream
  for (size_t i = 0; i < o.size(); ++i) {
    if (auto w = as<CudaStreamWaitEvent>(o[i])) {
      bool used = false;
      for (size_t k = i + 1; k < o.size() && !used; ++k) used = gpu_op_on_stream(o[k], w->stream());
      if (!used) {
        erase_positions(o, {i});
        return 1;
      }
    }
  }
  // R6a: wait on the stream that recorded the event, nothing launched there in between
  // R6b: second host sync of one event
  for (size_t i = 0; i < o.size(); ++i) {
    auto cer = as<CudaEventRecord>(o[i]);
    if (!cer) continue;
    const Consumers c = consumers_of(o, i, cer->event());
    if (c.hostSyncs.size() > 1) {
      erase_positions(o, {c.hostSyncs[1]});
      return 1;
    }
    for (size_t wpos : c.streamWaits) {
      auto w = as<CudaStreamWaitEvent>(o[wpos]);
      if (w->stream() == cer->stream()) { // in-order stream already guarantees it
        erase_positions(o, {wpos});
        return 1;
      }
    }
  }
  // R3: two StreamSyncs of one stream, no GPU op of that stream between -> drop the first
  for (size_t i = 0; i < o.size(); ++i) {
    auto s1 = as<StreamSync>(o[i]);
    if (!s1) continue;
    for (size_t k = i + 1; k < o.size(); ++k) {
      if (gpu_op_on_stream(o[k], s1->stream())) break;
      auto s2 = as<StreamSync>(o[k]);
      if (s2 && s2->stream() == s1->stream()) {
        erase_positions(o, {i});
        return 1;
      }
    }
  }
  // R4 / R5: pairs of records in one stream that are both consumed by exactly one host sync
  for (size_t i = 0; i < o.size(); ++i) {
    auto r1 = as<CudaEventRecord>(o[i]);
    if (!r1) continue;
    const Consumers c1 = consumers_of(o, i, r1->event());
    if (c1.hostSyncs.size() != 1 || !c1.streamWaits.empty()) continue;
    bool sameState = true; // no GPU op of the stream between the two records
    for (size_t k = i + 1; k < o.size(); ++k) {
      if (gpu_op_on_stream(o[k], r1->stream())) sameState = false;
      auto r2 = as<CudaEventRecord>(o[k]);
      if (!r2 || !(r2->stream() == r1->stream())) continue;
      const Consumers c2 = consumers_of(o, k, r2->event());
      if (c2.hostSyncs.size() != 1 || !c2.streamWaits.empty()) continue;
      const size_t w1 = c1.hostSyncs[0], w2 = c2.hostSyncs[0];
      if (w2 < w1) {
        // R5 (and R4 when sameState): r2 is at or after r1 in the stream and the host
        // already waited for it -> waiting for r1 afterwards is a no-op
        erase_positions(o, {i, w1});
        return 2;
      }
      if (sameState && w1 < w2) {
        // R4: identical stream state, host already waited at w1 -> second pair is a no-op
        erase_positions(o, {k, w2});
        return 2;
      }
    }
  }
  return 0;
}

// ---- enumeration of topological orders ----
struct Dag {
  std::vector<BoundOpPtr> ops;
  std::vector<std::vector<int>> succs;
  std::vector<int> indeg;
};

Dag to_dag(const Graph<OpBase> &g) {
  Dag d;
  std::unordered_map<uint32_t, int> idx;
  for (const auto &v : g.vertices()) {
    auto b = std::dynamic_pointer_cast<BoundOp>(v);
    if (!b) THROW_RUNTIME("make_schedules: " << v->desc() << " is not a BoundOp");
    idx[v->uid()] = int(d.ops.size());
    d.ops.push_back(b);
  }
  d.succs.resize(d.ops.size());
  d.indeg.assign(d.ops.size(), 0);
  for (const auto &v : g.vertices()) {
    for (const auto &s : g.succs(v)) {
      d.succs[idx[v->uid()]].push_back(idx[s->uid()]);
      d.indeg[idx[s->uid()]]++;
    }
  }
  return d;
}

void enumerate(const Dag &d, std::vector<int> &indeg, std::vector<char> &used, std::vector<int> &cur,
               std::vector<Schedule> &out, size_t maxCount) {
  if (maxCount && out.size() >= maxCount) return;
  if (cur.size() == d.ops.size()) {
    Schedule s;
    for (int i : cur) s.order.push_back(d.ops[i]);
    out.push_back(std::move(s));
    return;
  }
  for (int i = 0; i < int(d.ops.size()); ++i) {
    if (used[i] || indeg[i]) continue;
    used[i] = 1;
    for (int s : d.succs[i]) --indeg[s];
    cur.push_back(i);
    enumerate(d, indeg, used, cur, out, maxCount);
    cur.pop_back();
    for (int s : d.succs[i]) ++indeg[s];
    used[i] = 0;
  }
}

uint64_t sat_add(uint64_t a, uint64_t b) {
  const uint64_t m = std::numeric_limits<uint64_t>::max();
  return (a > m - b) ? m : a + b;
}

// number of completions from the set `done` (bitmask), memoised
uint64_t completions(const Dag &d, uint64_t done, std::unordered_map<uint64_t, uint64_t> &memo) {
  const int n = int(d.ops.size());
  if (done == ((n == 64) ? ~0ULL : ((1ULL << n) - 1))) return 1;
  auto it = memo.find(done);
  if (it != memo.end()) return it->second;
  uint64_t total = 0;
  for (int i = 0; i < n; ++i) {
    if (done >> i & 1) continue;
    bool ready = true;
    for (int p = 0; p < n && ready; ++p) {
      if (done >> p & 1) continue;
      for (int s : d.succs[p])
        if (s == i) {
          ready = false;
          break;
        }
    }
    if (ready) total = sat_add(total, completions(d, done | (1ULL << i), memo));
  }
  memo[done] = total;
  return total;
}

} // namespace

int Schedule::remove_redundant_syncs(Sequence<BoundOp> &order) {
  int removed = 0;
  while (int n = one_pass(order)) removed += n;
  return removed;
}

bool Schedule::predicate(const Schedule &a, const Schedule &b) {
  return bool(get_equivalence(a.order, b.order));
}

bool Schedule::by_op_typeid(const BoundOpPtr &a, const BoundOpPtr &b) {
  if (a->kind() != b->kind()) return a->kind() < b->kind();
  return a->key() < b->key();
}

std::vector<Schedule> make_schedules(const Graph<OpBase> &g, size_t maxCount) {
  const Dag d = to_dag(g);
  std::vector<int> indeg = d.indeg, cur;
  std::vector<char> used(d.ops.size(), 0);
  std::vector<Schedule> out;
  enumerate(d, indeg, used, cur, out, maxCount);
  return out;
}

uint64_t count_schedules(const Graph<OpBase> &g) {
  const Dag d = to_dag(g);
  if (d.ops.size() > 62) THROW_RUNTIME("count_schedules: more than 62 vertices");
  std::unordered_map<uint64_t, uint64_t> memo;
  return completions(d, 0, memo);
}

std::vector<Schedule> make_schedules_random(const Graph<OpBase> &g, size_t n, uint64_t seed) {
  const Dag d = to_dag(g);
  const int nv = int(d.ops.size());
  const bool exact = nv <= 62;
  std::unordered_map<uint64_t, uint64_t> memo;
  std::mt19937_64 rng(seed);
  std::vector<Schedule> out;
  for (size_t k = 0; k < n; ++k) {
    std::vector<int> indeg = d.indeg;
    std::vector<char> used(nv, 0);
    uint64_t done = 0;
    Schedule s;
    for (int step = 0; step < nv; ++step) {
      std::vector<int> ready;
      std::vector<double> weight;
      for (int i = 0; i < nv; ++i) {
        if (used[i] || indeg[i]) continue;
        ready.push_back(i);
        weight.push_back(exact ? double(completions(d, done | (1ULL << i), memo)) : 1.0);
      }
      std::discrete_distribution<int> pick(weight.begin(), weight.end());
      const int i = ready[pick(rng)];
      used[i] = 1;
      if (exact) done |= 1ULL << i;
      for (int t : d.succs[i]) --indeg[t];
      s.order.push_back(d.ops[i]);
    }
    out.push_back(std::move(s));
  }
  return out;
}

} // namespace tenzing
