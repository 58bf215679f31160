#include "tenzing/dfs.hpp"

#include <sstream>
#include <unordered_set>

#include "tenzing/communicator.hpp"
#include "tenzing/counters.hpp"
#include "tenzing/schedule.hpp"
#include "tenzing/state.hpp"
#include "tenzing/trap.hpp"

namespace tenzing {
namespace dfs {

void to_json(Json &j, const Opts &opts) {
  Json o = Json::object();
  o["maxSeqs"] = opts.maxSeqs;
  j = Json::object();
  j["dfs__Opts"] = o;
}

void Result::dump_csv(std::ostream &os) const {
  Json j;
  to_json(j, opts_);
  os << j.dump() << "\n";
  for (size_t i = 0; i < simResults.size(); ++i)
    os << csv_row(i, simResults[i].benchResult, simResults[i].seq) << "\n";
  os.flush();
}

std::string Result::csv() const {
  std::ostringstream ss;
  dump_csv(ss);
  return ss.str();
}

int Result::best() const {
  int b = -1;
  for (size_t i = 0; i < simResults.size(); ++i)
    if (b < 0 || simResults[i].benchResult.pct10 < simResults[b].benchResult.pct10) b = int(i);
  return b;
}

void Stop::bcast(int root, Communicator &comm) {
  int64_t v = value_ ? 1 : 0;
  v = comm.bcast_int(v, root);
  value_ = v != 0;
}

namespace {
// The set of digests seen so far.  One probe sequence in one flat array per look-up: with
// millions of states a node-based set spends its time in cache misses and allocations.
class KeySet {
  std::vector<StateKey> slots_; // {0,0} = empty
  size_t used_ = 0;
  static StateKey norm(StateKey k) {
    if (k.a == 0 && k.b == 0) k.b = 1;
    return k;
  }
  bool place(const StateKey &k) {
    const size_t mask = slots_.size() - 1;
    for (size_t i = StateKeyHasher{}(k) & mask;; i = (i + 1) & mask) {
      StateKey &s = slots_[i];
      if (s.a == 0 && s.b == 0) {
        s = k;
        return true;
      }
      if (s == k) return false;
    }
  }

public:
  KeySet() : slots_(1024) {}
  // true if the key was not there yet
  bool insert(StateKey k) {
    k = norm(k);
    if ((used_ + 1) * 2 > slots_.size()) {
      std::vector<StateKey> old(slots_.size() * 2);
      old.swap(slots_);
      for (const StateKey &o : old)
        if (o.a || o.b) place(o);
    }
    const bool fresh = place(k);
    used_ += fresh;
    return fresh;
  }
};
} // namespace

std::vector<Sequence<BoundOp>> get_all_sequences(const Graph<OpBase> &g, const Platform &plat,
                                                 int64_t maxSeqs, bool removeRedundantSyncs) {
  const double t0 = counters::now();
  std::vector<Sequence<BoundOp>> ret;
  KeySet seenStates, seenSeqs;
  std::vector<SDP::State> stack;
  stack.emplace_back(g);
  seenStates.insert(stack.back().canonical_hash());
  const uint64_t limit = maxSeqs < 0 ? UINT64_MAX : uint64_t(maxSeqs);
  while (!stack.empty() && ret.size() < limit) {
    SDP::State cur = std::move(stack.back());
    stack.pop_back();
    TENZING_COUNTER_OP(dfs, STATES_VISITED, += 1);
    if (cur.is_terminal()) {
      Sequence<BoundOp> seq = cur.sequence();
      if (removeRedundantSyncs) Schedule::remove_redundant_syncs(seq);
      const StateKey k = removeRedundantSyncs ? sequence_hash(seq) : cur.sequence_hash();
      if (seenSeqs.insert(k)) ret.push_back(std::move(seq));
      continue;
    }
    std::vector<SDP::State> next = cur.frontier(plat);
    // push in reverse so the first decision is explored first
    for (auto it = next.rbegin(); it != next.rend(); ++it) {
      if (seenStates.insert(it->canonical_hash())) {
        stack.push_back(std::move(*it));
      } else {
        TENZING_COUNTER_OP(dfs, STATES_PRUNED, += 1);
      }
    }
  }
  TENZING_COUNTER_OP(dfs, ENUMERATE_TIME, += counters::now() - t0);
  return ret;
}

Result explore(const Graph<OpBase> &g, Platform &plat, Benchmarker &benchmarker, const Opts &opts) {
  Communicator &comm = plat.comm();
  const int rank = comm.rank();
  Result res(opts);
  std::vector<Sequence<BoundOp>> seqs;
  if (rank == 0) seqs = get_all_sequences(g, plat, opts.maxSeqs, opts.removeRedundantSyncs);

  if (rank == 0) register_handler([&res](int) { res.dump_csv(); });
  size_t i = 0;
  while (true) {
    Stop stop;
    if (rank == 0 && i >= seqs.size()) stop = Stop(true, Stop::Reason::predicate);
    stop.bcast(0, comm);
    if (stop) break;
    Sequence<BoundOp> sut;
    if (rank == 0) sut = seqs[i++];
    sut = bcast(sut, g, comm, 0);
    plat.provision_events(sut);
    SimResult sr;
    sr.seq = sut;
    const double t0 = counters::now();
    sr.benchResult = benchmarker.benchmark(sut, plat, opts.benchOpts);
    TENZING_COUNTER_OP(dfs, BENCHMARK_TIME, += counters::now() - t0);
    res.simResults.push_back(std::move(sr));
  }
  if (rank == 0) unregister_handler();
  return res;
}

} // namespace dfs
} // namespace tenzing
