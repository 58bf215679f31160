#include "tenzing/benchmarker.hpp"

#include <algorithm>
#include <charconv>
#include <chrono>
#include <cmath>
#include <fstream>
#include <map>
#include <numeric>
#include <random>
#include <sstream>

#include "tenzing/communicator.hpp"
#include "tenzing/cuda/cuda_runtime.hpp"
#include "tenzing/cuda/device_barrier.hpp"
#include "tenzing/numeric.hpp"
#include "tenzing/operation_serdes.hpp"
#include "tenzing/randomness.hpp"

namespace tenzing {

using Benchmark::Opts;
using Benchmark::Result;

namespace {
double wall() {
  using clk = std::chrono::steady_clock;
  return std::chrono::duration<double>(clk::now().time_since_epoch()).count();
}
} // namespace

Result Result::from_samples(std::vector<double> s) {
  Result r;
  if (s.empty()) return r;
  std::sort(s.begin(), s.end());
  r.pct01 = percentile_sorted(s, 1);
  r.pct10 = percentile_sorted(s, 10);
  r.pct50 = percentile_sorted(s, 50);
  r.pct90 = percentile_sorted(s, 90);
  r.pct99 = percentile_sorted(s, 99);
  r.stddev = ::tenzing::stddev(s);
  return r;
}

void run_sequence(Sequence<BoundOp> &order, Platform &plat) {
  plat.next_epoch();
  for (auto &op : order) op->run(plat);
}

// ---- host-wall benchmarker ----
double EmpiricalBenchmarker::timed_run(Sequence<BoundOp> &order, Platform &plat, size_t nSamples) {
  plat.comm().barrier();
  const double t0 = wall();
  for (size_t i = 0; i < nSamples; ++i) run_sequence(order, plat);
  return wall() - t0;
}

EmpiricalBenchmarker::Measurement EmpiricalBenchmarker::measure(Sequence<BoundOp> &order,
                                                                Platform &plat, size_t hint,
                                                                double targetSecs) {
  Measurement m{std::max<size_t>(hint, 1), 0};
  while (true) {
    double elapsed = plat.comm().max_over_ranks(timed_run(order, plat, m.nSamples));
    if (elapsed >= targetSecs || m.nSamples >= (size_t(1) << 24)) {
      m.time = elapsed / double(m.nSamples);
      return m;
    }
    // step half-way towards 1.1x the estimated sample count (same rule on every rank,
    // elapsed is already the max over ranks)
    const double perSample = std::max(elapsed / double(m.nSamples), 1e-9);
    const double est = 1.1 * targetSecs / perSample;
    const double step = std::ceil((est - double(m.nSamples)) * 0.5);
    m.nSamples += size_t(std::max(1.0, step));
  }
}

Result EmpiricalBenchmarker::benchmark(Sequence<BoundOp> &order, Platform &plat, const Opts &opts) {
  std::vector<double> times;
  for (size_t attempt = 0; attempt <= opts.maxRetries; ++attempt) {
    times.clear();
    Measurement m = measure(order, plat, 1, opts.targetSecs); // calibrate
    for (size_t i = 0; i < opts.nIters; ++i) {
      m = measure(order, plat, m.nSamples, opts.targetSecs);
      times.push_back(m.time);
    }
    // every rank holds the same max-reduced series, so the verdict agrees everywhere
    if (!opts.checkRandomness || !randomness::compound_test(times)) break;
    if (attempt < opts.maxRetries) TENZING_INFO("measurement series failed the runs test, retrying");
  }
  return Result::from_samples(times);
}

std::vector<Result> EmpiricalBenchmarker::benchmark(std::vector<Schedule> &schedules,
                                                    Platform &plat, const Opts &opts) {
  Communicator &comm = plat.comm();
  std::vector<int> perm(schedules.size());
  std::iota(perm.begin(), perm.end(), 0);
  std::vector<std::vector<double>> times(schedules.size());
  std::mt19937 rng(12345);
  for (size_t i = 0; i < opts.nIters; ++i) {
    if (comm.rank() == 0) std::shuffle(perm.begin(), perm.end(), rng);
    comm.bcast(perm.data(), perm.size() * sizeof(int), 0);
    for (int si : perm) times[si].push_back(timed_run(schedules[si].order, plat, 1));
  }
  std::vector<Result> out;
  for (auto &t : times) {
    comm.allreduce_max(t.data(), t.size());
    out.push_back(Result::from_samples(t));
  }
  return out;
}

// ---- device-timed benchmarker ----
DeviceBenchmarker::~DeviceBenchmarker() {
  for (auto e : starts_) cudaEventDestroy(e);
  for (auto e : stops_) cudaEventDestroy(e);
  if (fence_) cudaEventDestroy(fence_);
}

double DeviceBenchmarker::timed_run(Sequence<BoundOp> &order, Platform &plat, size_t nSamples) {
  if (plat.is_null()) return EmpiricalBenchmarker::timed_run(order, plat, nSamples);
  const int ns = plat.num_streams();
  while (int(starts_.size()) < 1) {
    cudaEvent_t e;
    CUDA_RUNTIME(cudaEventCreate(&e));
    starts_.push_back(e);
  }
  while (int(stops_.size()) < ns + 1) {
    cudaEvent_t e;
    CUDA_RUNTIME(cudaEventCreate(&e));
    stops_.push_back(e);
  }
  plat.synchronize();
  plat.comm().barrier();
  // start: recorded on stream 0, every other stream starts behind it
  cudaStream_t s0 = plat.cuda_stream(plat.streams_[0]);
  if (plat.device_barrier()) plat.device_barrier()->enqueue(s0); // line the ranks up on the GPUs
  CUDA_RUNTIME(cudaEventRecord(starts_[0], s0));
  for (int i = 1; i < ns; ++i)
    CUDA_RUNTIME(cudaStreamWaitEvent(plat.cuda_stream(plat.streams_[i]), starts_[0], 0));
  for (size_t i = 0; i < nSamples; ++i) run_sequence(order, plat);
  // stop: stream 0 joins every stream, then records
  for (int i = 1; i < ns; ++i) {
    CUDA_RUNTIME(cudaEventRecord(stops_[i], plat.cuda_stream(plat.streams_[i])));
    CUDA_RUNTIME(cudaStreamWaitEvent(s0, stops_[i], 0));
  }
  CUDA_RUNTIME(cudaEventRecord(stops_[ns], s0));
  CUDA_RUNTIME(cudaEventSynchronize(stops_[ns]));
  float ms = 0;
  CUDA_RUNTIME(cudaEventElapsedTime(&ms, starts_[0], stops_[ns]));
  return double(ms) * 1e-3;
}

// ---- CSV replay ----
namespace {
// shortest text that reads back as exactly the same double: a results file must reproduce
// the search that wrote it when it is replayed (CachingBenchmarker)
std::string exact(double v) {
  char buf[40];
  auto res = std::to_chars(buf, buf + sizeof(buf), v);
  return std::string(buf, res.ptr);
}
} // namespace

std::string csv_row(size_t idx, const Result &r, const Sequence<BoundOp> &seq) {
  std::ostringstream ss;
  ss << idx << "|" << exact(r.pct01) << "|" << exact(r.pct10) << "|" << exact(r.pct50) << "|"
     << exact(r.pct90) << "|" << exact(r.pct99) << "|" << exact(r.stddev);
  for (const auto &op : seq) ss << "|" << op->json().dump();
  return ss.str();
}

void CsvBenchmarker::load(std::istream &is, const Graph<OpBase> &g) {
  std::string line;
  size_t lineNo = 0;
  while (std::getline(is, line)) {
    ++lineNo;
    if (line.empty() || line[0] == '#') continue;
    if (line[0] == '{') { // reproduce header (or another JSON header line, which is skipped)
      const reproduce::Header h = reproduce::parse_header(line);
      if (h.valid && !producer_.valid) {
        producer_ = h;
        if (!h.same_build())
          TENZING_INFO("csv written by build " << h.version() << ", this is " << reproduce::version_string());
      }
      continue;
    }
    std::vector<std::string> cols;
    {
      std::stringstream ss(line);
      std::string c;
      while (std::getline(ss, c, '|')) cols.push_back(c);
    }
    if (cols.size() < 8) continue;
    DataRow row;
    try {
      row.res.pct01 = std::stod(cols[1]);
      row.res.pct10 = std::stod(cols[2]);
      row.res.pct50 = std::stod(cols[3]);
      row.res.pct90 = std::stod(cols[4]);
      row.res.pct99 = std::stod(cols[5]);
      row.res.stddev = std::stod(cols[6]);
    } catch (const std::exception &) {
      if (data_.empty()) continue; // a textual column header
      THROW_RUNTIME("csv line " << lineNo << ": bad statistics columns");
    }
    // the row's ops are resolved together: alternatives of a ChoiceOp may share op names
    Json ops = Json::array();
    for (size_t c = 7; c < cols.size(); ++c) {
      if (cols[c].empty()) continue; // padded rows
      ops.push_back(Json::parse(cols[c]));
    }
    from_json(ops, g, row.seq);
    byKey_.emplace(canonical_key(row.seq), data_.size());
    data_.push_back(std::move(row));
  }
}

CsvBenchmarker::CsvBenchmarker(const std::string &path, const Graph<OpBase> &g) {
  std::ifstream is(path);
  if (!is) THROW_RUNTIME("cannot open " << path);
  load(is, g);
}

CsvBenchmarker CsvBenchmarker::from_text(const std::string &text, const Graph<OpBase> &g) {
  CsvBenchmarker b;
  std::istringstream is(text);
  b.load(is, g);
  return b;
}

// ---- result cache / replay ----
void CachingBenchmarker::preload(const Sequence<BoundOp> &seq, const Result &r) {
  auto ins = known_.emplace(canonical_key(seq), Entry{});
  Entry &e = ins.first->second;
  if (ins.second) e.first = r;
  e.recorded.push_back(r);
  ++pending_;
}

void CachingBenchmarker::preload(const CsvBenchmarker &csv) {
  for (const auto &row : csv.rows()) preload(row.seq, row.res);
}

Result CachingBenchmarker::benchmark(Sequence<BoundOp> &order, Platform &plat, const Opts &opts) {
  const std::string key = canonical_key(order);
  auto it = known_.find(key);
  if (it != known_.end()) {
    Entry &e = it->second;
    if (e.next < e.recorded.size()) { // retrace the recorded run
      ++hits_;
      --pending_;
      return e.recorded[e.next++];
    }
    if (reuse_) {
      ++hits_;
      return e.first;
    }
  }
  ++misses_;
  const Result r = inner_.benchmark(order, plat, opts);
  if (it == known_.end()) {
    Entry e;
    e.first = r;
    known_.emplace(key, std::move(e));
  }
  return r;
}

Result CsvBenchmarker::benchmark(Sequence<BoundOp> &order, Platform &, const Opts &) {
  auto it = byKey_.find(canonical_key(order));
  if (it == byKey_.end()) THROW_RUNTIME("no CSV row matches sequence: " << order.desc());
  return data_[it->second].res;
}

// ---- timeline simulation ----
double SimBenchmarker::simulate(const Sequence<BoundOp> &order) const {
  const std::vector<Span> t = timeline(order);
  double end = 0;
  for (const Span &s : t) end = std::max(end, s.end);
  // the sequence is over when the host has issued (and waited through) its last op
  return t.empty() ? 0 : t.back().end;
}

std::vector<SimBenchmarker::Span> SimBenchmarker::timeline(const Sequence<BoundOp> &order) const {
  std::vector<Span> spans;
  double host = 0;                               // host clock
  std::map<int, double> streamReady;             // stream id -> time its last op finishes
  std::map<int, double> eventTime;               // event id -> time it fires
  struct Running {
    double start, end;
  };
  std::vector<Running> running;                  // GPU ops placed so far (for contention)
  auto lookup = [](const std::unordered_map<std::string, double> &m, const std::string &k, double d) {
    auto it = m.find(k);
    return it == m.end() ? d : it->second;
  };
  for (const auto &op : order) {
    const double issued = host;
    host += model_.launchCost;
    if (auto g = std::dynamic_pointer_cast<BoundGpuOp>(op)) {
      const int s = g->stream().id();
      double start = std::max(host, streamReady[s]);
      double dur = lookup(model_.gpuSeconds, g->name(), model_.defaultGpuSeconds);
      if (model_.contention > 0) {
        int overlapping = 0;
        for (const auto &r : running)
          if (r.end > start) ++overlapping;
        dur *= 1.0 + model_.contention * overlapping;
      }
      streamReady[s] = start + dur;
      running.push_back(Running{start, start + dur});
      spans.push_back(Span{start, start + dur});
      continue;
    } else if (auto cer = std::dynamic_pointer_cast<CudaEventRecord>(op)) {
      eventTime[cer->event().id()] = std::max(host, streamReady[cer->stream().id()]);
    } else if (auto w = std::dynamic_pointer_cast<CudaStreamWaitEvent>(op)) {
      double &ready = streamReady[w->stream().id()];
      ready = std::max(ready, eventTime[w->event().id()]);
    } else if (auto ces = std::dynamic_pointer_cast<CudaEventSync>(op)) {
      host = std::max(host, eventTime[ces->event().id()]);
    } else if (auto ss = std::dynamic_pointer_cast<StreamSync>(op)) {
      host = std::max(host, streamReady[ss->stream().id()]);
    } else if (auto sw = std::dynamic_pointer_cast<StreamWait>(op)) {
      double &ready = streamReady[sw->waiter().id()];
      ready = std::max(ready, std::max(host, streamReady[sw->waitee().id()]));
    } else {
      host += lookup(model_.cpuSeconds, op->name(), model_.defaultCpuSeconds);
    }
    spans.push_back(Span{issued, host});
  }
  return spans;
}

Result SimBenchmarker::benchmark(Sequence<BoundOp> &order, Platform &, const Opts &) {
  const double t = simulate(order);
  Result r;
  r.pct01 = r.pct10 = r.pct50 = r.pct90 = r.pct99 = t;
  return r;
}

// ---- cost model ----
Result ModelBenchmarker::benchmark(Sequence<BoundOp> &order, Platform &, const Opts &) {
  const double t = model_(order);
  Result r;
  r.pct01 = r.pct10 = r.pct50 = r.pct90 = r.pct99 = t;
  r.stddev = 0;
  return r;
}

} // namespace tenzing
