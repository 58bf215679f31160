// Turning a sequence into a time.
//
// Reference: include/tenzing/benchmarker.hpp:14-57, src/benchmarker.cpp:21-223.
//  * EmpiricalBenchmarker: host-wall measurement with the reference's protocol
//    (calibrate samples/measurement to >= 10 ms, nIters measurements, max over
//    ranks, runs-test gate with retries, percentiles).
//  * DeviceBenchmarker: same protocol but each measurement is timed on the
//    device with CUDA events recorded on every platform stream (first start ->
//    last stop), max over ranks - what BASELINE.json calls "device-timed".
//  * CsvBenchmarker: replays a results CSV (both the DFS form with an opts line
//    and the header-less MCTS form).
//  * ModelBenchmarker: a cost model callback - toy problems and solver tests.
#pragma once

#include <functional>
#include <istream>
#include <unordered_map>
#include <string>
#include <vector>

#include "tenzing/graph.hpp"
#include "tenzing/platform.hpp"
#include "tenzing/schedule.hpp"
#include "tenzing/reproduce.hpp"
#include "tenzing/sequence.hpp"

namespace tenzing {

namespace Benchmark {
struct Result {
  double pct01 = 0, pct10 = 0, pct50 = 0, pct90 = 0, pct99 = 0, stddev = 0;
  static Result from_samples(std::vector<double> samples);
};
struct Opts {
  size_t nIters = 1000;     // measurements per sequence
  size_t maxRetries = 10;   // re-measure while the series fails the runs test
  double targetSecs = 0.01; // minimum duration of one measurement
  bool checkRandomness = true;
};
} // namespace Benchmark

class Benchmarker {
public:
  virtual ~Benchmarker() = default;
  virtual Benchmark::Result benchmark(Sequence<BoundOp> &order, Platform &plat,
                                      const Benchmark::Opts &opts = Benchmark::Opts()) = 0;
};

// run `order` once under a fresh epoch
void run_sequence(Sequence<BoundOp> &order, Platform &plat);

class EmpiricalBenchmarker : public Benchmarker {
public:
  Benchmark::Result benchmark(Sequence<BoundOp> &order, Platform &plat,
                              const Benchmark::Opts &opts = Benchmark::Opts()) override;
  // interleaved measurement of several schedules in a fresh random order per iteration
  std::vector<Benchmark::Result> benchmark(std::vector<Schedule> &schedules, Platform &plat,
                                           const Benchmark::Opts &opts = Benchmark::Opts());

protected:
  struct Measurement {
    size_t nSamples;
    double time; // seconds per sample, max over ranks
  };
  // one measurement of nSamples back-to-back executions
  virtual double timed_run(Sequence<BoundOp> &order, Platform &plat, size_t nSamples);
  Measurement measure(Sequence<BoundOp> &order, Platform &plat, size_t nSamplesHint,
                      double targetSecs);
};

class DeviceBenchmarker : public EmpiricalBenchmarker {
  std::vector<cudaEvent_t> starts_, stops_;
  cudaEvent_t fence_ = nullptr;

public:
  ~DeviceBenchmarker() override;

protected:
  double timed_run(Sequence<BoundOp> &order, Platform &plat, size_t nSamples) override;
};

class CsvBenchmarker : public Benchmarker {
public:
  struct DataRow {
    Benchmark::Result res;
    Sequence<BoundOp> seq;
  };
  CsvBenchmarker(const std::string &path, const Graph<OpBase> &g);
  static CsvBenchmarker from_text(const std::string &text, const Graph<OpBase> &g);
  Benchmark::Result benchmark(Sequence<BoundOp> &order, Platform &plat,
                              const Benchmark::Opts &opts = Benchmark::Opts()) override;
  const std::vector<DataRow> &rows() const { return data_; }
  // reproduce header found in the file (valid == false if there was none)
  const reproduce::Header &producer() const { return producer_; }

private:
  CsvBenchmarker() = default;
  reproduce::Header producer_;
  void load(std::istream &is, const Graph<OpBase> &g);
  std::vector<DataRow> data_;
  std::unordered_map<std::string, size_t> byKey_; // canonical key -> row
};

// Replays recorded results and (optionally) remembers new ones, asking the wrapped benchmarker
// only when it has to.  Sequences are identified up to stream/event renaming (canonical key).
//  * resume: preload the CSV of an interrupted search and run the solver again with the same
//    seed.  Recorded results are handed out per sequence in the order they were recorded, each
//    once, so the search retraces the interrupted run exactly without touching the GPU and
//    starts measuring where that run stopped (the reference cannot resume a search: SURVEY 5.4);
//  * reuse = true: a sequence that has a result (replayed or measured) is never measured
//    again.  MCTS rollouts that differ only in redundant synchronisation collapse to one
//    sequence after cleaning, so a budgeted search stops paying for the same schedule twice.
// Every rank must hold the same recorded results (same preloaded file), so that all ranks
// agree on whether a sequence is measured (a collective) or looked up.
class CachingBenchmarker : public Benchmarker {
public:
  explicit CachingBenchmarker(Benchmarker &inner, bool reuse = true) : inner_(inner), reuse_(reuse) {}
  void preload(const Sequence<BoundOp> &seq, const Benchmark::Result &r);
  void preload(const CsvBenchmarker &csv);
  Benchmark::Result benchmark(Sequence<BoundOp> &order, Platform &plat,
                              const Benchmark::Opts &opts = Benchmark::Opts()) override;
  size_t hits() const { return hits_; }       // answered without measuring
  size_t misses() const { return misses_; }   // measured
  size_t size() const { return known_.size(); } // distinct sequences with a result
  size_t pending() const { return pending_; } // recorded results not handed out yet

private:
  struct Entry {
    Benchmark::Result first;                 // what `reuse` hands out
    std::vector<Benchmark::Result> recorded; // preloaded results, in order
    size_t next = 0;                         // next recorded result to hand out
  };
  Benchmarker &inner_;
  bool reuse_;
  std::unordered_map<std::string, Entry> known_; // canonical key -> results
  size_t hits_ = 0, misses_ = 0, pending_ = 0;
};

class ModelBenchmarker : public Benchmarker {
  std::function<double(const Sequence<BoundOp> &)> model_;

public:
  explicit ModelBenchmarker(std::function<double(const Sequence<BoundOp> &)> model)
      : model_(std::move(model)) {}
  Benchmark::Result benchmark(Sequence<BoundOp> &order, Platform &plat,
                              const Benchmark::Opts &opts = Benchmark::Opts()) override;
};

// Timeline simulation of a sequence: a "fake backend" like the reference's CSV replay
// (benchmarker.hpp:43-57), but predictive.  The host issues ops one after the other
// (`launchCost` each; CudaEventSync blocks it), every stream executes its ops in order, a
// CudaEventRecord captures the time its stream reaches it, a CudaStreamWaitEvent delays its
// stream until that time.  Op durations come from a table keyed by op name (e.g. measured by
// scripts/microbench_halo.py).  Lets search strategies be studied and tested without a GPU.
class SimBenchmarker : public Benchmarker {
public:
  struct Model {
    std::unordered_map<std::string, double> gpuSeconds; // GPU op name -> duration
    std::unordered_map<std::string, double> cpuSeconds; // CPU op name -> host time
    double defaultGpuSeconds = 10e-6;
    double defaultCpuSeconds = 0;
    double launchCost = 3e-6; // host time to issue any op
    // when two GPU ops overlap in time their durations are stretched by this factor per
    // concurrent op beyond the first (0 = perfect overlap, 1 = they share the machine equally)
    double contention = 0;
  };
  explicit SimBenchmarker(Model m) : model_(std::move(m)) {}
  // simulated wall time of one execution of the sequence
  double simulate(const Sequence<BoundOp> &order) const;
  // per op of the sequence: (start, end) on its timeline - GPU ops on their stream, everything
  // else on the host.  Used to check that a schedule respects the program's dependencies.
  struct Span {
    double start, end;
  };
  std::vector<Span> timeline(const Sequence<BoundOp> &order) const;
  Benchmark::Result benchmark(Sequence<BoundOp> &order, Platform &plat,
                              const Benchmark::Opts &opts = Benchmark::Opts()) override;

private:
  Model model_;
};

// one CSV row: idx|pct01|pct10|pct50|pct90|pct99|stddev|<op json>|...
std::string csv_row(size_t idx, const Benchmark::Result &r, const Sequence<BoundOp> &seq);

} // namespace tenzing
