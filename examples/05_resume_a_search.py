"""An interrupted search continues from its results file.

The CSV a solver writes, plus the seed, is a checkpoint: CachingBenchmarker hands the recorded
results back in their recorded order, so the (deterministic) solver retraces the first run
without measuring anything and starts measuring where that run stopped.  With reuse=True the
same wrapper also never measures one schedule twice.  No GPU needed: a cost model stands in for
the hardware here."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")

import tenzing_b200 as tz  # noqa: E402
from tenzing_b200 import _C  # noqa: E402

# program: three pairs (fast_i, slow_i); running fast_i before slow_i is cheaper
g = tz.Graph()
for i in range(3):
    for n in (f"fast{i}", f"slow{i}"):
        op = tz.NoOp(n)
        g.start_then(op)
        g.then_finish(op)
calls = []


def cost(seq):
    calls.append(1)
    names = seq.names()
    return sum(5e-3 if names.index(f"fast{i}") < names.index(f"slow{i}") else 10e-3 for i in range(3))


def search(iters, bench):
    opts = _C.mcts.Opts()
    opts.n_iters, opts.seed = iters, 7
    return _C.mcts.explore(g, tz.Platform.make_null(1), bench, "FastMin", opts)


first = search(40, tz.PyBenchmarker(cost))                  # "interrupted" after 40 iterations
print(f"first run : {len(first.sim_results)} iterations, {len(calls)} measurements")

del calls[:]
resumed_bench = tz.CachingBenchmarker(tz.PyBenchmarker(cost), reuse=False)
resumed_bench.preload_csv(tz.CsvBenchmarker.from_text(first.csv(), g))
resumed = search(100, resumed_bench)
assert [r.path.names() for r in resumed.sim_results[:40]] == [r.path.names() for r in first.sim_results]
print(f"resumed   : {len(resumed.sim_results)} iterations, {len(calls)} measurements "
      f"({resumed_bench.hits()} replayed from the file)")
best = resumed.sim_results[resumed.best()]
print(f"best      : {best.bench_result.pct10 * 1e3:.1f} ms  {best.path.desc()}")
