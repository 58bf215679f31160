"""runs an endless MCTS on a cost model; test_trap sends SIGINT and expects the partial CSV"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")
import tenzing_b200 as tz  # noqa: E402
from tenzing_b200 import _C  # noqa: E402

g = tz.Graph()
for n in "abcdefgh":
    o = tz.NoOp(n)
    g.start_then(o)
    g.then_finish(o)
plat = tz.Platform.make_null(1)


def cost(seq):
    time.sleep(0.002)
    return float(len(seq))


opts = _C.mcts.Opts()
opts.n_iters = 1_000_000
print("READY", flush=True)
_C.mcts.explore(g, plat, tz.PyBenchmarker(cost), "Random", opts)
