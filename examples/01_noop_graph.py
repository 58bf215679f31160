"""The smallest program: Start -> op1 -> Finish (reference: test/test_noop_graph.cpp)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tenzing_b200 as tz  # noqa: E402

g = tz.Graph()
op1 = tz.NoOp("op1")
g.start_then(op1)
g.then_finish(op1)
plat = tz.Platform.make_null(1)          # ids only: runs without a GPU
state = tz.State(g)
print("sequence:", state.sequence().names())
for d in state.get_decisions(plat):
    print("decision:", d.desc())
print("all sequences:", [s.names() for s in tz.dfs.get_all_sequences(g, plat)])
