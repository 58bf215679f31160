"""2-rank CPU worker (gloo): DFS + MCTS in lock step over torch.distributed, halo/SpMV set-up
in host-memory mode is NOT possible across processes, so only the control plane is used here."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")
import torch.distributed as dist  # noqa: E402

import tenzing_b200 as tz  # noqa: E402
from tenzing_b200 import _C  # noqa: E402
from tenzing_b200.parallel import init_process_group_from_env, make_comm  # noqa: E402

rank, world, _ = init_process_group_from_env(backend="gloo")
comm = make_comm()
assert comm.size() == 2 and comm.rank() == rank
g = tz.Graph()
op1 = tz.NoOp("op1")
g.start_then(op1)
g.then_finish(op1)
plat = tz.Platform.make_null(2, comm)
res = tz.dfs.explore(g, plat, tz.PyBenchmarker(lambda s: 1.0))
assert [r.seq.names() for r in res.sim_results] == [["Start", "op1", "Finish"]]

# a graph with GPU ops: rank 0 searches, rank 1 re-resolves every broadcast op by name
k1, k2 = tz.EmptyGpuOp("k1"), tz.EmptyGpuOp("k2")
g2 = tz.Graph()
g2.start_then(k1)
g2.then(k1, k2)
g2.then_finish(k2)
opts = _C.mcts.Opts()
opts.n_iters = 6
res = _C.mcts.explore(g2, plat, tz.PyBenchmarker(lambda s: float(len(s))), "FastMin", opts)
keys = [r.path.canonical_key() for r in res.sim_results]
gathered = [None, None]
dist.all_gather_object(gathered, keys)
assert gathered[0] == gathered[1] and len(keys) > 0
dist.barrier()
if rank == 0:
    print("GLOO_WORKER_OK", flush=True)
dist.destroy_process_group()
