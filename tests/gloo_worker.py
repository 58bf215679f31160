"""2-rank CPU worker (gloo), launched like bench.py is (one process per rank): DFS + MCTS in
lock step over torch.distributed, then both workloads EXECUTED on the CPU - host-memory
symmetric heap mapped across the two processes through POSIX shared memory, host execution of
every op, the cross-process flag protocol - and checked against the oracles."""
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

# the workloads for real, one process per rank
from tenzing_b200.models import DistributedSpMV, HaloExchange, common  # noqa: E402
from tenzing_b200.models import halo as halo_mod, spmv as spmv_mod  # noqa: E402

_C.set_host_wait_timeout(20.0)
halo = HaloExchange(comm=comm, nX=6, nY=5, nZ=4, nQ=2, nGhost=2, host=True)
halo.platform.enable_host_streams()
spmv = DistributedSpMV(comm=comm, n=301, nnz_per_row=7, seed=3, host=True)
spmv.platform.set_host_execution(True)
for wl, styles in ((halo, halo_mod.STYLES), (spmv, spmv_mod.STYLES + spmv_mod.EXPERIMENTAL_STYLES)):
    for style in styles:
        seq = common.fixed_schedule(wl.graph(style), wl.platform)
        wl.fill()
        comm.barrier()
        wl.run(seq, 3)
        wl.platform.synchronize()
        comm.barrier()
        assert wl.check(), (style, rank)
        comm.barrier()
# the collective library itself as the transport (second baseline: "dist" style) - here gloo,
# on GPUs NCCL; a searched schedule of it, not just the fixed one
res = halo.search("mcts", style="dist", iters=4, benchmarker=_C.EmpiricalBenchmarker(),
                  bench=common.bench_opts(n_iters=2, target_secs=1e-5, max_retries=0,
                                          check_randomness=False))
seq, _ = common.best_of(res)
assert "he_dist_post" in seq.names() and "he_dist_wait" in seq.names()
halo.fill()
comm.barrier()
halo.run(seq, 3)
halo.platform.synchronize()
comm.barrier()
assert halo.check(), ("dist", rank)
comm.barrier()
for bw in (None, 10 ** 9):      # neighbours only / all-to-all
    sp = DistributedSpMV(comm=comm, n=301, bw=bw, nnz_per_row=7, seed=3, host=True)
    sp.platform.set_host_execution(True)
    seq = common.fixed_schedule(sp.graph("dist"), sp.platform)
    sp.fill()
    comm.barrier()
    sp.run(seq, 3)
    comm.barrier()
    assert sp.check(), ("spmv dist", rank, bw)
    comm.barrier()
# and a search whose every candidate is executed by both processes
res = halo.search("mcts", style="fused_mixed", iters=5, benchmarker=_C.EmpiricalBenchmarker(),
                  bench=common.bench_opts(n_iters=2, target_secs=1e-5, max_retries=0,
                                          check_randomness=False))
assert len(res.sim_results) == 5
dist.barrier()
if rank == 0:
    print("GLOO_WORKER_OK", flush=True)
dist.destroy_process_group()
