"""No GPU needed: execute both workloads on the CPU with four ranks as threads of this process.

The grid, staging buffers and flags live in host memory; every op runs its host restatement
(tenzing/host_exec.hpp): payload is stored into the neighbour's memory, flags are published with
release and polled with acquire semantics.  With enable_host_streams() every platform stream is a
worker thread and cudaEventRecord / cudaStreamWaitEvent / cudaEventSynchronize keep their CUDA
meaning, so a schedule that lacks a synchronisation really races (scripts/sanitize.sh runs such
executions under ThreadSanitizer).  A flag that nobody publishes makes the wait throw after a
time-out instead of hanging."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")

from tenzing_b200 import _C  # noqa: E402
from tenzing_b200.models import DistributedSpMV, HaloExchange, common  # noqa: E402

_C.set_host_wait_timeout(10.0)
report = {}


def body(rank, comm):
    halo = HaloExchange(comm=comm, nX=16, nQ=2, nGhost=2, host=True)
    halo.platform.enable_host_streams()
    halo.fill()
    bench = common.bench_opts(n_iters=3, target_secs=1e-4, max_retries=0, check_randomness=False)
    # a real search: every candidate schedule is executed by all four ranks and timed (wall clock)
    res = halo.search("mcts", style="choice", iters=20, benchmarker=_C.EmpiricalBenchmarker(), bench=bench)
    seq, stats = common.best_of(res)
    halo.fill()
    comm.barrier()
    halo.run(seq)
    halo.platform.synchronize()
    comm.barrier()
    ok_halo = halo.check()

    spmv = DistributedSpMV(comm=comm, n=20_000, nnz_per_row=10, host=True)
    spmv.platform.set_host_execution(True)
    s_res = spmv.search("dfs", style="fused", benchmarker=_C.EmpiricalBenchmarker(), bench=bench)
    s_seq, s_stats = common.best_of(s_res)
    spmv.fill()
    comm.barrier()
    spmv.run(s_seq)
    comm.barrier()
    report[rank] = (ok_halo, spmv.check(), stats.pct10, seq.desc(), s_stats.pct10, s_seq.desc())
    comm.barrier()


_C.loopback_run(4, body)
assert all(r[0] and r[1] for r in report.values()), report
ok_h, ok_s, t_h, d_h, t_s, d_s = report[0]
print(f"halo  best of 20 MCTS iterations: {t_h * 1e6:8.1f} us  {d_h}")
print(f"spmv  best of the DFS           : {t_s * 1e6:8.1f} us  {d_s}")
print("results match the oracles on all 4 ranks")
