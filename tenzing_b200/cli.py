"""Command-line front end mirroring the native drivers (reference: tenzing-dfs/examples/spmv.cu,
tenzing-mcts/examples/{spmv,halo}_*.cu):

  python -m tenzing_b200.cli spmv --solver dfs  -m 150000 -b 50
  python -m tenzing_b200.cli spmv --solver mcts --strategy Coverage -i 300 -b 50
  python -m tenzing_b200.cli halo --strategy FastMin -i 300 -b 50 -n 512
(under torch.distributed.run for several GPUs).  Prints the reproducibility header and the
results CSV on rank 0, like the reference's executables.
"""
import argparse
import sys

import torch

from . import _C
from .models import DistributedSpMV, HaloExchange, common
from .parallel import init_process_group_from_env, make_comm


def main(argv=None):
    ap = argparse.ArgumentParser(prog="tenzing_b200.cli")
    ap.add_argument("workload", choices=["spmv", "halo"])
    ap.add_argument("--solver", choices=["dfs", "mcts"], default="mcts")
    ap.add_argument("--strategy", "-s", default="FastMin", choices=_C.mcts.strategy_names())
    ap.add_argument("--mcts-iters", "-i", type=int, default=300)
    ap.add_argument("--benchmark-iters", "-b", type=int, default=50)
    ap.add_argument("--max-seqs", type=int, default=15000)
    ap.add_argument("--matrix-m", "-m", type=int, default=150000)
    ap.add_argument("--random", action="store_true", help="uniformly random columns (bw = m)")
    ap.add_argument("--matrix-file", default=None, help="MatrixMarket file instead of a synthetic matrix")
    ap.add_argument("--fp32", action="store_true")
    ap.add_argument("--n", "-n", type=int, default=512, help="halo: interior cells per dimension per rank")
    ap.add_argument("--nq", type=int, default=3)
    ap.add_argument("--ghost", type=int, default=3)
    ap.add_argument("--neighbors", type=int, default=26, choices=[6, 18, 26])
    ap.add_argument("--streams", type=int, default=2)
    ap.add_argument("--style", default="choice")
    ap.add_argument("--no-expand-rollout", action="store_true")
    ap.add_argument("--null", action="store_true", help="no GPU: enumerate/search on ids only")
    ap.add_argument("--host-exec", action="store_true",
                    help="no GPU, but execute the ops for real on the CPU (host memory)")
    ap.add_argument("--host-streams", action="store_true",
                    help="with --host-exec: one worker thread per stream, CUDA event semantics")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--resume", default=None, metavar="CSV",
                    help="results CSV of an interrupted run of the same command: with the same "
                         "--seed the search replays it from the file and continues measuring "
                         "where it stopped (every rank must see the same file)")
    ap.add_argument("--cache", action="store_true",
                    help="never measure the same schedule twice in this run")
    args = ap.parse_args(argv)

    _C.init()
    rank, world, _ = init_process_group_from_env()
    comm = make_comm()
    if rank == 0:
        print(_C.reproduce_header([sys.argv[0]] + (argv if argv is not None else sys.argv[1:])))
    host = args.null or args.host_exec or args.host_streams or not torch.cuda.is_available()
    if args.workload == "spmv":
        wl = DistributedSpMV(comm=comm, n=args.matrix_m, bw=args.matrix_m if args.random else None,
                             dtype=torch.float32 if args.fp32 else torch.float64, host=host,
                             n_streams=args.streams, matrix_file=args.matrix_file)
    else:
        wl = HaloExchange(comm=comm, nX=args.n, nQ=args.nq, nGhost=args.ghost,
                          neighbors=args.neighbors, host=host, n_streams=args.streams)
    if args.host_streams:
        wl.platform.enable_host_streams()
    elif args.host_exec:
        wl.platform.set_host_execution(True)
    bench = common.bench_opts(n_iters=args.benchmark_iters)
    benchmarker = _C.EmpiricalBenchmarker() if host else _C.DeviceBenchmarker()
    if host:
        bench = common.bench_opts(n_iters=min(args.benchmark_iters, 3), target_secs=1e-5,
                                  check_randomness=False)
    if args.solver == "dfs":
        res = wl.search("dfs", style=args.style, benchmarker=benchmarker, bench=bench,
                        max_seqs=args.max_seqs, resume_csv=args.resume, cache=args.cache)
    else:
        res = wl.search("mcts", style=args.style, benchmarker=benchmarker, bench=bench,
                        iters=args.mcts_iters, strategy=args.strategy,
                        expand_rollout=not args.no_expand_rollout, seed=args.seed,
                        resume_csv=args.resume, cache=args.cache)
    if rank == 0:
        sys.stdout.write(res.csv())
    if world > 1:
        torch.distributed.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
