#!/usr/bin/env python
"""Device time of every implementation style (fixed single-stream schedule) of both workloads.

Run alone (1 GPU) or under torch.distributed.run (N GPUs).  The `reference` rows are the
reference's op structure (pack / isend / wait / unpack; scatter / send / recv / product / add)
with copy-engine peer copies standing in for CUDA-aware MPI - the baseline the fused styles
are compared with.  Times: median of --reps steps, CUDA events, L2 flushed, max over ranks."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")
import torch  # noqa: E402

from tenzing_b200 import _C  # noqa: E402
from tenzing_b200.parallel import init_process_group_from_env, make_comm  # noqa: E402
from tenzing_b200.models import HaloExchange, DistributedSpMV, common  # noqa: E402
from tenzing_b200.models.halo import STYLES as HALO_STYLES  # noqa: E402
from tenzing_b200.models.spmv import STYLES as SPMV_STYLES  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=512)
    ap.add_argument("--nq", type=int, default=3)
    ap.add_argument("--spmv-n", type=int, default=1_000_000)
    ap.add_argument("--spmv-random-n", type=int, default=0, help="also run a random matrix of this size")
    ap.add_argument("--reps", type=int, default=15)
    ap.add_argument("--streams", type=int, default=2)
    ap.add_argument("--out", default="")
    ap.add_argument("--search-iters", type=int, default=0,
                    help="also report the best schedule of an MCTS (FastMin) search per halo style")
    ap.add_argument("--halo-styles", default=",".join(HALO_STYLES),
                    help='comma separated; add "dist" for the torch.distributed (NCCL) baseline')
    ap.add_argument("--hand", action="store_true", help="hand-written fused_groups pipelines + put-only times")
    ap.add_argument("--no-spmv", action="store_true")
    ap.add_argument("--mixed-hand", action="store_true", help="hand schedules over the fused_mixed ops")
    ap.add_argument("--rank-grid", default="", help="e.g. 1,2,1")
    args = ap.parse_args()
    rank, world, _ = init_process_group_from_env()
    comm = make_comm()
    dev = torch.device("cuda", torch.cuda.current_device())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []

    def measure(wl, style):
        seq = common.fixed_schedule(wl.graph(style), wl.platform)
        wl.fill()
        comm.barrier()
        wl.run(seq)
        wl.sync()
        comm.barrier()
        ok = wl.check()
        wl.platform.provision_events(seq)
        _C.device_time_steps(seq, wl.platform, 3, flush.data_ptr(), flush.numel())
        t = torch.tensor(_C.device_time_steps(seq, wl.platform, args.reps, flush.data_ptr(), flush.numel()),
                         dtype=torch.float64, device=dev)
        if world > 1:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return float(t.median()) * 1e6, ok, len(seq)

    rg = [int(v) for v in args.rank_grid.split(",")] if args.rank_grid else None
    halo = HaloExchange(comm=comm, nX=args.nx, nQ=args.nq, nGhost=3, n_streams=args.streams, rank_grid=rg)
    def measure_seq(wl, seq):
        wl.fill()
        comm.barrier()
        wl.run(seq)
        wl.sync()
        comm.barrier()
        ok = wl.check()
        wl.platform.provision_events(seq)
        _C.device_time_steps(seq, wl.platform, 3, flush.data_ptr(), flush.numel())
        t = torch.tensor(_C.device_time_steps(seq, wl.platform, args.reps, flush.data_ptr(), flush.numel()),
                         dtype=torch.float64, device=dev)
        if world > 1:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return float(t.median()) * 1e6, ok

    for style in args.halo_styles.split(","):
        us, ok, n = measure(halo, style)
        rows.append(dict(workload=f"halo {args.nx}^3 nQ={args.nq}", style=style, us=round(us, 2), ok=ok,
                         ops=n, n_gpus=world, bytes_out=halo.bytes_per_exchange()))
        if args.search_iters and style in ("fused_groups", "staged", "fused_staged", "fused_mixed", "fused_pipelined"):
            res = halo.search("mcts", style=style, iters=args.search_iters, strategy="FastMin",
                              bench=common.bench_opts(n_iters=3, target_secs=0.002, max_retries=0,
                                                      check_randomness=False))
            seq, _ = common.best_of(res)
            us, ok = measure_seq(halo, seq)
            rows.append(dict(workload=f"halo {args.nx}^3 nQ={args.nq}", style=style + "+mcts",
                             us=round(us, 2), ok=ok, ops=len(seq), n_gpus=world,
                             schedule=seq.desc(), bytes_out=halo.bytes_per_exchange()))
    if args.hand:
        # hand-written pipelines over the fused_groups ops: puts back to back on stream 0 (each
        # gets the whole NVLink), unpacks on stream 1 as soon as their group has been put
        g = halo.graph("flat:fused_groups")

        def pipeline(order):
            ops = [g.start()]
            ev = 0
            for grp in order:
                ops.append(_C.BoundGpuOp(g.find_name("he_packput_" + grp), 0))
                ops.append(_C.CudaEventRecord(ev, 0))
                ev += 1
            for i, grp in enumerate(order):
                ops.append(_C.CudaStreamWaitEvent(1, i))
                ops.append(_C.BoundGpuOp(g.find_name("he_waitunpack_" + grp), 1))
            ops += [_C.CudaEventRecord(ev, 1), _C.CudaEventSync(ev), g.finish()]
            return _C.Sequence(ops)

        for order in ("xyz", "zyx", "yxz"):
            seq = pipeline(order)
            us, ok = measure_seq(halo, seq)
            rows.append(dict(workload=f"halo {args.nx}^3 nQ={args.nq}", style="groups pipeline " + order,
                             us=round(us, 2), ok=ok, ops=len(seq), n_gpus=world))
        # put kernels alone (no waits): how long does the NVLink side take?
        for name in ("he_packput_x", "he_packput_y", "he_packput_z"):
            seq = _C.Sequence([_C.BoundGpuOp(g.find_name(name), 0)])
            t = torch.tensor(_C.device_time_steps(seq, halo.platform, args.reps, flush.data_ptr(), flush.numel()),
                             dtype=torch.float64, device=dev)
            if world > 1:
                torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
            rows.append(dict(workload=f"halo {args.nx}^3 nQ={args.nq}", style=name + " alone",
                             us=round(float(t.median()) * 1e6, 2), ok=True, ops=1, n_gpus=world))
        g2 = halo.graph("flat:fused_staged")
        seq = _C.Sequence([_C.BoundGpuOp(g2.find_name("he_packput_all"), 0)])
        t = torch.tensor(_C.device_time_steps(seq, halo.platform, args.reps, flush.data_ptr(), flush.numel()),
                         dtype=torch.float64, device=dev)
        if world > 1:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        rows.append(dict(workload=f"halo {args.nx}^3 nQ={args.nq}", style="he_packput_all alone",
                         us=round(float(t.median()) * 1e6, 2), ok=True, ops=1, n_gpus=world))
    if args.mixed_hand:
        g = halo.graph("flat:fused_mixed")
        B = lambda n, st: _C.BoundGpuOp(g.find_name(n), st)  # noqa: E731
        hands = {
            "thin first, unpack beside thick put": [
                g.start(), B("he_packput_thin", 0), _C.CudaEventRecord(0, 0), B("he_put_thick", 0),
                _C.CudaStreamWaitEvent(1, 0), B("he_waitunpack_thin", 1), B("he_wait_thick", 0),
                _C.CudaEventRecord(1, 0), _C.CudaEventRecord(2, 1), _C.CudaEventSync(1),
                _C.CudaEventSync(2), g.finish()],
            "both puts concurrent (DFS best)": [
                g.start(), B("he_put_thick", 0), B("he_packput_thin", 1), B("he_wait_thick", 0),
                B("he_waitunpack_thin", 1), _C.CudaEventRecord(1, 0), _C.CudaEventRecord(2, 1),
                _C.CudaEventSync(1), _C.CudaEventSync(2), g.finish()],
        }
        for label, ops in hands.items():
            us, ok = measure_seq(halo, _C.Sequence(ops))
            rows.append(dict(workload=f"halo {args.nx}^3 nQ={args.nq}", style="mixed: " + label,
                             us=round(us, 2), ok=ok, ops=len(ops), n_gpus=world,
                             items_per_block=os.environ.get("TZB_ITEMS_PER_BLOCK", "resident")))
    del halo
    torch.cuda.empty_cache()
    cases = [] if args.no_spmv else [("spmv band", args.spmv_n, None)]
    if args.spmv_random_n:
        cases.append(("spmv random", args.spmv_random_n, args.spmv_random_n))
    for label, n, bw in cases:
        s = DistributedSpMV(comm=comm, n=n, bw=bw, nnz_per_row=10, n_streams=args.streams)
        for style in SPMV_STYLES:
            us, ok, nops = measure(s, style)
            rows.append(dict(workload=f"{label} n={n}", style=style, us=round(us, 2), ok=ok, ops=nops,
                             n_gpus=world, hbm_bytes=s.hbm_bytes(), recv_bytes=s.recv_bytes()))
        del s
    if rank == 0:
        for r in rows:
            print(f"{r['workload']:26s} {r['style']:18s} {r['us']:10.2f} us  ok={r['ok']}  ops={r['ops']}")
            if "schedule" in r:
                print("      " + r["schedule"])
        if args.out:
            json.dump(rows, open(args.out, "w"), indent=1)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
