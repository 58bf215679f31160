"""Worker run under torch.distributed.run by test_gpu_multi.py: one rank per GPU.

Checks, on every rank, that each implementation style of both workloads produces the oracle
result when the neighbours are OTHER GPUs (peer stores + device flags over NVLink), for several
consecutive epochs (both staging parities), then runs a short search."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from tenzing_b200.parallel import init_process_group_from_env, make_comm  # noqa: E402
from tenzing_b200.models import HaloExchange, DistributedSpMV, common  # noqa: E402


def all_ok(ok):
    t = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(t.item() > 0.5)




def main():
    rank, world, _ = init_process_group_from_env()
    comm = make_comm()
    failures = []
    halo_styles = ("reference", "staged", "direct", "fused_all", "fused_staged", "fused_groups",
                   "fused_mixed", "fused_pipelined", "fused_pull", "dist")
    from tenzing_b200.ops import nvls
    have_nvls = nvls.available()
    if rank == 0:
        print("NVLS multicast available:", have_nvls, flush=True)
    spmv_styles = ("reference", "fused", "fused_separate", "fused_all", "fused_pull", "dist") + \
        (("nvls",) if have_nvls else ())
    for order in ("XYZQ", "QXYZ"):
        h = HaloExchange(comm=comm, nX=24, nY=10, nZ=12, nQ=2, nGhost=3, storage_order=order)
        for style in halo_styles:
            seq = common.fixed_schedule(h.graph(style), h.platform)
            if rank == 0 and os.environ.get("TZ_WORKER_VERBOSE"):
                print("halo", order, style, flush=True)
            for _ in range(3):
                h.fill()
                comm.barrier()
                h.run(seq)
                h.sync()
                comm.barrier()
                if not all_ok(h.check()):
                    failures.append(f"halo {order} {style}")
                    break
    for dtype in (torch.float64, torch.float32):
        for n, bw in ((40_000, None), (30_011, 30_011)):  # band (neighbour peers) and random (all peers)
            s = DistributedSpMV(comm=comm, n=n, bw=bw, nnz_per_row=10, dtype=dtype)
            for style in spmv_styles:
                seq = common.fixed_schedule(s.graph(style), s.platform)
                if rank == 0 and os.environ.get("TZ_WORKER_VERBOSE"):
                    print("spmv", dtype, n, style, flush=True)
                for _ in range(3):
                    s.fill()
                    comm.barrier()
                    s.run(seq)
                    s.sync()
                    comm.barrier()
                    if not all_ok(s.check()):
                        failures.append(f"spmv {dtype} n={n} {style}")
                        break
    # the fused schedules replayed as CUDA graphs across ranks (device-side epoch)
    from tenzing_b200 import _C
    for wl, style in ((HaloExchange(comm=comm, nX=24, nY=10, nZ=12, nQ=2, nGhost=3), "fused_mixed"),
                      (DistributedSpMV(comm=comm, n=40_000, nnz_per_row=10), "fused_all"),
                      (DistributedSpMV(comm=comm, n=30_011, bw=30_011, nnz_per_row=10), "fused")):
        seq = common.fixed_schedule(wl.graph(style), wl.platform)
        wl.platform.provision_events(seq)
        wl.fill()
        comm.barrier()
        wl.run(seq)
        wl.sync()
        comm.barrier()
        replay = _C.GraphReplay(seq, wl.platform)
        for _ in range(2):
            wl.fill()
            comm.barrier()
            for _ in range(3):
                replay.launch()
            wl.sync()
            comm.barrier()
            if not all_ok(wl.check()):
                failures.append(f"graph replay {style}")
                break
        del replay
    # Ialltoallv between real GPUs: rank r sends the 16 values 1000 r + p + (0..15) to every peer p
    from tenzing_b200.utils.tensor import device_tensor
    heap = _C.SymmetricHeap(comm, 1 << 20)
    k = 16
    off_src, off_dst, off_flags = heap.alloc(k * 8 * world), heap.alloc(k * 8 * world), heap.alloc(8 * world)
    src = device_tensor(heap.local_ptr() + off_src, (world, k), torch.float64, owner=heap)
    dst = device_tensor(heap.local_ptr() + off_dst, (world, k), torch.float64, owner=heap)
    tr = _C.P2PTransport()
    plat = _C.Platform.make_n_streams(1, comm, False)
    sreq = _C.Request()
    a2a = _C.Ialltoallv(src.data_ptr(), [k * 8 * p for p in range(world)], [k * 8] * world,
                        [heap.peer_ptr(p) + off_dst + k * 8 * rank for p in range(world)],
                        [heap.peer_ptr(p) + off_flags + 8 * rank for p in range(world)], sreq, tr, "a2a")
    rreqs = [_C.Request() for _ in range(world)]
    irecvs = [_C.Irecv(heap.local_ptr() + off_flags + 8 * p, rreqs[p], tr, f"a2a_irecv{p}") for p in range(world)]
    g = _C.Graph()
    waits = [_C.Wait(sreq, "a2a_wait_send"), _C.MultiWait(rreqs, "a2a_wait_recv")]
    for p_ in [a2a] + irecvs:
        g.start_then(p_)
        for w in waits:
            g.then(p_, w)
    for w in waits:
        g.then_finish(w)
    seq = _C.dfs.get_all_sequences(g, plat, 2, True)[0]
    for epoch in range(3):
        src.copy_((1000.0 * rank + torch.arange(world, device="cuda", dtype=torch.float64)[:, None]
                   + torch.arange(k, device="cuda", dtype=torch.float64)[None, :]) + 0.5 * epoch)
        dst.zero_()
        torch.cuda.synchronize()
        comm.barrier()
        seq.run(plat)
        plat.synchronize()
        torch.cuda.synchronize()
        comm.barrier()
        exp = (1000.0 * torch.arange(world, device="cuda", dtype=torch.float64)[:, None] + rank
               + torch.arange(k, device="cuda", dtype=torch.float64)[None, :]) + 0.5 * epoch
        if not all_ok(bool(torch.equal(dst, exp))):
            failures.append("Ialltoallv on devices")
            break
    # a short search across ranks: rank 0 drives, all ranks execute in lock step
    h = HaloExchange(comm=comm, nX=32, nQ=1, nGhost=2, neighbors=6)
    res = h.search("mcts", style="choice", iters=8,
                   bench=common.bench_opts(n_iters=2, target_secs=0.0005, check_randomness=False))
    seq, br = common.best_of(res)
    h.fill()
    comm.barrier()
    h.run(seq)
    h.sync()
    comm.barrier()
    if not all_ok(h.check()):
        failures.append("halo after mcts")
    s = DistributedSpMV(comm=comm, n=50_000, nnz_per_row=10)
    res = s.search("dfs", style="fused",
                   bench=common.bench_opts(n_iters=2, target_secs=0.0005, check_randomness=False))
    seq, br = common.best_of(res)
    s.fill()
    comm.barrier()
    s.run(seq)
    s.sync()
    comm.barrier()
    if not all_ok(s.check()):
        failures.append("spmv after dfs")
    if rank == 0:
        print("MULTI_GPU_RESULT", "OK" if not failures else "FAIL " + "; ".join(failures), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 1 if failures else 0


if __name__ == "__main__":
    sys.exit(main())
