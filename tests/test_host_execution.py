"""The workloads EXECUTED on the CPU (tenzing/host_exec.hpp): host-memory symmetric heap, Null
platform with host execution on, ranks as threads of one process.  Payload really moves into the
neighbour's memory, epoch flags are published with release and polled with acquire semantics,
staging is double buffered by epoch parity - the device protocol, minus the device.  Every
implementation style, several searched schedules per style, several epochs per schedule, checked
against the same oracles as the GPU tests."""
import random

import pytest
import torch

import tenzing_b200 as tz
from tenzing_b200 import _C
from tenzing_b200.models import DistributedSpMV, HaloExchange, common
from tenzing_b200.models import halo as halo_mod
from tenzing_b200.models import spmv as spmv_mod


def _schedules(wl, style, n, seed):
    """the fixed schedule plus up to n-1 random complete walks through the decision tree"""
    g = wl.graph(style)
    seqs = [common.fixed_schedule(g, wl.platform)]
    rng = random.Random(seed)          # same seed on every rank: same walks
    for _ in range(n - 1):
        st = tz.State(g)
        while True:
            ds = st.get_decisions(wl.platform)
            if not ds:
                break
            st = st.apply(rng.choice(ds))
        seq = st.sequence()
        _C.remove_redundant_syncs(seq)
        seqs.append(seq)
    return seqs


def _run_world(size, body):
    _C.set_host_wait_timeout(10.0)
    errors = []

    def guarded(rank, comm):
        try:
            body(rank, comm)
        except BaseException as e:     # noqa: BLE001 - reported on the main thread
            errors.append((rank, repr(e)))
            raise

    try:
        _C.loopback_run(size, guarded)
    finally:
        _C.set_host_wait_timeout(20.0)
    assert not errors, errors


@pytest.mark.timeout(300)
@pytest.mark.parametrize("size,grid", [(1, None), (2, None), (4, None), (8, None), (6, (3, 2, 1))])
def test_halo_exchange_executes_on_the_cpu_in_every_style(size, grid):
    def body(rank, comm):
        h = HaloExchange(comm=comm, nX=6, nY=5, nZ=4, nQ=2, nGhost=2, host=True, rank_grid=grid)
        h.platform.set_host_execution(True)
        for style in halo_mod.STYLES:
            for k, seq in enumerate(_schedules(h, style, 3, seed=11)):
                h.fill()
                comm.barrier()
                h.run(seq, 3)          # three epochs: both staging parities, flags keep growing
                comm.barrier()
                assert h.check(), (style, k, rank)
                comm.barrier()

    _run_world(size, body)


@pytest.mark.timeout(300)
@pytest.mark.parametrize("neighbors,order", [(6, "XYZQ"), (18, "QXYZ"), (26, "QXYZ")])
def test_halo_variants_execute_on_the_cpu(neighbors, order):
    def body(rank, comm):
        h = HaloExchange(comm=comm, nX=5, nY=4, nZ=6, nQ=3, nGhost=1, neighbors=neighbors,
                         storage_order=order, dtype=torch.float32, host=True)
        h.platform.set_host_execution(True)
        for style in ("reference", "staged", "direct", "fused_mixed", "fused_pipelined", "fused_pull"):
            h.fill()
            comm.barrier()
            h.run(common.fixed_schedule(h.graph(style), h.platform), 2)
            comm.barrier()
            assert h.check(), (style, rank)
            comm.barrier()

    _run_world(4, body)


@pytest.mark.timeout(300)
@pytest.mark.parametrize("size", [1, 2, 3, 5])
def test_distributed_spmv_executes_on_the_cpu_in_every_style(size):
    def body(rank, comm):
        for bw in (None, 10 ** 9):     # band (neighbours only) and random columns (all-to-all)
            s = DistributedSpMV(comm=comm, n=403, bw=bw, nnz_per_row=7, seed=3, host=True)
            s.platform.set_host_execution(True)
            for style in spmv_mod.STYLES + spmv_mod.EXPERIMENTAL_STYLES:
                for k, seq in enumerate(_schedules(s, style, 3, seed=5)):
                    s.fill()
                    comm.barrier()
                    s.run(seq, 3)
                    comm.barrier()
                    assert s.check(), (style, k, rank, bw)
                    comm.barrier()

    _run_world(size, body)


@pytest.mark.timeout(300)
@pytest.mark.parametrize("size", [1, 4])
def test_streams_as_threads_run_the_schedules_concurrently_and_correctly(size):
    """enable_host_streams(): every platform stream is a worker thread and events behave as in
    CUDA, so ops bound to different streams really overlap unless the schedule orders them.
    Random schedules of every style must still produce the oracle's result."""
    def body(rank, comm):
        h = HaloExchange(comm=comm, nX=6, nY=5, nZ=4, nQ=2, nGhost=2, host=True)
        h.platform.enable_host_streams()
        s = DistributedSpMV(comm=comm, n=301, nnz_per_row=7, seed=3, host=True)
        s.platform.enable_host_streams()
        for wl, styles in ((h, halo_mod.STYLES), (s, spmv_mod.STYLES + spmv_mod.EXPERIMENTAL_STYLES)):
            for style in styles:
                for k, seq in enumerate(_schedules(wl, style, 3, seed=23)):
                    wl.fill()
                    comm.barrier()
                    wl.run(seq, 3)
                    wl.platform.synchronize()
                    comm.barrier()
                    assert wl.check(), (style, k, rank)
                    comm.barrier()

    _run_world(size, body)


@pytest.mark.timeout(120)
def test_the_searched_choice_graph_executes_and_a_real_search_runs_on_the_cpu():
    """MCTS over the ChoiceOp of all styles with the wall-clock benchmarker: every schedule the
    search tries is really executed by both ranks (host execution), and the winner is correct."""
    def body(rank, comm):
        h = HaloExchange(comm=comm, nX=6, nQ=1, nGhost=1, neighbors=6, host=True)
        h.platform.set_host_execution(True)
        h.fill()
        bench = common.bench_opts(n_iters=2, target_secs=1e-5, max_retries=0, check_randomness=False)
        res = h.search("mcts", style="choice", iters=12, benchmarker=_C.EmpiricalBenchmarker(),
                       bench=bench, seed=2)
        seq, _ = common.best_of(res)
        h.fill()
        comm.barrier()
        h.run(seq)
        comm.barrier()
        assert h.check()
        comm.barrier()

    _run_world(2, body)


@pytest.mark.timeout(120)
def test_searching_the_spmv_choice_graph_resolves_same_named_ops_of_different_styles():
    """Every SpMV style has ops called "yl" / "yr" / "PackPut", and they are different ops (other
    flag sets, other outputs).  A broadcast sequence names ops only by name, so the receiving
    ranks must pick the alternative the SET of names identifies - otherwise ranks execute
    different kernels and wait for flags nobody publishes.  (Found by host execution: the
    multi-rank search over the choice graph timed out in a flag wait.)"""
    def body(rank, comm):
        s = DistributedSpMV(comm=comm, n=301, nnz_per_row=7, seed=3, host=True)
        s.platform.set_host_execution(True)
        s.fill()
        bench = common.bench_opts(n_iters=2, target_secs=1e-5, max_retries=0, check_randomness=False)
        res = s.search("mcts", style="choice", iters=24, benchmarker=_C.EmpiricalBenchmarker(),
                       bench=bench, seed=4)
        styles_seen = set()
        for r in res.sim_results:
            names = set(r.path.names())
            styles_seen.add("fused_all" if "SpMVAll" in names else "reference" if "PostSend" in names
                            else "fused_separate" if "y" in names else "fused")
        assert len(styles_seen) >= 3, styles_seen       # the search did try several styles
        seq, _ = common.best_of(res)
        s.fill()
        comm.barrier()
        s.run(seq, 2)
        comm.barrier()
        assert s.check()
        comm.barrier()

    _run_world(3, body)


def test_sequence_json_round_trip_picks_the_right_alternative():
    s = DistributedSpMV(n=120, nnz_per_row=5, host=True)
    g = s.graph("choice")
    plat = s.platform
    for style in spmv_mod.STYLES:
        want = common.fixed_schedule(s.graph(style), plat)
        back = _C.sequence_from_json(want.to_json(s.graph(style)), g)    # resolved in the choice graph
        assert back.names() == want.names()
        # the resolved ops are the chosen alternative's own ops: executing them gives the answer
        plat.set_host_execution(True)
        s.fill()
        s.run(back, 2)
        assert s.check(), style
        plat.set_host_execution(False)


@pytest.mark.timeout(120)
def test_a_wait_nobody_satisfies_raises_instead_of_hanging():
    """dead-lock detection: a rank that waits for a flag its partner never publishes gets an
    exception after the time-out, and the partner is released from its barrier with one too"""
    seen = {}

    def body(rank, comm):
        h = HaloExchange(comm=comm, nX=4, nQ=1, nGhost=1, neighbors=6, host=True)
        h.platform.set_host_execution(True)
        seq = common.fixed_schedule(h.graph("fused_staged"), h.platform)
        _C.set_host_wait_timeout(0.5)
        try:
            if rank == 0:
                h.run(seq)             # rank 1 never runs: its flags never arrive
            comm.barrier()
        except RuntimeError as e:
            seen[rank] = str(e)
            raise

    with pytest.raises(RuntimeError):
        _C.loopback_run(2, body)
    _C.set_host_wait_timeout(20.0)
    assert "did not reach epoch" in seen[0]
    assert "another rank failed" in seen.get(1, "another rank failed")


@pytest.mark.timeout(120)
def test_the_host_posted_comm_op_library_moves_data_between_ranks():
    """Isend / Irecv / OwningIsend / OwningIrecv / Wait / MultiWait / OwningWaitall / Ialltoallv
    (reference: mpi/ops_mpi.hpp, ops_halo_exchange.hpp:68-92) executed on host memory: three
    thread ranks, a ring exchange with the plain ops, one with the owning ops, and a personalised
    all-to-all, scheduled by the solver and repeated over several epochs."""
    import ctypes
    import numpy as np

    n_ranks, n = 3, 16

    def body(rank, comm):
        heap = _C.SymmetricHeap(comm, 1 << 16, True)
        off_src, off_dst = heap.alloc(n * 8), heap.alloc(n * 8)
        off_dst2 = heap.alloc(n * 8)
        off_a2a_src, off_a2a_dst = heap.alloc(n_ranks * 8), heap.alloc(n_ranks * 8)
        off_flags = heap.alloc(8 * 8)

        def view(r, off, count):
            buf = (ctypes.c_double * count).from_address(heap.peer_ptr(r) + off)
            return np.frombuffer(buf, dtype=np.float64)

        right, left = (rank + 1) % n_ranks, (rank - 1) % n_ranks
        tr = _C.P2PTransport(True)
        plat = tz.Platform.make_null(1, comm)
        plat.set_host_execution(True)
        sreq, rreq = _C.Request(), _C.Request()
        isend = _C.Isend(heap.local_ptr() + off_src, n * 8, heap.peer_ptr(right) + off_dst,
                         heap.peer_ptr(right) + off_flags, sreq, tr, "isend")
        irecv = _C.Irecv(heap.local_ptr() + off_flags, rreq, tr, "irecv")
        wait = _C.Wait(sreq, "wait_send")
        multi = _C.MultiWait([rreq], "wait_recv")
        oisend, osreq = _C.OwningIsend(heap.local_ptr() + off_src, n * 8, heap.peer_ptr(left) + off_dst2,
                                       heap.peer_ptr(left) + off_flags + 8, tr, "owning_isend")
        oirecv, orreq = _C.OwningIrecv(heap.local_ptr() + off_flags + 8, tr, "owning_irecv")
        owait = _C.OwningWaitall([osreq, orreq], "wait_owned")
        a2areq = _C.Request()
        a2a = _C.Ialltoallv(heap.local_ptr() + off_a2a_src, [8 * p for p in range(n_ranks)],
                            [8] * n_ranks, [heap.peer_ptr(p) + off_a2a_dst + 8 * rank for p in range(n_ranks)],
                            [heap.peer_ptr(p) + off_flags + 16 + 8 * rank for p in range(n_ranks)],
                            a2areq, tr, "a2a")
        a2a_recv = [_C.Irecv(heap.local_ptr() + off_flags + 16 + 8 * p, _C.Request(), tr, f"a2a_irecv{p}")
                    for p in range(n_ranks)]
        g = tz.Graph()
        posts = [isend, irecv, oisend, oirecv, a2a] + a2a_recv
        waits = [wait, multi, owait]
        for p in posts:
            g.start_then(p)
            for w in waits:
                g.then(p, w)
        for w in waits:
            g.then_finish(w)
        seqs = tz.dfs.get_all_sequences(g, plat, 6, True)
        assert seqs
        for k, seq in enumerate(seqs[:3]):
            for epoch in range(3):
                view(rank, off_src, n)[:] = 100 * rank + k + epoch + np.arange(n)
                view(rank, off_a2a_src, n_ranks)[:] = [1000 * rank + p + epoch for p in range(n_ranks)]
                comm.barrier()
                seq.run(plat)
                comm.barrier()          # everybody's sends have landed (waits only cover my own)
                assert np.array_equal(view(rank, off_dst, n), 100 * left + k + epoch + np.arange(n))
                assert np.array_equal(view(rank, off_dst2, n), 100 * right + k + epoch + np.arange(n))
                assert list(view(rank, off_a2a_dst, n_ranks)) == [1000 * p + rank + epoch for p in range(n_ranks)]
                comm.barrier()

    _run_world(n_ranks, body)


@pytest.mark.timeout(300)
@pytest.mark.parametrize("size", [2, 4, 8])
def test_consecutive_exchanges_without_any_barrier_between_them(size):
    """An application alternates exchange and compute with no rank-wide barrier in between, so a
    fast rank starts exchange e+1 while a slow neighbour is still reading the ghost cells of
    exchange e.  Interior values change every epoch and every rank checks its ghosts right after
    each exchange: staged styles rely on their two staging buffers, the styles that put straight
    into ghost cells on the permission flags (TzbBox::ready) - a put that arrives an epoch early
    shows up as a wrong ghost value here."""
    import time

    def body(rank, comm):
        h = HaloExchange(comm=comm, nX=5, nY=4, nZ=3, nQ=2, nGhost=2, host=True)
        h.platform.set_host_execution(True)
        rng = random.Random(1000 + rank)       # different delays on every rank
        g = h.args.nGhost
        for style in halo_mod.STYLES:
            seq = common.fixed_schedule(h.graph(style), h.platform)
            h.platform.provision_events(seq)
            h.fill()
            base = h.expected()
            comm.barrier()                      # only between styles
            for epoch in range(1, 7):
                interior = h.cells()[:, g:-g, g:-g, g:-g]
                interior += 1.0e6               # "compute": every interior cell changes
                seq.run(h.platform)
                # a slow consumer: the neighbours may long be in their next exchange by now
                time.sleep(rng.choice((0.0, 0.0, 0.006)))
                # "consume": every cell, ghosts included, must carry this epoch's values
                assert torch.equal(h.cells(), base + 1.0e6 * epoch), (style, epoch, rank)
            comm.barrier()

    _run_world(size, body)
