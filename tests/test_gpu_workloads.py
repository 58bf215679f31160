"""Single-GPU end-to-end checks of both workloads: every implementation style must
produce exactly the oracle result, for both epoch parities (double buffers)."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("order", ["XYZQ", "QXYZ"])
@pytest.mark.parametrize("style", ["reference", "staged", "direct", "fused_all", "fused_staged", "fused_groups", "fused_mixed", "fused_pipelined", "fused_pull"])
def test_halo_styles_single_rank_periodic(style, order):
    from tenzing_b200.models import HaloExchange, common
    h = HaloExchange(nX=20, nY=12, nZ=9, nQ=3, nGhost=3, storage_order=order)
    seq = common.fixed_schedule(h.graph(style), h.platform)
    for _ in range(3):  # odd and even epochs
        h.fill()
        assert not h.check()
        h.run(seq)
        assert h.check(), f"{style}/{order} wrong after exchange"


@pytest.mark.parametrize("neighbors", [6, 18])
def test_halo_fewer_neighbors(neighbors):
    from tenzing_b200.models import HaloExchange, common
    h = HaloExchange(nX=16, nQ=2, nGhost=2, neighbors=neighbors)
    seq = common.fixed_schedule(h.graph("staged"), h.platform)
    h.fill()
    h.run(seq)
    assert h.check()


def test_halo_fp32_and_thick_rows():
    from tenzing_b200.models import HaloExchange, common
    h = HaloExchange(nX=200, nY=10, nZ=10, nQ=1, nGhost=3, dtype=torch.float32)
    seq = common.fixed_schedule(h.graph("fused_staged"), h.platform)
    h.fill()
    h.run(seq)
    assert h.check()


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("style", ["reference", "fused", "fused_separate", "fused_all"])
def test_spmv_styles_single_rank(style, dtype):
    from tenzing_b200.models import DistributedSpMV, common
    s = DistributedSpMV(n=20011, nnz_per_row=10, dtype=dtype)
    seq = common.fixed_schedule(s.graph(style), s.platform)
    for _ in range(2):
        s.fill()
        s.run(seq)
        assert s.check(), f"{style} wrong"


def test_gpu_graph_stream_assignment_equivalence():
    """reference test/test_gpu_graph.cu:41-118"""
    import tenzing_b200 as tz
    from tenzing_b200 import _C
    k1, k2, k3 = (tz.EmptyGpuOp(n) for n in ("k1", "k2", "k3"))
    g = tz.Graph()
    g.start_then(k1)
    g.then(k1, k2)
    g.then(k1, k3)
    g.then_finish(k2)
    g.then_finish(k3)
    plat = tz.Platform.make_n_streams(2)
    st = tz.State(g)
    ds = st.get_decisions(plat)
    assigns = [d for d in ds if d.kind() == "AssignOpStream"]
    assert sorted(d.stream for d in assigns) == [0, 1] and len(ds) == 2
    s0, s1 = (st.apply(d) for d in assigns)
    assert bool(_C.state_equivalence(s0, s1))
    assert bool(_C.graph_equivalence(s0.graph(), s1.graph()))
    assert s0.graph().contains(tz.BoundGpuOp(k1, assigns[0].stream))
    execs = [d for d in s0.get_decisions(plat) if d.kind() == "ExecuteOp"]
    assert len(execs) == 1 and execs[0].op.name() == "k1"


def test_dfs_search_on_device_finds_valid_best():
    from tenzing_b200.models import DistributedSpMV, common
    s = DistributedSpMV(n=50000, nnz_per_row=10)
    res = s.search("dfs", style="fused", bench=common.bench_opts(n_iters=3, target_secs=0.001,
                                                                 check_randomness=False))
    assert len(res.sim_results) == 8
    seq, br = common.best_of(res)
    assert br.pct10 > 0
    s.fill()
    s.run(seq)
    assert s.check()
    assert res.csv().splitlines()[0].startswith('{"dfs__Opts"')


def test_mcts_search_on_device_halo():
    from tenzing_b200.models import HaloExchange, common
    h = HaloExchange(nX=32, nQ=1, nGhost=2, neighbors=6)
    res = h.search("mcts", style="choice", iters=12,
                   bench=common.bench_opts(n_iters=3, target_secs=0.001, check_randomness=False))
    assert len(res.sim_results) == 12
    seq, br = common.best_of(res)
    h.fill()
    h.run(seq)
    assert h.check()


def test_host_posted_p2p_ops_in_a_searched_graph():
    """Isend / Irecv / Wait (reference: ops_mpi.hpp) over the peer-memory transport: a producer
    kernel fills a buffer, the host posts the send to "the peer" (this rank itself), the receive
    completes through the flag, a consumer kernel reads the landed data.  The solver must insert
    the GPU->CPU sync before the Isend and nothing after the Wait."""
    import tenzing_b200 as tz
    from tenzing_b200 import _C
    from tenzing_b200.utils.tensor import device_tensor
    comm = _C.SelfComm()
    heap = _C.SymmetricHeap(comm, 1 << 20)
    n = 4096
    src = device_tensor(heap.local_ptr() + heap.alloc(n * 8), (n,), torch.float64, owner=heap)
    dst = device_tensor(heap.peer_ptr(0) + heap.alloc(n * 8), (n,), torch.float64, owner=heap)
    flag_off = heap.alloc(8)
    out = torch.zeros(n, dtype=torch.float64, device="cuda")
    tr = _C.P2PTransport()
    sreq, rreq = _C.Request(), _C.Request()
    plat = tz.Platform.make_n_streams(2)

    def produce(stream):
        with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
            src.copy_(torch.arange(n, dtype=torch.float64, device="cuda") + float(plat.epoch()))

    def consume(stream):
        with torch.cuda.stream(torch.cuda.ExternalStream(stream)):
            out.copy_(dst * 2)

    fill = tz.PyGpuOp("fill", produce)
    use = tz.PyGpuOp("use", consume)
    isend = _C.Isend(src.data_ptr(), n * 8, dst.data_ptr(), heap.peer_ptr(0) + flag_off, sreq, tr, "isend")
    irecv = _C.Irecv(heap.local_ptr() + flag_off, rreq, tr, "irecv")
    wsend, wrecv = _C.Wait(sreq, "wait_send"), _C.Wait(rreq, "wait_recv")
    g = tz.Graph()
    g.start_then(fill)
    g.then(fill, isend)
    g.start_then(irecv)
    g.then(isend, wsend)
    g.then(isend, wrecv)
    g.then(irecv, wrecv)
    g.then(irecv, wsend)
    g.then(wrecv, use)
    g.then_finish(use)
    g.then_finish(wsend)
    seqs = tz.dfs.get_all_sequences(g, plat, 50, True)
    assert seqs
    for seq in seqs[:10]:
        names = seq.names()
        i = names.index("isend")
        assert any(o.kind() == "CudaEventSync" for o in list(seq)[:i])  # fill -> isend needs the host
        plat.provision_events(seq)
        seq.run(plat)
        plat.synchronize()
        torch.cuda.synchronize()
        exp = (torch.arange(n, dtype=torch.float64, device="cuda") + float(plat.epoch())) * 2
        assert torch.equal(out, exp)


def test_bulk_copy_path_matches_the_oracle():
    """box_bulk_kernel (cp.async.bulk + mbarrier, opt-in with TZB_BULK=1: validated on B200s, not
    faster than the LSU path there) on the thick regions of the halo exchange.
    Runs in a subprocess under a hard time limit: a pipeline bug would hang the kernel."""
    import subprocess
    import sys
    code = (
        "import os, torch\n"
        "os.environ['TENZING_ACK_NOTICE'] = '1'\n"
        "from tenzing_b200.models import HaloExchange, common\n"
        "from tenzing_b200 import ops\n"
        "for nx, dtype in ((64, torch.float64), (128, torch.float32), (512, torch.float64)):\n"
        "    h = HaloExchange(nX=nx, nQ=2, nGhost=3, dtype=dtype)\n"
        "    for style in ('fused_mixed', 'fused_all', 'direct'):\n"
        "        seq = common.fixed_schedule(h.graph(style), h.platform)\n"
        "        h.fill(); h.run(seq, 3); assert h.check(), (nx, style)\n"
        "print('BULK_OK')\n")
    # same check for the other opt-in variants: 256-bit accesses on thick rows, sector path everywhere
    for env in (dict(os.environ, TZB_BULK="1"), dict(os.environ, TZB_WIDE256="1"), dict(os.environ, TZB_THIN_SECTOR="2")):
        p = subprocess.run(["timeout", "120", sys.executable, "-c", code], capture_output=True, text=True,
                           env=env, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        assert p.returncode == 0 and "BULK_OK" in p.stdout, (p.stdout + p.stderr)[-3000:]


def test_spmv_pull_style_matches_the_oracle():
    """spmv_csr_pull_kernel: the remote x entries are read from their owners inside the product.
    Subprocess under a hard time limit (a protocol bug would spin in a flag wait)."""
    import subprocess
    import sys
    code = (
        "import os, torch\n"
        "os.environ['TENZING_ACK_NOTICE'] = '1'\n"
        "from tenzing_b200.models import DistributedSpMV, common\n"
        "for n, dtype in ((5000, torch.float64), (200000, torch.float32)):\n"
        "    s = DistributedSpMV(n=n, nnz_per_row=10, dtype=dtype)\n"
        "    seq = common.fixed_schedule(s.graph('fused_pull'), s.platform)\n"
        "    s.fill(); s.run(seq, 3); assert s.check(), n\n"
        "print('PULL_OK')\n")
    p = subprocess.run(["timeout", "120", sys.executable, "-c", code], capture_output=True, text=True,
                       cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert p.returncode == 0 and "PULL_OK" in p.stdout, (p.stdout + p.stderr)[-3000:]


def test_ialltoallv_on_the_device():
    """Ialltoallv (reference: include/tenzing/mpi/ops_mpi.hpp:82-119, src/mpi/ops_mpi.cpp:29-34)
    executed on a GPU: copy-engine copies into "every peer's" receive slot (a world of one: this
    rank itself) plus one flag per peer; the receive side completes through Irecv + Wait."""
    import tenzing_b200 as tz
    from tenzing_b200 import _C
    from tenzing_b200.utils.tensor import device_tensor
    comm = _C.SelfComm()
    heap = _C.SymmetricHeap(comm, 1 << 20)
    n = 1000
    src = device_tensor(heap.local_ptr() + heap.alloc(n * 8), (n,), torch.float64, owner=heap)
    dst = device_tensor(heap.peer_ptr(0) + heap.alloc(n * 8), (n,), torch.float64, owner=heap)
    flag_off = heap.alloc(8)
    tr = _C.P2PTransport()
    plat = tz.Platform.make_n_streams(1)
    sreq, rreq = _C.Request(), _C.Request()
    a2a = _C.Ialltoallv(src.data_ptr(), [0], [n * 8], [dst.data_ptr()], [heap.peer_ptr(0) + flag_off],
                        sreq, tr, "a2a")
    irecv = _C.Irecv(heap.local_ptr() + flag_off, rreq, tr, "irecv")
    g = tz.Graph()
    wsend, wrecv = _C.Wait(sreq, "wait_send"), _C.Wait(rreq, "wait_recv")
    for p in (a2a, irecv):
        g.start_then(p)
        g.then(p, wsend)
        g.then(p, wrecv)
    g.then_finish(wsend)
    g.then_finish(wrecv)
    seq = tz.dfs.get_all_sequences(g, plat, 4, True)[0]
    for epoch in range(3):
        src.copy_(torch.arange(n, dtype=torch.float64, device="cuda") * (epoch + 1))
        dst.zero_()
        torch.cuda.synchronize()
        seq.run(plat)
        plat.synchronize()
        torch.cuda.synchronize()
        assert torch.equal(dst, src)
