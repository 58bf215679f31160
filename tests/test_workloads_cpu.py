"""Workload set-up logic on the CPU: symmetric heap in host mode, SpMV partition / need-list
exchange across an in-process world, halo geometry and graph structure."""
import numpy as np
import pytest
import torch

import tenzing_b200 as tz
from tenzing_b200 import _C
from tenzing_b200.models import HaloExchange, DistributedSpMV, common


def _csr(d, dtype=np.float64):
    import scipy.sparse as sp
    return sp.csr_matrix((np.frombuffer(d["val"], dtype=dtype), np.frombuffer(d["col_ind"], dtype=np.int32),
                          np.frombuffer(d["row_ptr"], dtype=np.int32)), shape=(d["rows"], max(d["cols"], 1)))


def test_generator_is_reproducible_per_row_block():
    n = 500
    whole = _csr(_C.spmv.random_band_rowsF64(n, 40, 10, 0, n, 5))
    part = _csr(_C.spmv.random_band_rowsF64(n, 40, 10, 100, 250, 5))
    assert (whole[100:250] != part).nnz == 0
    other = _csr(_C.spmv.random_band_rowsF64(n, 40, 10, 0, n, 6))
    assert (whole != other).nnz > 0
    coo = whole.tocoo()
    assert np.all(np.abs(coo.row - coo.col) <= 40)
    assert 6 <= whole.nnz / n <= 10 and whole.data.min() >= 0.5 and whole.data.max() < 1.5


@pytest.mark.parametrize("world,n,bw", [(2, 101, None), (4, 403, None), (3, 200, 200)])
def test_row_part_spmv_setup_matches_oracle(world, n, bw):
    seed = 11
    band = bw if bw is not None else max(1, n // world)
    A = _csr(_C.spmv.random_band_rowsF64(n, band, 8, 0, n, seed))
    x = DistributedSpMV.x_value(np.arange(n, dtype=np.int64))
    y_ref = A @ x
    info = {}

    def body(rank, comm):
        s = DistributedSpMV(comm=comm, n=n, bw=band, nnz_per_row=8, seed=seed, host=True)
        lb, ub = s.spmv.row_range()
        assert (lb, ub) == _C.spmv.get_partition(n, rank, world)
        la, ra = _csr(s.spmv.local_csr()), _csr(s.spmv.remote_csr())
        gl = np.frombuffer(s.spmv.remote_globals(), dtype=np.int32)
        # split: local columns shifted to 0, remote columns compacted in ascending global order
        assert np.all(np.diff(gl) > 0) and not np.any((gl >= lb) & (gl < ub))
        assert la.nnz + ra.nnz == A[lb:ub].nnz
        y = la @ x[lb:ub] + (ra @ x[gl] if len(gl) else 0)
        assert np.allclose(y, y_ref[lb:ub], rtol=1e-13)
        assert np.allclose(s.expected(), y_ref[lb:ub], rtol=1e-13)
        info[rank] = dict(lb=lb, ub=ub, gl=gl, send=s.spmv.send_params(), recv=s.spmv.recv_params(),
                          nsend=s.spmv.num_send())
        comm.barrier()

    _C.loopback_run(world, body)
    # what r receives from p is exactly what p sends to r, at the displacement r expects
    for r in range(world):
        for peer, displ, count in info[r]["recv"]:
            seg = info[r]["gl"][displ:displ + count]
            assert np.all((seg >= info[peer]["lb"]) & (seg < info[peer]["ub"]))
            match = [s for s in info[peer]["send"] if s[0] == r]
            assert len(match) == 1 and match[0][2] == count and match[0][3] == displ
        assert sum(c for _, _, c in info[r]["recv"]) == len(info[r]["gl"])
        assert sum(s[2] for s in info[r]["send"]) == info[r]["nsend"]


def test_spmv_graphs_have_the_reference_shape():
    s = DistributedSpMV(n=300, nnz_per_row=6, host=True)
    g = s.graph("reference")
    comp = g.find_name("spmv")
    inner = comp.graph()
    names = sorted(v.name() for v in inner.vertices())
    assert names == sorted(["Start", "yl", "yr", "Pack", "PostSend", "WaitSend", "PostRecv",
                            "WaitRecv", "y", "Finish"])  # reference: ops_spmv.cuh:396-417
    succ = {v.name(): sorted(o.name() for o in inner.succs(v)) for v in inner.vertices()}
    assert succ["Start"] == ["Pack", "PostRecv", "yl"]
    assert succ["Pack"] == ["PostSend"]
    assert succ["PostSend"] == succ["PostRecv"] == ["WaitRecv", "WaitSend"]
    assert succ["WaitRecv"] == ["yr"] and succ["yl"] == succ["yr"] == ["y"]
    assert succ["y"] == succ["WaitSend"] == ["Finish"]
    # expanding the compound gives exactly the inner graph (test_expand_spmv.cu:15-51)
    e = g.clone_but_expand(comp, inner)
    assert e.vertex_size() == inner.vertex_size()
    assert s.graph("fused").find_name("spmv").graph().vertex_size() == 5
    ch = s.graph("choice").find_name("spmv")
    assert [c.name() for c in ch.choices()] == ["spmv_reference", "spmv_fused", "spmv_fused_separate", "spmv_fused_all"]


def test_halo_geometry_and_graph_structure():
    h = HaloExchange(nX=8, nY=6, nZ=4, nQ=2, nGhost=2, host=True)
    d = h.domain
    # 12 cells + 0 origin elements (ghost radius 2 is already 16-byte aligned), padded to 128 B
    assert d.extent() == [12, 10, 8] and d.row_elems() == 16 and d.x_origin() == 0
    odd = HaloExchange(nX=8, nQ=1, nGhost=3, host=True).domain
    assert odd.x_origin() == 1 and (odd.index(3, 0, 0, 0) * 8) % 16 == 0
    assert len(d.dirs()) == 26
    # faces + edges + corners, all quantities, 8 bytes each
    faces = 2 * (2 * 6 * 4 + 8 * 2 * 4 + 8 * 6 * 2)
    edges = 4 * (2 * 2 * 4 + 2 * 6 * 2 + 8 * 2 * 2)
    corners = 8 * 8
    assert d.bytes_per_exchange() == (faces + edges + corners) * 2 * 8
    for style, nv in (("reference", 2 + 26 * 5 + 1), ("staged", 2 + 52), ("direct", 2 + 28),  # 26 puts + he_ready + he_wait_puts
                      ("fused_all", 4), ("fused_staged", 4)):
        g = h.graph("flat:" + style)
        assert g.vertex_size() == nv, style
    g = h.graph("flat:staged")
    put = g.find_name("he_packput_dx1_dy0_dz0")
    assert len(g.succs(put)) == 26  # every put precedes every wait
    g = h.graph("flat:reference")
    isend = g.find_name("he_isend_dx1_dy0_dz0")
    assert len([s for s in g.succs(isend) if s.name().startswith("he_waitrecv")]) == 26
    assert [p.name() for p in g.preds(isend)] == ["he_pack_dx1_dy0_dz0"]
    six = HaloExchange(nX=8, nQ=1, nGhost=1, neighbors=6, host=True)
    assert len(six.domain.dirs()) == 6


def test_halo_oracle_on_a_multi_rank_world_is_self_consistent():
    """fill() + the closed-form expectation agree with a brute-force neighbour copy (host memory,
    4 logical ranks in one process)"""
    grids = {}

    def body(rank, comm):
        h = HaloExchange(comm=comm, nX=4, nY=3, nZ=2, nQ=2, nGhost=1, host=True)
        h.fill()
        grids[rank] = (h, h.cells().clone(), h.expected().clone(), h.domain.rank_coord())
        comm.barrier()

    _C.loopback_run(4, body)
    h0 = grids[0][0]
    rg = h0.domain.rank_grid()
    assert rg == [2, 2, 1]
    by_coord = {tuple(c): r for r, (_, _, _, c) in grids.items()}
    nx, ny, nz, g = 4, 3, 2, 1
    for r, (h, cells, exp, coord) in grids.items():
        for dx, dy, dz in h.domain.dirs():
            nb = by_coord[((coord[0] + dx) % rg[0], (coord[1] + dy) % rg[1], (coord[2] + dz) % rg[2])]
            assert nb == h.domain.neighbor(dx, dy, dz)
            src = grids[nb][1]

            def rng(d, n, ghost):
                if ghost:
                    return slice(0, g) if d < 0 else (slice(g + n, g + n + g) if d > 0 else slice(g, g + n))
                # the neighbour's interior cells adjacent to its opposite face
                return slice(n, n + g) if d < 0 else (slice(g, 2 * g) if d > 0 else slice(g, g + n))

            want = src[:, rng(dz, nz, False), rng(dy, ny, False), rng(dx, nx, False)]
            got = exp[:, rng(dz, nz, True), rng(dy, ny, True), rng(dx, nx, True)]
            assert torch.equal(got, want)


def test_schedules_round_trip_through_json_for_workload_graphs():
    h = HaloExchange(nX=8, nQ=1, nGhost=1, neighbors=6, host=True)
    g = h.graph("choice")
    plat = h.platform
    opts = _C.mcts.Opts()
    opts.n_iters = 5
    res = _C.mcts.explore(g, plat, tz.PyBenchmarker(lambda s: float(len(s))), "Random", opts)
    for r in res.sim_results:
        text = r.path.to_json(g)
        back = _C.sequence_from_json(text, g)
        assert back.canonical_key() == r.path.canonical_key()
    # the CSV of a search over the choice graph replays against the same graph
    replay = tz.CsvBenchmarker.from_text(res.csv(), g)
    assert replay.num_rows() == 5
    seq = res.sim_results[0].path
    assert replay.benchmark(seq, plat).pct10 == float(len(seq))
    assert common.fixed_schedule(h.graph("staged"), plat).names()[-1] == "Finish"


def test_matrix_market_file_is_read_on_root_and_distributed(tmp_path):
    import scipy.io
    import scipy.sparse as sp
    rng = np.random.default_rng(3)
    n = 57
    A = sp.random(n, n, density=0.15, random_state=rng, format="coo") + sp.identity(n)
    path = str(tmp_path / "a.mtx")
    scipy.io.mmwrite(path, A.tocoo())
    A = A.tocsr()
    x = DistributedSpMV.x_value(np.arange(n, dtype=np.int64))
    y_ref = A @ x
    seen = {}

    def body(rank, comm):
        s = DistributedSpMV(comm=comm, matrix_file=path, host=True)
        lb, ub = s.spmv.row_range()
        assert s.n == n and (lb, ub) == _C.spmv.get_partition(n, rank, 3)
        assert np.allclose(s.expected(), y_ref[lb:ub], rtol=1e-12)
        seen[rank] = s.spmv.local_nnz() + s.spmv.remote_nnz()
        comm.barrier()

    _C.loopback_run(3, body)
    assert sum(seen.values()) == A.nnz


def test_every_workload_graph_and_rewrite_keeps_the_graph_invariants():
    """Graph.validate() on every implementation style, and along a random walk through the
    decision tree (expand / choose / assign-stream rewrites) of the choice graphs."""
    import random
    h = HaloExchange(nX=8, nQ=1, nGhost=1, neighbors=6, host=True)
    s = DistributedSpMV(n=300, nnz_per_row=6, host=True)
    from tenzing_b200.models import halo as halo_mod, spmv as spmv_mod
    for wl, styles in ((h, halo_mod.STYLES), (s, spmv_mod.STYLES)):
        for style in styles + ("choice",):
            g = wl.graph(style)
            assert g.validate() == "", (style, g.validate())
        plat = tz.Platform.make_null(2)
        rng = random.Random(7)
        for walk in range(4):
            st = tz.State(wl.graph("choice"))
            for _ in range(400):
                ds = st.get_decisions(plat)
                if not ds:
                    break
                st = st.apply(rng.choice(ds))
                assert st.graph().validate() == ""
            assert not st.get_decisions(plat), "walk did not reach a terminal state"


def test_which_sequences_can_be_replayed_as_a_cuda_graph():
    """GraphReplay::supported: device-only schedules (kernels, event records, stream waits, the
    trailing host sync in front of Finish) yes; anything that runs an op on the host no"""
    from tenzing_b200 import _C
    from tenzing_b200.models import DistributedSpMV, HaloExchange, common
    h = HaloExchange(nX=8, nQ=1, nGhost=2, host=True)
    for style in ("fused_all", "fused_staged", "fused_mixed", "staged", "direct", "fused_pull"):
        ok, why = _C.GraphReplay.supported(common.fixed_schedule(h.graph(style), h.platform))
        assert ok, (style, why)
    ok, why = _C.GraphReplay.supported(common.fixed_schedule(h.graph("reference"), h.platform))
    assert not ok and "host" in why
    s = DistributedSpMV(n=2000, host=True)
    assert _C.GraphReplay.supported(common.fixed_schedule(s.graph("fused_all"), s.platform))[0]
    assert not _C.GraphReplay.supported(common.fixed_schedule(s.graph("reference"), s.platform))[0]
