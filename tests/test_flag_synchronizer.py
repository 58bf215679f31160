"""FlagSynchronizer: every schedule the search can produce for the fused halo styles is
deadlock free (each flag wait follows its publish in the shared sequence), a hand-made
schedule that waits first is rejected, and the publish -> wait edges can be derived."""
import tenzing_b200 as tz
from tenzing_b200 import _C
from tenzing_b200.models import HaloExchange

FS = _C.FlagSynchronizer


def test_every_searched_schedule_is_deadlock_free():
    h = HaloExchange(nX=8, nQ=1, nGhost=1, neighbors=6, host=True)
    for style in ("staged", "direct", "fused_all", "fused_staged", "fused_groups", "fused_mixed",
                  "fused_pipelined", "fused_pull"):
        seqs = tz.dfs.get_all_sequences(h.graph(style), h.platform, 300, True)
        assert seqs
        for s in seqs:
            assert FS.is_deadlock_free(s), (style, s.desc())


def test_a_wait_before_its_publish_is_reported():
    h = HaloExchange(nX=8, nQ=1, nGhost=1, neighbors=6, host=True)
    g = h.graph("flat:fused_staged")
    put = _C.BoundGpuOp(g.find_name("he_packput_all"), 0)
    fin = _C.BoundGpuOp(g.find_name("he_waitunpack_all"), 0)
    good = tz.Sequence([g.start(), put, fin, g.finish()])
    bad = tz.Sequence([g.start(), fin, put, g.finish()])
    assert FS.is_deadlock_free(good)
    v = FS.check(bad)
    assert len(v) == 6 and all(pos == 1 and name == "he_waitunpack_all" for pos, name, _ in v)
    assert not FS.is_ready(fin, tz.Sequence([g.start()]))
    assert FS.is_ready(fin, tz.Sequence([g.start(), put]))
    assert FS.is_ready(put, tz.Sequence([g.start()]))


def test_edges_can_be_derived_from_the_flag_channels():
    h = HaloExchange(nX=8, nQ=1, nGhost=1, neighbors=6, host=True)
    g = h.graph("flat:staged")
    # strip the publish -> wait edges the builder added, then let the synchronizer re-derive them
    puts = [v for v in g.vertices() if v.name().startswith("he_packput")]
    waits = [v for v in g.vertices() if v.name().startswith("he_waitunpack")]
    g2 = tz.Graph(g)
    for p in puts:
        for w in waits:
            g2.erase_edge_only(p, w)
        g2.then_finish(p)
    for w in waits:
        g2.start_then(w)
    added = FS.add_edges(g2)
    assert added == 6  # one per channel: the put towards d feeds the unpack of side -d
    for w in waits:
        assert len([p for p in g2.preds(w) if p.name().startswith("he_packput")]) == 1
    assert FS.add_edges(g2) == 0
    for s in tz.dfs.get_all_sequences(g2, tz.Platform.make_null(1), 200, True):
        assert FS.is_deadlock_free(s)


def test_spmv_ops_declare_their_flag_channels_too():
    """every SpMV style (validated and experimental): all enumerated schedules pass, the waits
    really name channels, and a hand-made schedule that waits before the publish is reported"""
    from tenzing_b200.models import DistributedSpMV
    from tenzing_b200.models import spmv as spmv_mod

    def body(rank, comm):
        s = DistributedSpMV(comm=comm, n=120, nnz_per_row=5, host=True)
        for style in spmv_mod.STYLES + spmv_mod.EXPERIMENTAL_STYLES:
            seqs = tz.dfs.get_all_sequences(s.graph(style), s.platform, 200, True)
            assert seqs
            for q in seqs:
                assert FS.is_deadlock_free(q), (style, q.desc())
        inner = s.graph("fused").find_name("spmv").graph()
        put = _C.BoundGpuOp(inner.find_name("PackPut"), 0)
        yl = _C.BoundGpuOp(inner.find_name("yl"), 0)
        yr = _C.BoundGpuOp(inner.find_name("yr"), 0)
        g = s.graph("fused")
        assert FS.is_deadlock_free(tz.Sequence([g.start(), put, yl, yr, g.finish()]))
        v = FS.check(tz.Sequence([g.start(), yl, yr, put, g.finish()]))
        assert len(v) == comm.size() and all(name == "yr" for _, name, _ in v)   # one channel per source rank
        # the one-launch product publishes before it waits: fine on its own
        one = s.graph("fused_all").find_name("spmv").graph().find_name("SpMVAll")
        assert FS.is_deadlock_free(tz.Sequence([g.start(), _C.BoundGpuOp(one, 1), g.finish()]))
        # the pull style: ready -> pull -> free is the only legal order of the three
        pull = s.graph("fused_pull").find_name("spmv").graph()
        ready, yrp, free = (_C.BoundGpuOp(pull.find_name(n), 0) for n in ("XReady", "yrPull", "XFree"))
        assert FS.is_deadlock_free(tz.Sequence([g.start(), ready, yrp, free, g.finish()]))
        assert FS.check(tz.Sequence([g.start(), yrp, ready, free, g.finish()]))
        assert FS.check(tz.Sequence([g.start(), ready, free, yrp, g.finish()]))

    _C.loopback_run(3, body)
