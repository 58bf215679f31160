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
