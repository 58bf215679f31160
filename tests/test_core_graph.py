"""Graph / op / sequence unit tests.

Mirrors the assertions of the reference's in-source doctest cases
(src/graph.cpp:422-501, src/operation.cpp:87-101, src/sequence.cpp:169-176).
"""
import json

import pytest

import tenzing_b200 as tz
from tenzing_b200 import _C


def chain(*names):
    g = tz.Graph()
    ops = [tz.NoOp(n) for n in names]
    g.start_then(ops[0])
    for a, b in zip(ops, ops[1:]):
        g.then(a, b)
    g.then_finish(ops[-1])
    return g, ops


def test_empty_graph_has_start_and_finish():
    g = tz.Graph()
    assert g.vertex_size() == 2
    assert g.start().name() == "Start"
    assert g.finish().name() == "Finish"


def test_then_builds_edges():
    g, (a, b) = chain("a", "b")
    assert g.vertex_size() == 4
    assert [o.name() for o in g.succs(a)] == ["b"]
    assert [o.name() for o in g.preds(b)] == ["a"]
    assert [o.name() for o in g.start_vertices()] == ["a"]
    assert [o.name() for o in g.finish_vertices()] == ["b"]


def test_then_rejects_start_finish():
    g = tz.Graph()
    with pytest.raises(RuntimeError):
        g.then(g.start(), tz.NoOp("x"))
    with pytest.raises(RuntimeError):
        g.then(tz.NoOp("x"), g.finish())


def test_noop_identity_is_by_name():
    assert tz.NoOp("a").eq(tz.NoOp("a"))
    assert not tz.NoOp("a").eq(tz.NoOp("b"))
    assert tz.NoOp("a").lt(tz.NoOp("b"))
    assert _C.Start().eq(_C.Start())
    assert not _C.Start().eq(_C.Finish())


def test_clone_is_deep_but_equal():
    g, (a, b) = chain("a", "b")
    c = g.clone()
    assert c.vertex_size() == g.vertex_size()
    assert c.contains(a) and c.contains(b)
    assert c.find(a) is not a  # cloned op object
    assert bool(_C.graph_equivalence(g, c))


def test_clone_but_replace():
    g, (a, b) = chain("a", "b")
    r = g.clone_but_replace(tz.NoOp("z"), a)
    assert r.vertex_size() == g.vertex_size()
    assert not r.contains(a)
    z = r.find_name("z")
    assert [o.name() for o in r.succs(z)] == ["b"]
    assert [o.name() for o in r.preds(z)] == ["Start"]
    # the original is untouched
    assert g.contains(a)


def test_clone_but_expand_vertex_count():
    inner, _ = chain("i1", "i2", "i3")
    comp = tz.CompoundOp("comp", inner)
    g = tz.Graph()
    pre, post = tz.NoOp("pre"), tz.NoOp("post")
    g.start_then(pre)
    g.then(pre, comp)
    g.then(comp, post)
    g.then_finish(post)
    e = g.clone_but_expand(comp, comp.graph())
    # comp replaced by 3 inner vertices
    assert e.vertex_size() == g.vertex_size() - 1 + 3
    assert [o.name() for o in e.succs(pre)] == ["i1"]
    assert [o.name() for o in e.preds(post)] == ["i3"]
    assert not e.contains(comp)


def test_expand_whole_graph_equals_inner():
    """test_expand_spmv.cu:15-51: Start->compound->Finish expands to the inner vertex count"""
    inner = tz.Graph()
    a, b, c = tz.NoOp("a"), tz.NoOp("b"), tz.NoOp("c")
    inner.start_then(a)
    inner.start_then(b)
    inner.then(a, c)
    inner.then(b, c)
    inner.then_finish(c)
    comp = tz.CompoundOp("comp", inner)
    g = tz.Graph()
    g.start_then(comp)
    g.then_finish(comp)
    e = g.clone_but_expand(comp, inner)
    assert e.vertex_size() == inner.vertex_size()
    assert e.edge_size() == inner.edge_size()


def test_frontier_ignores_stream_binding():
    k1, k2 = tz.EmptyGpuOp("k1"), tz.EmptyGpuOp("k2")
    g = tz.Graph()
    g.start_then(k1)
    g.then(k1, k2)
    g.then_finish(k2)
    done = [g.start(), tz.BoundGpuOp(k1, 1)]
    assert [o.name() for o in g.frontier(done)] == ["k2"]


def test_erase_and_edges():
    g, (a, b, c) = chain("a", "b", "c")
    g.erase(b)
    assert g.vertex_size() == 4
    assert g.succs(a) == []
    g.add_edge(a, c)
    g.erase_edge_only(a, c)
    assert g.preds(c) == []


def test_graphviz_text():
    g, _ = chain("a")
    dot = g.graphviz()
    assert dot.startswith("digraph") and "a" in dot and "->" in dot


def test_empty_sequence_and_unique_event():
    s = tz.Sequence()
    assert len(s) == 0
    assert s.new_unique_event() == 0
    s.push_back(tz.CudaEventRecord(0, 0))
    s.push_back(tz.CudaEventSync(2))
    assert s.new_unique_event() == 1


def test_sync_op_json_and_names():
    cer = tz.CudaEventRecord(3, 1, "CER-after-yl")
    assert json.loads(cer.json()) == {"event": 3, "kind": "CudaEventRecord",
                                      "name": "CER-after-yl", "stream": 1}
    # keys are emitted alphabetically, compact
    assert cer.json() == '{"event":3,"kind":"CudaEventRecord","name":"CER-after-yl","stream":1}'
    cswe = tz.CudaStreamWaitEvent(0, 3)
    assert json.loads(cswe.json())["kind"] == "CudaStreamWaitEvent"
    assert cswe.name() == "CSWE-anon"
    ces = tz.CudaEventSync(3)
    assert json.loads(ces.json()) == {"event": 3, "kind": "CudaEventSync", "name": "CES-anon"}
    # identity is (event, stream), not the name
    assert tz.CudaEventRecord(3, 1, "other").eq(cer)
    assert not tz.CudaEventRecord(4, 1).eq(cer)
    assert tz.NoOp("n").json() == '{"kind":"NoOp","name":"n"}'
    k = tz.EmptyGpuOp("k")
    assert tz.BoundGpuOp(k, 1).json() == '{"name":"k","stream":1}'


def test_sequence_equivalence_up_to_renaming():
    k1, k2 = tz.EmptyGpuOp("k1"), tz.EmptyGpuOp("k2")
    a = tz.Sequence([tz.BoundGpuOp(k1, 0), tz.CudaEventRecord(0, 0), tz.CudaStreamWaitEvent(1, 0),
                     tz.BoundGpuOp(k2, 1)])
    b = tz.Sequence([tz.BoundGpuOp(k1, 1), tz.CudaEventRecord(5, 1), tz.CudaStreamWaitEvent(0, 5),
                     tz.BoundGpuOp(k2, 0)])
    c = tz.Sequence([tz.BoundGpuOp(k1, 1), tz.CudaEventRecord(5, 1), tz.CudaStreamWaitEvent(1, 5),
                     tz.BoundGpuOp(k2, 1)])
    eq = _C.sequence_equivalence(a, b)
    assert bool(eq) and eq.stream_map() == {0: 1, 1: 0} and eq.event_map() == {0: 5}
    assert not bool(_C.sequence_equivalence(a, c))
    assert a.canonical_key() == b.canonical_key() != c.canonical_key()
    # a true bijection: 0->1, 1->2 is accepted (the reference's involution rejects it, D17)
    d = tz.Sequence([tz.BoundGpuOp(k1, 1), tz.CudaEventRecord(0, 1), tz.CudaStreamWaitEvent(2, 0),
                     tz.BoundGpuOp(k2, 2)])
    assert bool(_C.sequence_equivalence(a, d))


def test_serdes_roundtrip_resolves_by_name():
    k1, k2 = tz.EmptyGpuOp("k1"), tz.EmptyGpuOp("k2")
    n = tz.NoOp("host")
    g = tz.Graph()
    g.start_then(k1)
    g.then(k1, k2)
    g.then(k2, n)
    g.then_finish(n)
    seq = tz.Sequence([g.start(), tz.BoundGpuOp(k1, 0), tz.CudaEventRecord(0, 0, "CER-b4-k2"),
                       tz.CudaStreamWaitEvent(1, 0, "CSWE-after-k1"), tz.BoundGpuOp(k2, 1),
                       tz.CudaEventRecord(1, 1), tz.CudaEventSync(1), n, g.finish()])
    text = seq.to_json(g)
    for e in json.loads(text):
        assert "in_graph" in e
    back = _C.sequence_from_json(text, g)
    assert back.names() == seq.names()
    assert back.canonical_key() == seq.canonical_key()
    assert back[1].stream() == 0 and back[4].stream() == 1
    # ops inside compound / choice ops are found too
    inner = tz.Graph()
    inner.start_then(k1)
    inner.then_finish(k1)
    outer = tz.Graph()
    ch = tz.ChoiceOp("pick", [tz.CompoundOp("c", inner), k2])
    outer.start_then(ch)
    outer.then_finish(ch)
    op = _C.op_from_json('{"name":"k1","stream":1}', outer)
    assert op.name() == "k1" and op.stream() == 1


def test_use_streams_enumerators():
    k = [tz.EmptyGpuOp(f"k{i}") for i in range(3)]
    g = tz.Graph()
    for op in k:
        g.start_then(op)
        g.then_finish(op)
    assert len(_C.use_streams(g, [0, 1])) == 8
    # canonical: op i may open at most one new stream -> set partitions into <=2 blocks = 4
    canon = _C.use_streams2(g, [0, 1])
    assert len(canon) == 4
    keys = set()
    for a in canon:
        for b in canon:
            if a is not b:
                assert not _C.is_equivalent_stream_mapping(a, b)
    # every full assignment is equivalent to exactly one canonical one
    for full in _C.use_streams(g, [0, 1]):
        assert sum(_C.is_equivalent_stream_mapping(full, c) for c in canon) == 1
    del keys


def test_validate_reports_structural_violations():
    """Graph.validate(): the invariants TENZING_RUNTIME_SANITY_CHECKS builds assert after rewrites."""
    g = tz.Graph()
    a, b, c = tz.NoOp("a"), tz.NoOp("b"), tz.NoOp("c")
    g.start_then(a)
    g.then(a, b)
    g.then_finish(b)
    assert g.validate() == ""
    assert g.preds_find_name("b") == (True, [a])
    assert g.succs_find_name("a") == (True, [b])
    assert g.preds_find_name("nope") == (False, [])
    # a vertex that leads nowhere
    g.then(a, c)
    assert "does not lead to Finish" in g.validate()
    g.then_finish(c)
    assert g.validate() == ""
    # expansion and replacement keep the invariants
    sub = tz.Graph()
    x, y = tz.NoOp("x"), tz.NoOp("y")
    sub.start_then(x)
    sub.then(x, y)
    sub.then_finish(y)
    assert g.clone_but_expand(b, sub).validate() == ""
    assert g.clone_but_replace(tz.NoOp("b2"), b).validate() == ""
    # a cycle
    g.then(b, a)
    assert g.validate() == "the graph has a cycle"
    # erasing the only path to Finish
    h = tz.Graph()
    h.start_then(a)
    h.then_finish(a)
    h.erase(a)
    assert h.validate() != ""
