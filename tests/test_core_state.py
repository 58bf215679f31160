"""Decisions, state transitions and synchronisation insertion.

Mirrors test/test_noop_graph.cpp:10-42 and the rule table of
include/tenzing/event_synchronizer.hpp."""
import tenzing_b200 as tz
from tenzing_b200 import _C


def gpu_chain(names, extra_edges=()):
    ops = {n: tz.EmptyGpuOp(n) for n in names}
    return ops


def run_to_end(state, plat, pick=lambda ds: ds[0]):
    while not state.is_terminal():
        state = state.apply(pick(state.get_decisions(plat)))
    return state


def test_noop_graph_decisions():
    g = tz.Graph()
    op1 = tz.NoOp("op1")
    g.start_then(op1)
    g.then_finish(op1)
    plat = tz.Platform.make_null(1)
    st = tz.State(g)
    assert len(st.sequence()) == 1 and st.sequence()[0].name() == "Start"
    ds = st.get_decisions(plat)
    assert len(ds) == 1 and ds[0].kind() == "ExecuteOp" and ds[0].op.name() == "op1"
    for d in ds:
        assert len(st.apply(d).sequence()) == 2
    assert not st.is_terminal()
    end = run_to_end(st, plat)
    assert end.sequence().names() == ["Start", "op1", "Finish"]
    assert end.get_decisions(plat) == []


def test_gpu_op_offers_one_assignment_per_stream():
    k = tz.EmptyGpuOp("k")
    g = tz.Graph()
    g.start_then(k)
    g.then_finish(k)
    for n in (1, 2, 4):
        plat = tz.Platform.make_null(n)
        ds = tz.State(g).get_decisions(plat)
        assert [d.kind() for d in ds] == ["AssignOpStream"] * n
        assert sorted(d.stream for d in ds) == list(range(n))
    plat = tz.Platform.make_null(2)
    st = tz.State(g).apply(tz.State(g).get_decisions(plat)[1])
    assert len(st.sequence()) == 1  # binding does not execute anything
    assert st.graph().find_name("k").stream() == 1


def test_same_stream_needs_no_sync_and_gpu_to_cpu_needs_record_then_sync():
    a, b = tz.EmptyGpuOp("a"), tz.EmptyGpuOp("b")
    g = tz.Graph()
    g.start_then(a)
    g.then(a, b)
    g.then_finish(b)
    plat = tz.Platform.make_null(1)
    end = run_to_end(tz.State(g), plat)
    assert end.sequence().names() == ["Start", "a", "b", "CER-after-b", "CES-b4-Finish", "Finish"]
    cer, ces = end.sequence()[3], end.sequence()[4]
    assert cer.kind() == "CudaEventRecord" and cer.stream() == 0
    assert ces.kind() == "CudaEventSync" and ces.event() == cer.event()


def test_cross_stream_needs_record_then_stream_wait():
    a, b = tz.EmptyGpuOp("a"), tz.EmptyGpuOp("b")
    g = tz.Graph()
    g.start_then(a)
    g.then(a, b)
    g.then_finish(b)
    g = _C.apply_assignment(g, [("a", 0), ("b", 1)])
    plat = tz.Platform.make_null(2)
    st = tz.State(g)
    st = st.apply(st.get_decisions(plat)[0])  # a
    bb = g.find_name("b")
    assert not tz.EventSynchronizer.is_synced(bb, g, st.sequence())
    syncs = st.get_syncs_before_op(bb)
    assert [s.kind() for s in syncs] == ["CudaEventRecord"]
    assert syncs[0].name() == "CER-b4-b" and syncs[0].stream() == 0
    st = st.apply(tz.ExecuteOp(syncs[0]))
    syncs = st.get_syncs_before_op(bb)
    assert [s.kind() for s in syncs] == ["CudaStreamWaitEvent"]
    assert syncs[0].name() == "CSWE-after-a" and syncs[0].stream() == 1
    st = st.apply(tz.ExecuteOp(syncs[0]))
    assert tz.EventSynchronizer.is_synced(bb, g, st.sequence())
    ds = st.get_decisions(plat)
    assert [d.op.name() for d in ds] == ["b"]


def test_cpu_to_gpu_and_cpu_to_cpu_need_nothing():
    c1, c2, k = tz.NoOp("c1"), tz.NoOp("c2"), tz.EmptyGpuOp("k")
    g = tz.Graph()
    g.start_then(c1)
    g.then(c1, c2)
    g.then(c2, k)
    g.then_finish(k)
    plat = tz.Platform.make_null(1)
    end = run_to_end(tz.State(g), plat)
    assert end.sequence().names()[:4] == ["Start", "c1", "c2", "k"]


def test_event_ids_are_never_reused_in_a_sequence():
    ks = [tz.EmptyGpuOp(f"k{i}") for i in range(3)]
    host = [tz.NoOp(f"h{i}") for i in range(3)]
    g = tz.Graph()
    for k, h in zip(ks, host):
        g.start_then(k)
        g.then(k, h)
        g.then_finish(h)
    plat = tz.Platform.make_null(2)
    end = run_to_end(tz.State(g), plat, pick=lambda ds: ds[-1])
    events = [op.event() for op in end.sequence() if op.kind() == "CudaEventRecord"]
    assert len(events) == len(set(events)) and len(events) >= 1


def test_sync_decisions_are_offered_next_to_other_ready_ops():
    """sync placement is itself searched: the record can be issued while other ops are ready"""
    a, b, c = tz.EmptyGpuOp("a"), tz.NoOp("b"), tz.NoOp("c")
    g = tz.Graph()
    g.start_then(a)
    g.then(a, b)
    g.then_finish(b)
    g.start_then(c)
    g.then_finish(c)
    plat = tz.Platform.make_null(1)
    st = tz.State(g)
    st = st.apply([d for d in st.get_decisions(plat) if d.kind() == "AssignOpStream"][0])
    st = st.apply([d for d in st.get_decisions(plat) if d.desc().startswith("Execute a")][0])
    kinds = sorted(d.op.name() for d in st.get_decisions(plat))
    assert kinds == ["CER-after-a", "c"]


def test_expand_and_choose_decisions():
    inner = tz.Graph()
    i1, i2 = tz.NoOp("i1"), tz.NoOp("i2")
    inner.start_then(i1)
    inner.then(i1, i2)
    inner.then_finish(i2)
    comp = tz.CompoundOp("comp", inner)
    alt = tz.NoOp("alt")
    choice = tz.ChoiceOp("pick", [comp, alt])
    g = tz.Graph()
    g.start_then(choice)
    g.then_finish(choice)
    plat = tz.Platform.make_null(1)
    st = tz.State(g)
    ds = st.get_decisions(plat)
    assert [d.kind() for d in ds] == ["ChooseOp", "ChooseOp"]
    s_comp = st.apply(ds[0])
    d2 = s_comp.get_decisions(plat)
    assert [d.kind() for d in d2] == ["ExpandOp"]
    s_exp = s_comp.apply(d2[0])
    assert s_exp.graph().vertex_size() == 4
    assert run_to_end(s_exp, plat).sequence().names() == ["Start", "i1", "i2", "Finish"]
    assert run_to_end(st.apply(ds[1]), plat).sequence().names() == ["Start", "alt", "Finish"]


def test_state_frontier_drops_equivalent_siblings():
    k = tz.EmptyGpuOp("k")
    g = tz.Graph()
    g.start_then(k)
    g.then_finish(k)
    plat = tz.Platform.make_null(3)
    # binding k to stream 0, 1 or 2 gives three states that differ only by a stream renaming
    assert len(tz.State(g).get_decisions(plat)) == 3
    assert len(tz.State(g).frontier(plat)) == 1


def test_get_frontier_legacy():
    a, b = tz.EmptyGpuOp("a"), tz.EmptyGpuOp("b")
    g = tz.Graph()
    g.start_then(a)
    g.then(a, b)
    g.then_finish(b)
    plat = tz.Platform.make_null(2)
    fr = _C.mcts.get_frontier(plat, g, [g.start()])
    assert sorted(o.desc() for o in fr) == ["a(s0)", "a(s1)"]
    fr = _C.mcts.get_frontier(plat, g, [g.start(), tz.BoundGpuOp(a, 0)])
    assert sorted(o.kind() for o in fr) == ["BoundGpuOp", "CudaEventRecord"]


def _rollouts_agree_with_fresh_states(graph, plat, seed, rollouts=6):
    """A state derived step by step carries caches from its ancestors (sequence digest, executed
    and synchronised vertices, the path index); a state built from (graph, sequence) has none.
    Both must offer the same decisions and have the same digest at every step."""
    import random
    rng = random.Random(seed)
    for _ in range(rollouts):
        st = tz.State(graph)
        steps = 0
        while not st.is_terminal():
            fresh = tz.State(st.graph(), st.sequence())
            got = [d.desc() for d in st.get_decisions(plat)]
            want = [d.desc() for d in fresh.get_decisions(plat)]
            assert got == want, (steps, got, want)
            assert st.canonical_hash() == fresh.canonical_hash()
            assert st.canonical_key() == fresh.canonical_key()
            ds = st.get_decisions(plat)
            st = st.apply(ds[rng.randrange(len(ds))])
            steps += 1
        assert steps > 0


def test_derived_states_agree_with_fresh_ones_on_a_gpu_diamond():
    a, b, c, d = (tz.EmptyGpuOp(n) for n in "abcd")
    h = tz.NoOp("h")
    g = tz.Graph()
    g.start_then(a)
    g.then(a, b)
    g.then(a, c)
    g.then(b, d)
    g.then(c, d)
    g.then(c, h)
    g.then_finish(d)
    g.then_finish(h)
    for n_streams in (1, 2, 3):
        _rollouts_agree_with_fresh_states(g, tz.Platform.make_null(n_streams), seed=n_streams)


def test_derived_states_agree_with_fresh_ones_on_the_workload_graphs():
    from tenzing_b200.models import DistributedSpMV, HaloExchange
    s = DistributedSpMV(n=2000, nnz_per_row=6, host=True)
    for style in ("reference", "choice"):
        _rollouts_agree_with_fresh_states(s.graph(style), s.platform, seed=7, rollouts=3)
    h = HaloExchange(nX=8, nQ=1, nGhost=1, n_streams=2, host=True)
    for style in ("direct", "choice"):
        _rollouts_agree_with_fresh_states(h.graph(style), h.platform, seed=11, rollouts=2)


def test_equal_digests_iff_equal_keys_over_a_whole_enumeration():
    a, b, c = (tz.EmptyGpuOp(n) for n in "abc")
    g = tz.Graph()
    g.start_then(a)
    g.start_then(b)
    g.then(a, c)
    g.then(b, c)
    g.then_finish(c)
    plat = tz.Platform.make_null(2)
    seen = {}
    stack = [tz.State(g)]
    while stack:
        st = stack.pop()
        k, h = st.canonical_key(), st.canonical_hash()
        assert seen.setdefault(k, h) == h
        if not st.is_terminal():
            stack.extend(st.apply(d) for d in st.get_decisions(plat))
    assert len(set(seen.values())) == len(seen) > 50
