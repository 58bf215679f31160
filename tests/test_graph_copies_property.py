"""Graph copies share their edge lists until one of them changes an edge
(include/tenzing/graph.hpp: `adj_`).  Property: whatever is done to a copy - edges added or
removed, vertices erased, ops rebound, compounds expanded - the graph it was copied from keeps
its vertices and edges, and both stay well formed; and a model kept in plain Python agrees with
the copy after every step."""
import random

from hypothesis import given, settings, strategies as st

import tenzing_b200 as tz
from tenzing_b200 import _C


def edges_of(g):
    out = set()
    for v in g.vertices():
        for s in g.succs(v):
            out.add((v.desc(), s.desc()))
    return out


def random_dag(rng, n):
    ops = [tz.EmptyGpuOp(f"k{i}") if rng.random() < 0.5 else tz.NoOp(f"n{i}") for i in range(n)]
    g = tz.Graph()
    for i, op in enumerate(ops):
        preds = [j for j in range(i) if rng.random() < 0.35]
        if not preds:
            g.start_then(op)
        for j in preds:
            g.then(ops[j], op)
    for i, op in enumerate(ops):
        if not g.succs(op):
            g.then_finish(op)
    assert g.validate() == ""
    return g, ops


@settings(max_examples=60, deadline=None)
@given(seed=st.integers(0, 10_000), n=st.integers(2, 14), steps=st.integers(1, 12))
def test_changing_a_copy_never_changes_the_original(seed, n, steps):
    rng = random.Random(seed)
    g, ops = random_dag(rng, n)
    want_edges, want_vertices = edges_of(g), sorted(v.desc() for v in g.vertices())

    copy = tz.Graph(g)
    model = set(want_edges)  # what the copy's edges should be
    live = {op.desc(): op for op in ops}
    for _ in range(steps):
        what = rng.choice(["bind", "edge", "erase_edge", "erase", "copy_again"])
        names = sorted(live)
        if what == "bind":
            unbound = [k for k in names if isinstance(live[k], _C.GpuOp)]
            if not unbound:
                continue
            k = rng.choice(unbound)
            bound = _C.BoundGpuOp(live[k], rng.randrange(3))
            copy = copy.clone_but_replace(bound, live[k])
            model = {(bound.desc() if a == k else a, bound.desc() if b == k else b) for a, b in model}
            del live[k]
            live[bound.desc()] = bound
        elif what == "edge" and len(names) >= 2:
            a, b = sorted(rng.sample(range(len(names)), 2))
            # only forward in creation order, so the graph stays acyclic
            order = {d: i for i, d in enumerate(sorted(live, key=lambda d: int(d.split("(")[0][1:])))}
            x, y = sorted((names[a], names[b]), key=order.get)
            copy.add_edge(live[x], live[y])
            model.add((x, y))
        elif what == "erase_edge":
            inner = [(a, b) for a, b in model if a in live and b in live]
            if not inner:
                continue
            a, b = rng.choice(sorted(inner))
            # keep every vertex on a Start..Finish path: only drop an edge that is not the last
            # one into b or out of a
            if sum(1 for e in model if e[1] == b) < 2 or sum(1 for e in model if e[0] == a) < 2:
                continue
            copy.erase_edge_only(live[a], live[b])
            model.discard((a, b))
        elif what == "erase" and len(names) > 1:
            k = rng.choice(names)
            copy.erase(live[k])
            model = {e for e in model if k not in e}
            del live[k]
            # keep the copy well formed: whoever lost its last predecessor / successor is
            # attached to Start / Finish (builds with TENZING_RUNTIME_SANITY_CHECKS insist)
            for d, op in live.items():
                if not any(b == d for _a, b in model):
                    copy.start_then(op)
                    model.add(("Start", d))
                if not any(a == d for a, _b in model):
                    copy.then_finish(op)
                    model.add((d, "Finish"))
        elif what == "copy_again":
            other = tz.Graph(copy)       # a copy of the copy, changed and thrown away
            if len(names) > 1:
                other.erase(live[rng.choice(names)])
        assert edges_of(copy) == model
    # the original is what it was
    assert edges_of(g) == want_edges
    assert sorted(v.desc() for v in g.vertices()) == want_vertices
    assert g.validate() == ""


@settings(max_examples=30, deadline=None)
@given(seed=st.integers(0, 10_000))
def test_expanding_a_compound_in_a_copy_leaves_the_original_alone(seed):
    from tenzing_b200.models import DistributedSpMV
    rng = random.Random(seed)
    s = DistributedSpMV(n=500, nnz_per_row=4, host=True)
    g = s.graph("choice")
    before = (edges_of(g), g.vertex_size(), g.edge_size())
    state = tz.State(g)
    for _ in range(rng.randrange(1, 12)):
        if state.is_terminal():
            break
        ds = state.get_decisions(s.platform)
        state = state.apply(ds[rng.randrange(len(ds))])
        assert state.graph().validate() == ""
    assert (edges_of(g), g.vertex_size(), g.edge_size()) == before
