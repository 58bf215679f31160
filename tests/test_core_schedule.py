"""remove_redundant_syncs rule by rule (reference: src/schedule.cpp:19-321) and the legacy
order enumerators, checked against brute force."""
import itertools

import tenzing_b200 as tz
from tenzing_b200 import _C

K = {n: tz.EmptyGpuOp(n) for n in "abcd"}


def B(n, s):
    return tz.BoundGpuOp(K[n], s)


def names(seq):
    return [f"{o.kind()}:{','.join(map(str, o.events()))}:{','.join(map(str, o.streams()))}"
            if o.kind().startswith("Cuda") or o.kind().startswith("Stream") else o.name() for o in seq]


def clean(ops):
    seq = tz.Sequence(ops)
    removed = _C.remove_redundant_syncs(seq)
    return removed, seq


def test_r1_unused_record_is_dropped():
    removed, seq = clean([B("a", 0), tz.CudaEventRecord(0, 0), B("b", 0)])
    assert removed == 1 and seq.names() == ["a", "b"]


def test_r2_stream_wait_without_later_gpu_op_is_dropped_and_its_record_follows():
    removed, seq = clean([B("a", 0), tz.CudaEventRecord(0, 0), tz.CudaStreamWaitEvent(1, 0),
                          tz.NoOp("host")])
    assert removed == 2 and seq.names() == ["a", "host"]
    # ... but kept when a GPU op follows in the waiting stream
    removed, seq = clean([B("a", 0), tz.CudaEventRecord(0, 0), tz.CudaStreamWaitEvent(1, 0), B("b", 1)])
    assert removed == 0


def test_r3_back_to_back_stream_syncs():
    removed, seq = clean([B("a", 0), tz.StreamSync(0), tz.NoOp("h"), tz.StreamSync(0)])
    assert removed == 1 and [o.kind() for o in seq] == ["BoundGpuOp", "NoOp", "StreamSync"]
    removed, _ = clean([B("a", 0), tz.StreamSync(0), B("b", 0), tz.StreamSync(0)])
    assert removed == 0
    removed, _ = clean([B("a", 0), tz.StreamSync(0), tz.StreamSync(1)])
    assert removed == 0


def test_r4_two_records_of_the_same_stream_state():
    # both records capture "after a"; the host already waited at the first sync
    ops = [B("a", 0), tz.CudaEventRecord(0, 0), tz.CudaEventRecord(1, 0), tz.CudaEventSync(0),
           tz.NoOp("h1"), tz.CudaEventSync(1), tz.NoOp("h2")]
    removed, seq = clean(ops)
    assert removed == 2
    assert [o.kind() for o in seq] == ["BoundGpuOp", "CudaEventRecord", "CudaEventSync", "NoOp", "NoOp"]
    assert seq[1].event() == 0 and seq[2].event() == 0  # the earlier sync is the one that stays


def test_r5_later_record_synced_first_makes_the_earlier_pair_redundant():
    ops = [B("a", 0), tz.CudaEventRecord(0, 0), B("b", 0), tz.CudaEventRecord(1, 0),
           tz.CudaEventSync(1), tz.NoOp("h1"), tz.CudaEventSync(0), tz.NoOp("h2")]
    removed, seq = clean(ops)
    assert removed == 2
    assert [o.events() for o in seq if o.kind().startswith("Cuda")] == [[1], [1]]
    # the other order is NOT redundant: sync(0) only covers a, sync(1) is still needed for b
    ops = [B("a", 0), tz.CudaEventRecord(0, 0), B("b", 0), tz.CudaEventRecord(1, 0),
           tz.CudaEventSync(0), tz.NoOp("h1"), tz.CudaEventSync(1), tz.NoOp("h2")]
    assert clean(ops)[0] == 0


def test_r6_duplicate_host_sync_and_self_stream_wait():
    removed, seq = clean([B("a", 0), tz.CudaEventRecord(0, 0), tz.CudaEventSync(0), tz.CudaEventSync(0)])
    assert removed == 1
    removed, seq = clean([B("a", 0), tz.CudaEventRecord(0, 0), tz.CudaStreamWaitEvent(0, 0), B("b", 0)])
    assert removed == 2 and seq.names() == ["a", "b"]


def test_cleaning_never_breaks_solver_output():
    """every sequence the solver produces stays executable-equivalent: same program ops in the
    same order, and every remaining wait still has its record before it"""
    g = tz.Graph()
    a, b, c = K["a"], K["b"], K["c"]
    g.start_then(a)
    g.then(a, b)
    g.then(a, c)
    g.then_finish(b)
    g.then_finish(c)
    plat = tz.Platform.make_null(2)
    for seq in tz.dfs.get_all_sequences(g, plat):
        before = [o.desc() for o in seq if not o.kind().startswith("Cuda")]
        _C.remove_redundant_syncs(seq)
        after = [o.desc() for o in seq if not o.kind().startswith("Cuda")]
        assert before == after
        recorded = set()
        for o in seq:
            if o.kind() == "CudaEventRecord":
                recorded.add(o.event())
            elif o.kind() in ("CudaEventSync", "CudaStreamWaitEvent"):
                assert o.event() in recorded
        # Finish is still host-synchronised with both leaves
        assert seq[-1].name() == "Finish"


def _linear_extensions(n, edges):
    count = 0
    for perm in itertools.permutations(range(n)):
        pos = {v: i for i, v in enumerate(perm)}
        if all(pos[a] < pos[b] for a, b in edges):
            count += 1
    return count


def test_make_schedules_enumerates_all_topological_orders():
    ops = [tz.NoOp(f"n{i}") for i in range(5)]
    edges = [(0, 1), (0, 2), (1, 3), (2, 3)]  # n4 is free
    g = tz.Graph()
    for o in ops:
        g.start_then(o)
        g.then_finish(o)
    for a, b in edges:
        g.then(ops[a], ops[b])
    want = _linear_extensions(5, edges)
    scheds = _C.make_schedules(g)
    assert len(scheds) == want == _C.count_schedules(g)
    assert len({tuple(s.names()) for s in scheds}) == want
    for s in scheds:
        n = s.names()
        assert n[0] == "Start" and n[-1] == "Finish"
        for a, b in edges:
            assert n.index(f"n{a}") < n.index(f"n{b}")
    assert len(_C.make_schedules(g, 3)) == 3


def test_make_schedules_random_is_valid_and_roughly_uniform():
    ops = [tz.NoOp(f"n{i}") for i in range(4)]
    g = tz.Graph()
    for o in ops:
        g.start_then(o)
        g.then_finish(o)
    g.then(ops[0], ops[1])  # 12 linear extensions
    seen = {}
    for s in _C.make_schedules_random(g, 2400, 7):
        n = tuple(s.names())
        assert n.index("n0") < n.index("n1")
        seen[n] = seen.get(n, 0) + 1
    assert len(seen) == 12
    assert min(seen.values()) > 120 and max(seen.values()) < 290  # ~200 each


def test_insert_synchronization_makes_edges_explicit():
    g = tz.Graph()
    a, b = K["a"], K["b"]
    h = tz.NoOp("h")
    g.start_then(a)
    g.then(a, b)
    g.then(b, h)
    g.then_finish(h)
    g = _C.apply_assignment(g, [("a", 0), ("b", 1)])
    s = _C.insert_synchronization(g)
    kinds = sorted(v.kind() for v in s.vertices())
    assert kinds.count("CudaEventRecord") == 2 and kinds.count("CudaStreamWaitEvent") == 1 \
        and kinds.count("CudaEventSync") == 1
    assert s.vertex_size() == g.vertex_size() + 4
