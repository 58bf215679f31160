"""Property test of the semantic heart: for random programs (mixed CPU / GPU ops, random DAGs,
1-3 streams) EVERY sequence the solver enumerates - before and after the redundant-sync
peephole pass - must respect every dependency of the program when executed with arbitrary op
durations.  Execution is the timeline simulation (streams in order, events, host syncs), so a
missing or wrongly removed synchronisation shows up as a successor starting before its
predecessor has finished for some choice of durations."""
import random

from hypothesis import HealthCheck, given, settings, strategies as st

import tenzing_b200 as tz
from tenzing_b200 import _C

US = 1e-6


@st.composite
def programs(draw):
    n = draw(st.integers(min_value=2, max_value=5))
    gpu = [draw(st.booleans()) for _ in range(n)]
    edges = [(i, j) for i in range(n) for j in range(i + 1, n) if draw(st.integers(0, 99)) < 45]
    streams = draw(st.integers(min_value=1, max_value=3))
    seed = draw(st.integers(min_value=0, max_value=10**6))
    return n, gpu, edges, streams, seed


@settings(max_examples=40, deadline=None, suppress_health_check=[HealthCheck.too_slow])
@given(programs())
def test_every_enumerated_schedule_respects_the_program_order(prog):
    n, gpu, edges, n_streams, seed = prog
    ops = [tz.EmptyGpuOp(f"g{i}") if gpu[i] else tz.NoOp(f"c{i}") for i in range(n)]
    g = tz.Graph()
    has_pred, has_succ = set(), set()
    for a, b in edges:
        g.then(ops[a], ops[b])
        has_succ.add(a)
        has_pred.add(b)
    for i in range(n):
        if i not in has_pred:
            g.start_then(ops[i])
        if i not in has_succ:
            g.then_finish(ops[i])
    plat = tz.Platform.make_null(n_streams)
    rng = random.Random(seed)
    seqs = tz.dfs.get_all_sequences(g, plat, 120)
    assert seqs
    for seq in seqs:
        for cleaned in (False, True):
            if cleaned:
                _C.remove_redundant_syncs(seq)
            names = seq.names()
            assert names[0] == "Start" and names[-1] == "Finish"
            # three random duration assignments per schedule
            for _ in range(3):
                gpu_t = {o.name(): rng.choice([1, 5, 50, 200]) * US for o in ops}
                cpu_t = {o.name(): rng.choice([0, 1, 20]) * US for o in ops}
                sim = tz.SimBenchmarker(gpu_t, cpu_t, launch_cost=rng.choice([0.5, 3.0]) * US)
                spans = dict(zip(names, sim.timeline(seq)))  # program op names are unique
                for a, b in edges:
                    ea, sb = spans[ops[a].name()][1], spans[ops[b].name()][0]
                    assert ea <= sb + 1e-12, (f"{ops[b].name()} starts before {ops[a].name()} ends "
                                              f"in {seq.desc()} (cleaned={cleaned})")
                finish_start = spans["Finish"][0]
                for i in range(n):
                    assert spans[ops[i].name()][1] <= finish_start + 1e-12, \
                        f"Finish before {ops[i].name()} completed in {seq.desc()} (cleaned={cleaned})"


def test_schedules_respect_the_program_order_when_really_executed_on_stream_threads():
    """The same property, checked by execution instead of simulation: streams are worker threads
    with CUDA's event semantics (Platform.enable_host_streams), ops are Python callables that
    log when they start and end.  A successor must never start before its predecessor ended."""
    import threading
    import time
    rng = random.Random(2024)
    checked = 0
    for trial in range(12):
        n = rng.randint(3, 6)
        gpu = [rng.random() < 0.7 for _ in range(n)]
        edges = [(i, j) for i in range(n) for j in range(i + 1, n) if rng.random() < 0.4]
        n_streams = rng.randint(2, 3)
        log, lock = {}, threading.Lock()

        def body(name, dur):
            def run(*_):
                t0 = time.perf_counter()
                time.sleep(dur)
                with lock:
                    log[name] = (t0, time.perf_counter())
            return run

        ops = []
        for i in range(n):
            dur = rng.choice([0.0002, 0.001, 0.003])   # predecessors are often the slow ones
            name = f"{'g' if gpu[i] else 'c'}{i}"
            ops.append(tz.PyGpuOp(name, body(name, dur)) if gpu[i] else tz.PyCpuOp(name, body(name, dur)))
        g = tz.Graph()
        has_pred, has_succ = set(), set()
        for a, b in edges:
            g.then(ops[a], ops[b])
            has_succ.add(a)
            has_pred.add(b)
        for i in range(n):
            if i not in has_pred:
                g.start_then(ops[i])
            if i not in has_succ:
                g.then_finish(ops[i])
        plat = tz.Platform.make_null(n_streams)
        plat.enable_host_streams()
        seqs = tz.dfs.get_all_sequences(g, plat, 400)
        for seq in rng.sample(seqs, min(5, len(seqs))):
            _C.remove_redundant_syncs(seq)
            log.clear()
            seq.run(plat)
            plat.synchronize()
            assert len(log) == n
            for a, b in edges:
                assert log[ops[a].name()][1] <= log[ops[b].name()][0], (seq.names(), a, b)
            checked += 1
    assert checked >= 30
