"""DFS and MCTS on cost models (no GPU): enumeration counts against brute force, the
fast/slow toy with a known optimum (reference: tenzing-mcts/src_mcts_test/mcts.cpp:23-67),
all nine strategies, CSV dump + replay (CsvBenchmarker), tree bookkeeping."""
import itertools
import math

import pytest

import tenzing_b200 as tz
from tenzing_b200 import _C


def diamond_cpu(n_free=0):
    ops = [tz.NoOp(f"n{i}") for i in range(4 + n_free)]
    edges = [(0, 1), (0, 2), (1, 3), (2, 3)]
    g = tz.Graph()
    for o in ops:
        g.start_then(o)
        g.then_finish(o)
    for a, b in edges:
        g.then(ops[a], ops[b])
    return g, ops, edges


def brute_force_orders(n, edges):
    out = []
    for perm in itertools.permutations(range(n)):
        pos = {v: i for i, v in enumerate(perm)}
        if all(pos[a] < pos[b] for a, b in edges):
            out.append(perm)
    return out


def test_dfs_noop_graph_world_of_one():
    g = tz.Graph()
    op1 = tz.NoOp("op1")
    g.start_then(op1)
    g.then_finish(op1)
    plat = tz.Platform.make_null(1)
    seqs = tz.dfs.get_all_sequences(g, plat)
    assert [s.names() for s in seqs] == [["Start", "op1", "Finish"]]


def test_dfs_cpu_graph_counts_match_linear_extensions():
    g, ops, edges = diamond_cpu(n_free=2)
    plat = tz.Platform.make_null(2)
    seqs = tz.dfs.get_all_sequences(g, plat)
    want = brute_force_orders(6, edges)
    assert len(seqs) == len(want)
    got = {tuple(int(n[1:]) for n in s.names()[1:-1]) for s in seqs}
    assert got == set(want)
    assert len(tz.dfs.get_all_sequences(g, plat, 5)) == 5


def test_dfs_gpu_graph_sequences_are_valid_and_distinct():
    k1, k2, k3 = (tz.EmptyGpuOp(n) for n in ("k1", "k2", "k3"))
    g = tz.Graph()
    g.start_then(k1)
    g.then(k1, k2)
    g.then(k1, k3)
    g.then_finish(k2)
    g.then_finish(k3)
    one = tz.dfs.get_all_sequences(g, tz.Platform.make_null(1))
    two = tz.dfs.get_all_sequences(g, tz.Platform.make_null(2))
    assert len(two) > len(one) >= 2
    keys = [s.canonical_key() for s in two]
    assert len(set(keys)) == len(keys)
    for s in two:
        names = s.names()
        assert names[0] == "Start" and names[-1] == "Finish"
        assert names.index("k1") < names.index("k2") and names.index("k1") < names.index("k3")
        ops = list(s)
        stream = {o.name(): o.stream() for o in ops if o.kind() == "BoundGpuOp"}
        # canonical stream use: the first GPU op is always on stream 0
        assert stream["k1"] == 0
        for child in ("k2", "k3"):
            if stream[child] != stream["k1"]:
                i = names.index(child)
                waits = [o for o in ops[:i] if o.kind() == "CudaStreamWaitEvent"
                         and o.stream() == stream[child]]
                assert waits, f"{child} runs on another stream without a stream wait: {s.desc()}"
    # with the peephole pass the set can only shrink
    cleaned = tz.dfs.get_all_sequences(g, tz.Platform.make_null(2), -1, True)
    assert len(cleaned) <= len(two)


def fast_slow_cost(seq):
    """ops f0..f2 are cheap when run before their slow partner s_i, expensive after"""
    names = seq.names()
    t = 0.0
    for i in range(3):
        t += 5e-3 if names.index(f"f{i}") < names.index(f"s{i}") else 10e-3
    return t


def fast_slow_graph():
    g = tz.Graph()
    for i in range(3):
        for n in (f"f{i}", f"s{i}"):
            o = tz.NoOp(n)
            g.start_then(o)
            g.then_finish(o)
    return g


def test_dfs_explore_finds_the_optimum_and_dumps_csv():
    g = fast_slow_graph()
    plat = tz.Platform.make_null(1)
    bench = tz.PyBenchmarker(fast_slow_cost)
    res = tz.dfs.explore(g, plat, bench)
    assert len(res.sim_results) == math.factorial(6)
    best = res.sim_results[res.best()]
    assert best.bench_result.pct10 == pytest.approx(15e-3)
    lines = res.csv().splitlines()
    assert lines[0] == '{"dfs__Opts":{"maxSeqs":-1}}'
    cols = lines[1].split("|")
    assert cols[0] == "0" and len(cols) == 7 + 8 and cols[7] == '{"name":"Start"}'
    # replay: the CSV is a benchmarker
    replay = tz.CsvBenchmarker.from_text(res.csv(), g)
    assert replay.num_rows() == 720
    seq = res.sim_results[17].seq
    r = replay.benchmark(seq, plat)
    assert r.pct10 == pytest.approx(res.sim_results[17].bench_result.pct10)


@pytest.mark.parametrize("strategy", _C.mcts.strategy_names())
def test_every_mcts_strategy_runs_and_improves_on_random_first_guess(strategy):
    g = fast_slow_graph()
    plat = tz.Platform.make_null(1)
    opts = _C.mcts.Opts()
    opts.n_iters = 150
    opts.seed = 3
    res = _C.mcts.explore(g, plat, tz.PyBenchmarker(fast_slow_cost), strategy, opts)
    assert len(res.sim_results) == 150
    times = [r.bench_result.pct10 for r in res.sim_results]
    assert min(times) == pytest.approx(15e-3), f"{strategy} never found the optimum"
    for r in res.sim_results:
        assert r.path.names()[0] == "Start" and r.path.names()[-1] == "Finish"
    # header-less CSV rows
    assert res.csv().splitlines()[0].startswith("0|")


def test_fastmin_concentrates_on_fast_subtrees():
    g = fast_slow_graph()
    plat = tz.Platform.make_null(1)
    opts = _C.mcts.Opts()
    opts.n_iters = 300
    opts.seed = 1

    def mean_tail(strategy):
        res = _C.mcts.explore(g, plat, tz.PyBenchmarker(fast_slow_cost), strategy, opts)
        t = [r.bench_result.pct10 for r in res.sim_results][100:]
        return sum(t) / len(t)

    assert mean_tail("FastMin") < mean_tail("Random")


def test_mcts_stops_when_the_tree_is_fully_visited():
    g, ops, edges = diamond_cpu()
    plat = tz.Platform.make_null(1)
    opts = _C.mcts.Opts()
    opts.n_iters = 0  # until fully visited
    res = _C.mcts.explore(g, plat, tz.PyBenchmarker(lambda s: 1.0), "Coverage", opts)
    assert len({tuple(r.path.names()) for r in res.sim_results}) == len(brute_force_orders(4, edges)) == 2
    assert len(res.sim_results) <= 4


def test_mcts_tree_bookkeeping_parent_links_and_graphviz():
    g = fast_slow_graph()
    plat = tz.Platform.make_null(1)
    root = _C.mcts.Node(g)
    strat = _C.mcts.make_strategy("FastMin")
    for i in range(25):
        seq = _C.mcts.iterate(root, plat, strat, fast_slow_cost, i, True)
        assert seq.names()[-1] == "Finish"
    assert root.n == 25 and len(root.times()) == 25
    assert root.size() > 25 and not root.fully_visited
    kids = root.children()
    assert len(kids) == 6 and sum(k.n for k in kids) == 25
    for k in kids:
        assert k.parent().desc() == "Start"
        if k.n:
            deep = k
            while deep.children() and any(c.n for c in deep.children()):
                deep = [c for c in deep.children() if c.n][0]
            assert deep.get_sequence().names()[:2] == ["Start", k.op().name()]
            assert deep.root().desc() == "Start"
    dot = root.graphviz()
    assert dot.startswith("digraph D {") and "rollouts: 25" in dot


def test_mcts_without_rollout_expansion_keeps_the_tree_small():
    g = fast_slow_graph()
    plat = tz.Platform.make_null(1)
    big, small = _C.mcts.Node(g), _C.mcts.Node(g)
    s1, s2 = _C.mcts.make_strategy("FastMin"), _C.mcts.make_strategy("FastMin")
    for i in range(20):
        _C.mcts.iterate(big, plat, s1, fast_slow_cost, i, True)
        _C.mcts.iterate(small, plat, s2, fast_slow_cost, i, False)
    assert small.size() < big.size()
    assert small.n == big.n == 20


def test_histogram_and_strategy_names():
    assert _C.mcts.histogram([0.0, 0.1, 0.55, 1.0], 0.0, 1.0) == [1, 1, 0, 0, 0, 1, 0, 0, 0, 1]
    assert len(_C.mcts.strategy_names()) == 9
    with pytest.raises(RuntimeError):
        _C.mcts.make_strategy("nope")


def test_search_with_gpu_ops_on_a_null_platform_uses_streams_as_ids():
    """choice + compound + stream binding + sync insertion, all under MCTS, no GPU needed"""
    inner = tz.Graph()
    a, b = tz.EmptyGpuOp("a"), tz.EmptyGpuOp("b")
    inner.start_then(a)
    inner.start_then(b)
    inner.then_finish(a)
    inner.then_finish(b)
    fused = tz.EmptyGpuOp("ab_fused")
    choice = tz.ChoiceOp("impl", [tz.CompoundOp("split", inner), fused])
    g = tz.Graph()
    g.start_then(choice)
    g.then_finish(choice)
    plat = tz.Platform.make_null(2)

    def cost(seq):
        n = seq.names()
        if "ab_fused" in n:
            return 2.0
        streams = {o.name(): o.stream() for o in seq if o.kind() == "BoundGpuOp"}
        return 1.0 if streams["a"] != streams["b"] else 3.0

    opts = _C.mcts.Opts()
    opts.n_iters = 60
    res = _C.mcts.explore(g, plat, tz.PyBenchmarker(cost), "FastMin", opts)
    best = res.sim_results[res.best()]
    assert best.bench_result.pct10 == 1.0
    allseq = tz.dfs.get_all_sequences(g, plat, -1, True)
    assert min(cost(s) for s in allseq) == 1.0


def test_mcts_stops_on_a_large_tree():
    g = fast_slow_graph()
    plat = tz.Platform.make_null(1)
    opts = _C.mcts.Opts()
    opts.n_iters = 10_000
    opts.max_tree_nodes = 60
    res = _C.mcts.explore(g, plat, tz.PyBenchmarker(fast_slow_cost), "FastMin", opts)
    assert 0 < len(res.sim_results) < 60


def test_interrupt_dumps_the_partial_results():
    """SIGINT during a search makes rank 0 print the rows collected so far and exit(1)
    (reference: src/trap.cpp:11-35 + the solvers' register_handler calls)"""
    import os
    import signal
    import subprocess
    import sys
    import time as _t
    here = os.path.dirname(os.path.abspath(__file__))
    p = subprocess.Popen([sys.executable, os.path.join(here, "trap_worker.py")], stdout=subprocess.PIPE,
                         stderr=subprocess.PIPE, text=True)
    assert p.stdout.readline().strip() == "READY"
    _t.sleep(1.0)
    p.send_signal(signal.SIGINT)
    out, _ = p.communicate(timeout=60)
    assert p.returncode == 1
    rows = [line for line in out.splitlines() if "|" in line]
    assert len(rows) > 10 and rows[0].startswith("0|") and '{"name":"Start"}' in rows[0]


def test_tree_dumps_graph_dumps_and_csv_file_replay(tmp_path):
    g = fast_slow_graph()
    plat = tz.Platform.make_null(1)
    opts = _C.mcts.Opts()
    opts.n_iters = 30
    opts.dump_tree = True
    opts.dump_tree_prefix = str(tmp_path / "tree_")
    res = _C.mcts.explore(g, plat, tz.PyBenchmarker(fast_slow_cost), "AvgTime", opts)
    # iterations 0..9, 10, 20 (reference: mcts.hpp:302-303)
    dumped = sorted(int(p.name[5:-4]) for p in tmp_path.glob("tree_*.dot"))
    assert dumped == list(range(10)) + [10, 20]
    assert (tmp_path / "tree_20.dot").read_text().startswith("digraph D {")
    g.dump_graphviz(str(tmp_path / "graph.dot"))
    assert "f0" in (tmp_path / "graph.dot").read_text()
    # a results file on disk, with the reproducibility header in front, replays
    path = tmp_path / "results.csv"
    path.write_text(_C.reproduce_header(["prog"]) + "\n" + res.csv())
    replay = tz.CsvBenchmarker(str(path), g)
    assert replay.num_rows() == 30
    prod = replay.producer()
    assert prod["args"] == ["prog"] and prod["same_build"] and prod["version"] == _C.version_string()
    assert tz.CsvBenchmarker.from_text(res.csv(), g).producer() is None
    assert _C.parse_reproduce_header('{"not": "a header"}') is None
    assert _C.parse_reproduce_header("12|3|4") is None
    other = _C.parse_reproduce_header('{"args":["a","b"],"hash":"deadbeef","major":9,"minor":0,"patch":1}')
    assert other["version"] == "9.0.1-deadbeef" and not other["same_build"]
    r0 = res.sim_results[0]
    assert replay.benchmark(r0.path, plat).pct10 == pytest.approx(r0.bench_result.pct10)
    with pytest.raises(RuntimeError):
        replay.benchmark(tz.Sequence([g.start(), g.finish()]), plat)


def test_an_interrupted_search_resumes_from_its_csv_without_remeasuring():
    """CachingBenchmarker: preload the CSV of a 25-iteration run and search again with the same
    seed for 60 iterations.  The measurement is noisy (every call returns a different time, as on
    hardware) and the search visits some schedules more than once; the first 25 iterations must
    still retrace the interrupted run exactly, without calling the benchmark function."""
    g = fast_slow_graph()
    plat = tz.Platform.make_null(1)
    calls = []

    def noisy(seq):
        calls.append(1)
        return fast_slow_cost(seq) * (1.0 + 0.013 * ((len(calls) * 7919) % 101) / 101.0)

    def run(iters, bench):
        opts = _C.mcts.Opts()
        opts.n_iters = iters
        opts.seed = 5
        return _C.mcts.explore(g, plat, bench, "FastMin", opts)

    first = run(25, tz.PyBenchmarker(noisy))
    assert len(calls) == 25
    rows = lambda res: [(r.path.names(), r.bench_result.pct10) for r in res.sim_results]

    # resume without reuse: recorded results are replayed in order, everything else is measured
    replay = tz.CachingBenchmarker(tz.PyBenchmarker(noisy), reuse=False)
    replay.preload_csv(tz.CsvBenchmarker.from_text(first.csv(), g))
    assert replay.pending() == 25
    resumed = run(60, replay)
    assert rows(resumed)[:25] == rows(first)            # exact, including the float values
    assert replay.hits() == 25 and replay.misses() == 35 and replay.pending() == 0
    assert len(calls) == 25 + 35

    # reuse: no schedule is ever measured twice
    del calls[:]
    cache = tz.CachingBenchmarker(tz.PyBenchmarker(noisy), reuse=True)
    deduped = run(60, cache)
    distinct = {tuple(n) for n, _ in rows(deduped)}
    assert len(calls) == cache.misses() == len(distinct) == cache.size()
    assert cache.hits() == 60 - len(distinct)
    by_seq = {}
    for n, t in rows(deduped):
        assert by_seq.setdefault(tuple(n), t) == t      # repeated schedules report the first result
