"""Broadcast format vs. ChoiceOps whose alternatives share op names (every SpMV style has a "yl"):
a sequence that went through JSON must come back as the ops of the alternatives that were
actually chosen.  Checked by execution: every op logs which alternative it belongs to."""
import random

import tenzing_b200 as tz
from tenzing_b200 import _C


def _variant(tag, names, log, gpu_names=()):
    """compound op `tag` whose ops are called `names` (shared with the other variants)"""
    g = tz.Graph()
    prev = None
    for n in names:
        body = (lambda t=tag, nm=n: (lambda *_: log.append((t, nm))))()
        op = tz.PyGpuOp(n, body) if n in gpu_names else tz.PyCpuOp(n, body)
        if prev is None:
            g.start_then(op)
        else:
            g.then(prev, op)
        prev = op
    g.then_finish(prev)
    return tz.CompoundOp(tag, g)


def _program(log):
    # two choices in sequence; inside each, the alternatives reuse names; one alternative is a
    # strict superset of another (the SpMV "fused" / "fused_separate" situation)
    first = tz.ChoiceOp("first", [_variant("f_short", ["load", "mul"], log, gpu_names=("mul",)),
                                  _variant("f_long", ["load", "mul", "add"], log, gpu_names=("mul",)),
                                  _variant("f_other", ["load", "fma"], log)])
    second = tz.ChoiceOp("second", [_variant("s_a", ["pack", "send"], log, gpu_names=("pack",)),
                                    _variant("s_b", ["pack", "send", "ack"], log, gpu_names=("pack",))])
    g = tz.Graph()
    g.start_then(first)
    g.then(first, second)
    g.then_finish(second)
    return g


def test_round_trip_through_json_keeps_the_chosen_alternatives():
    rng = random.Random(11)
    plat = tz.Platform.make_null(2)
    seen = set()
    for trial in range(60):
        log = []
        g = _program(log)                  # what rank 0 searches
        st = tz.State(g)
        while True:
            ds = st.get_decisions(plat)
            if not ds:
                break
            st = st.apply(rng.choice(ds))
        seq = st.sequence()
        seq.run(plat)
        want = list(log)
        tags = {t for t, _ in want}
        assert len(tags) == 2              # one alternative of each choice ran
        seen |= tags

        # "another rank": its own copy of the program, ops resolved from the JSON text
        log2 = []
        g2 = _program(log2)
        back = _C.sequence_from_json(seq.to_json(st.graph()), g2)
        assert back.names() == seq.names()
        back.run(plat)
        assert log2 == want, (seq.names(), log2, want)
    assert seen == {"f_short", "f_long", "f_other", "s_a", "s_b"}


def test_csv_replay_resolves_rows_jointly_too():
    log = []
    g = _program(log)
    plat = tz.Platform.make_null(1)
    res = tz.dfs.explore(g, plat, tz.PyBenchmarker(lambda s: float(len(s))))
    assert len(res.sim_results) >= 6       # 3 x 2 combinations of alternatives at least
    replay = tz.CsvBenchmarker.from_text(res.csv(), _program([]))
    for r in res.sim_results:
        assert replay.benchmark(r.seq, plat).pct10 == r.bench_result.pct10
