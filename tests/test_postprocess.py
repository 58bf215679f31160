"""Design-rule extraction on synthetic search results with a known rule."""
import tenzing_b200 as tz
from tenzing_b200 import _C, postprocess


def test_rules_recover_the_planted_structure(tmp_path):
    # six host ops; the time is decided by one ordering (f0 before s0) and, for the GPU ops,
    # by whether a and b share a stream
    a, b = tz.EmptyGpuOp("a"), tz.EmptyGpuOp("b")
    g = tz.Graph()
    for n in ("f0", "s0", "x"):
        o = tz.NoOp(n)
        g.start_then(o)
        g.then_finish(o)
    for o in (a, b):
        g.start_then(o)
        g.then_finish(o)
    plat = tz.Platform.make_null(2)

    def cost(seq):
        n = seq.names()
        st = {o.name(): o.stream() for o in seq if o.kind() == "BoundGpuOp"}
        t = 1.0 if n.index("f0") < n.index("s0") else 2.0
        if st["a"] == st["b"]:
            t += 4.0
        return t * 1e-3

    opts = _C.mcts.Opts()
    opts.n_iters = 400
    opts.seed = 5
    res = _C.mcts.explore(g, plat, tz.PyBenchmarker(cost), "Coverage", opts)
    csv = res.csv()
    times, seqs = postprocess.read_results(csv)
    assert times.shape == (400, 6) and len(seqs) == 400
    out = postprocess.analyse(csv, prominence_pct=90.0)
    assert len(set(out["labels"])) == 4
    assert out["accuracy"] == 1.0
    text = out["text"]
    assert "f0 before s0" in text and "a same stream as b" in text
    # irrelevant op x never appears in a rule
    assert " x" not in text.replace("x before", "@").replace("before x", "@") or "@" not in text
    ev = postprocess.evaluate_rules(csv, prefixes=(50, 100), prominence_pct=90.0)
    assert ev and max(ev.values()) == 1.0
    p = tmp_path / "r.csv"
    p.write_text(csv)
    assert postprocess.main([str(p), str(tmp_path / "out.")]) == 0
    assert (tmp_path / "out.rules.txt").exists() and (tmp_path / "out.tree.dot").exists()
    # the picture of the classes (the reference draws it with matplotlib; here a plain SVG)
    import xml.etree.ElementTree as ET
    svg = ET.parse(tmp_path / "out.classes.svg").getroot()
    ns = "{http://www.w3.org/2000/svg}"
    dashed = [l for l in svg.iter(ns + "line") if l.get("stroke-dasharray")]
    assert len(dashed) == 3                               # 4 classes -> 3 boundaries
    assert len(next(svg.iter(ns + "polyline")).get("points").split()) == 400


def test_feature_pruning_drops_constant_and_duplicate_columns():
    import numpy as np
    X = np.array([[1, 0, 1, 0], [1, 1, 1, 0], [1, 0, 1, 1]], dtype=bool)
    Xp, cols = postprocess.prune_features(X, ["const", "u", "const2", "v"])
    assert cols == ["u", "v"] and Xp.shape == (3, 2)
    X = np.array([[1, 0], [0, 1]], dtype=bool)  # a column and its negation
    assert postprocess.prune_features(X, ["p", "not p"])[1] == ["p"]


def test_cli_runs_without_a_gpu(capsys):
    """python -m tenzing_b200.cli: reproducibility header + CSV, like the reference's executables"""
    import json
    from tenzing_b200 import cli
    assert cli.main(["spmv", "--solver", "dfs", "-m", "300", "--max-seqs", "4", "--style", "fused",
                     "--null"]) == 0
    out = capsys.readouterr().out.splitlines()
    assert set(json.loads(out[0])) == {"args", "hash", "major", "minor", "patch"}
    assert out[1] == '{"dfs__Opts":{"maxSeqs":4}}'
    assert len([l for l in out[2:] if "|" in l]) == 4


def test_cli_resumes_an_interrupted_search(capsys, tmp_path):
    """--resume: the second run retraces the first one's rows (same sequences, same numbers)
    from the file and only measures the additional iterations."""
    from tenzing_b200 import cli
    base = ["halo", "--solver", "mcts", "-n", "8", "--nq", "1", "--ghost", "1", "--neighbors", "6",
            "--style", "fused_mixed", "--null", "--seed", "3", "-b", "2"]
    assert cli.main(base + ["-i", "6"]) == 0
    first = capsys.readouterr().out
    path = tmp_path / "first.csv"
    path.write_text(first)
    assert cli.main(base + ["-i", "9", "--resume", str(path)]) == 0
    second = capsys.readouterr().out
    rows = lambda text: [l.split("|", 1)[1] for l in text.splitlines() if l[:1].isdigit()]
    assert len(rows(first)) == 6 and len(rows(second)) == 9
    assert rows(second)[:6] == rows(first)
