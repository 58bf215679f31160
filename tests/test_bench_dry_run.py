"""bench.py end to end without a GPU: `--host-dry-run` runs the same control flow (search with a
wall-clock budget agreed between ranks, timed steps, end-to-end loop, clock sampling, the JSON
line) with both workloads executed on the CPU, launched under torchrun + gloo exactly as the
driver launches the GPU run.  Its number is not a result; the point is that the script the
round is judged by cannot break unnoticed on a box without a GPU."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ARGS = ["--host-dry-run", "--nx", "8", "--spmv-n", "3000", "--spmv-big-n", "4000", "--steps", "4", "--warmup", "3",
        "--search-budget", "5", "--search-iters", "5"]
KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
        "scaling", "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches"}


def _check(line, n):
    d = json.loads(line)
    assert KEYS <= set(d), KEYS - set(d)
    assert d["n_gpus"] == n and d["steps"] == 4 and d["warmup"] == 3 and d["correct"] is True
    assert d["higher_is_better"] is False and d["scaling"] == "strong" and d["value"] > 0
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step", "median", "p10", "p90"} <= set(d["e2e"])
    # rank 0's block of x goes in and its block of y comes out, every step
    assert d["e2e"]["h2d_bytes_per_step"] == 3000 // n * 8 and d["e2e"]["d2h_bytes_per_step"] == 3000 // n * 8
    assert d["schedule"][0] == "Start" and d["schedule"][-1] == "Finish"
    # the other BASELINE configs ride along as sub-records
    assert d["halo"]["correct"] is True and d["halo"]["schedule"][-1] == "Finish" and d["halo"]["us"] > 0
    assert d["spmv_1m"]["correct"] is True and d["spmv"]["fp32"]["us"] > 0
    assert "not a result" in d["impl"]


@pytest.mark.timeout(600)
def test_bench_runs_end_to_end_on_one_rank():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", TENZING_ACK_NOTICE="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + ARGS, capture_output=True,
                       text=True, timeout=500, env=env, cwd=ROOT)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    _check([l for l in p.stdout.splitlines() if l.startswith("{")][-1], 1)


@pytest.mark.timeout(600)
def test_bench_runs_end_to_end_on_two_ranks_launched_like_the_driver_does():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", TENZING_ACK_NOTICE="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + os.getpid() % 100),
           os.path.join(ROOT, "bench.py"), "--gpus", "2"] + ARGS
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=500, env=env, cwd=ROOT)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1            # rank 0 only
    _check(lines[0], 2)


def test_reference_arm_reports_unavailable_and_exits_zero():
    # without a GPU the reference's driver fails at once; the arm still has to answer cleanly
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference",
                        "--ref-seconds", "20", "--ref-grace", "5"],
                       capture_output=True, text=True, timeout=400, cwd=ROOT)
    assert p.returncode == 0
    d = json.loads(p.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["unavailable"]


def test_reference_arm_reads_the_references_csv():
    sys.path.insert(0, ROOT)
    import bench
    text = [
        '{"args":["spmv"],"hash":"abc","major":0,"minor":1,"patch":0}\n',
        '{"dfs__Opts":{"maxSeqs":15000}}\n',
        '0|6.6e-05|6.7e-05|6.8e-05|7e-05|7.1e-05|1e-06|{"name":"Start"}|{"name":"yl","stream":0}|{"name":"Finish"}\n',
        '1|6.4e-05|6.486e-05|6.5e-05|6.9e-05|7e-05|1e-06|{"name":"Start"}|{"name":"Pack","stream":1}|{"name":"Finish"}\n',
        'some log line | with | pipes\n',
    ]
    rows = bench._ref_result_rows(text)
    assert len(rows) == 2 and min(r[1] for r in rows) == 6.486e-05


def test_reference_arm_reads_the_end_to_end_line():
    sys.path.insert(0, ROOT) if ROOT not in sys.path else None
    import bench
    got = bench._ref_parse_e2e("noise\nREF_E2E median_us=447.461 p10_us=442.611 p90_us=551.057 h2d_bytes=300000 "
                               "d2h_bytes=300000 schedule_pct10_us=511.083\n")
    assert got["median_us"] == pytest.approx(447.461) and got["h2d_bytes"] == 300000
    assert bench._ref_parse_e2e("REF_E2E none\n") is None
    assert bench._ref_parse_e2e("") is None


def test_side_records_that_do_not_finish_cannot_take_the_line_down():
    """with no time at all for the side records the line is still printed, says so, and the exit
    code is 0 (two ranks: the one that is behind must leave as well)"""
    env = dict(os.environ, TENZING_ACK_NOTICE="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + os.getpid() % 100),
           os.path.join(ROOT, "bench.py"), "--gpus", "2"] + ARGS + ["--side-budget", "0.001"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=500, env=env, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["value"] > 0 and "halo" in rec and "spmv_1m" in rec
    assert "records_incomplete" in rec or "oracle_pass" in rec
