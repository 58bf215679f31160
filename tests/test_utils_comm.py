"""L0 utilities and control-plane communicators (no GPU)."""
import json
import math
import multiprocessing as mp
import os
import random
import subprocess
import sys

import pytest

import tenzing_b200 as tz
from tenzing_b200 import _C

HERE = os.path.dirname(os.path.abspath(__file__))


def test_numeric_helpers():
    v = [3.0, 1.0, 2.0, 10.0]
    assert _C.avg(v) == 4.0
    assert _C.med(v) == 2.5 and _C.med([5.0, 1.0, 3.0]) == 3.0
    assert _C.var(v) == pytest.approx(sum((x - 4.0) ** 2 for x in v) / 4)
    assert _C.stddev(v) == pytest.approx(math.sqrt(_C.var(v)))
    assert _C.corr([1, 2, 3, 4], [2, 4, 6, 8]) == pytest.approx(1.0)
    assert _C.corr([1, 2, 3, 4], [8, 6, 4, 2]) == pytest.approx(-1.0)
    assert _C.corr([1, 1, 1], [1, 2, 3]) == 0.0
    assert _C.prime_factors(360) == [5, 3, 3, 2, 2, 2]
    assert _C.prime_factors(8) == [2, 2, 2] and _C.prime_factors(1) == [] and _C.prime_factors(13) == [13]
    assert _C.round_up(13, 8) == 16 and _C.round_up(16, 8) == 16


def test_runs_test_accepts_noise_and_rejects_drift():
    rng = random.Random(0)
    noise = [rng.random() for _ in range(200)]
    assert not _C.runs_test(noise)
    drift = [i * 0.01 + rng.random() * 0.001 for i in range(200)]
    assert _C.runs_test(drift)
    assert _C.runs_test([1.0, 2.0, 3.0])  # too few samples: "does not apply" counts as reject
    assert math.isnan(_C.runs_test_z([1.0, 2.0, 3.0]))
    assert _C.runs_test_z(drift) > 1.96


def test_reproduce_header_and_version():
    h = json.loads(_C.reproduce_header(["prog", "-i", "3"]))
    assert h["args"] == ["prog", "-i", "3"] and set(h) == {"args", "hash", "major", "minor", "patch"}
    assert _C.version_string().startswith(f'{h["major"]}.{h["minor"]}.{h["patch"]}-')


def test_bench_result_percentiles_follow_the_reference_index_rule():
    s = [float(i) for i in range(100, 0, -1)]  # 100 samples, unsorted
    r = tz.BenchResult.from_samples(s)
    assert (r.pct01, r.pct10, r.pct50, r.pct90, r.pct99) == (2.0, 11.0, 51.0, 91.0, 100.0)
    assert r.stddev == pytest.approx(_C.stddev(s))


def test_partition_and_rank_grid_rules():
    sp = _C.spmv
    parts = [sp.get_partition(10, i, 4) for i in range(4)]
    assert parts == [(0, 3), (3, 6), (6, 8), (8, 10)]  # remainder rows go to the low ranks
    for i in range(10):
        lo, hi = parts[sp.get_owner(10, i, 4)]
        assert lo <= i < hi
    assert _C.halo.rank_grid(8) == [2, 2, 2] and _C.halo.rank_grid(4) == [2, 2, 1]
    assert _C.halo.rank_grid(6) == [3, 2, 1] and _C.halo.rank_grid(1) == [1, 1, 1]
    assert _C.halo.dir_to_tag(-1, -1, -1) == 0 and _C.halo.dir_to_tag(1, 1, 1) == 26


def test_empirical_benchmarker_on_null_platform_measures_host_ops():
    import time
    g = tz.Graph()
    op = tz.PyCpuOp("sleepy", lambda: time.sleep(0.002))
    g.start_then(op)
    g.then_finish(op)
    plat = tz.Platform.make_null(1)
    seq = tz.dfs.get_all_sequences(g, plat)[0]
    r = tz.EmpiricalBenchmarker().benchmark(
        seq, plat, tz.BenchOpts(n_iters=5, max_retries=0, target_secs=0.004, check_randomness=False))
    assert 0.0018 < r.pct50 < 0.01
    assert plat.epoch() >= 5


def test_loopback_world_collectives_and_lockstep_dfs():
    """config 1 of BASELINE.json: the noop-graph DFS with a world of two (no MPI: two threads)"""
    results = {}

    def body(rank, comm):
        assert comm.size() == 2 and comm.rank() == rank and comm.kind() == "loopback"
        assert comm.bcast_string("hello" if rank == 0 else "", 0) == "hello"
        assert comm.max_over_ranks(float(rank + 1)) == 2.0
        parts = comm.allgather_bytes(bytes([rank]) * (rank + 1))
        assert parts == [b"\x00", b"\x01\x01"]
        g = tz.Graph()
        op1 = tz.NoOp("op1")
        g.start_then(op1)
        g.then_finish(op1)
        plat = tz.Platform.make_null(1, comm)
        res = tz.dfs.explore(g, plat, tz.PyBenchmarker(lambda s: 1.0 + rank))
        results[rank] = [r.seq.names() for r in res.sim_results]
        comm.barrier()

    _C.loopback_run(2, body)
    assert results[0] == results[1] == [["Start", "op1", "Finish"]]


def _socket_rank(rank, size, port, q):
    os.environ["TENZING_ACK_NOTICE"] = "1"
    import tenzing_b200 as tz2
    from tenzing_b200 import _C as C
    comm = C.SocketComm(rank, size, "127.0.0.1", port, 30.0)
    out = [comm.bcast_string("x" * 1000 if rank == 1 else "", 1) == "x" * 1000,
           comm.max_over_ranks(float(rank)) == float(size - 1),
           comm.allgather_bytes(bytes([65 + rank])) == [bytes([65 + r]) for r in range(size)]]
    comm.barrier()
    # a search in lock step: rank 0 enumerates, every rank "benchmarks"
    g = tz2.Graph()
    a, b = tz2.NoOp("a"), tz2.NoOp("b")
    for o in (a, b):
        g.start_then(o)
        g.then_finish(o)
    plat = tz2.Platform.make_null(1, comm)
    res = tz2.dfs.explore(g, plat, tz2.PyBenchmarker(lambda s: 1.0))
    out.append(sorted(tuple(r.seq.names()) for r in res.sim_results) ==
               [("Start", "a", "b", "Finish"), ("Start", "b", "a", "Finish")])
    q.put((rank, out))


def test_socket_comm_three_processes():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31000 + os.getpid() % 2000
    procs = [ctx.Process(target=_socket_rank, args=(r, 3, port, q)) for r in range(3)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(all(v) for v in got.values()), got


def test_gloo_world_of_two_runs_the_noop_dfs():
    """same configuration through torch.distributed (gloo), launched like bench.py is"""
    script = os.path.join(HERE, "gloo_worker.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(29800 + os.getpid() % 100), script]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    assert "GLOO_WORKER_OK" in p.stdout, (p.stdout + p.stderr)[-3000:]


def test_native_cpp_test_runner():
    """the C++ checks that mirror the reference's doctest cases (src/apps/tenzing-cpu-tests.cpp)"""
    exe = os.path.join(os.path.dirname(HERE), "build", "bin", "tenzing-cpu-tests")
    if not os.path.exists(exe):
        import build
        build.build()
    p = subprocess.run([exe, "-tc=*[cpu]*"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "0 failed" in p.stdout and "[FAIL]" not in p.stdout
    listed = subprocess.run([exe, "-ltc"], capture_output=True, text=True, timeout=60).stdout.split("\n")
    assert len([l for l in listed if l.strip()]) >= 20
    one = subprocess.run([exe, "-tc=numeric*"], capture_output=True, text=True, timeout=60)
    assert one.returncode == 0 and "1 cases" in one.stdout


def test_json_value_round_trips_and_sorts_keys():
    rt = _C.json_roundtrip
    assert rt('{"b": 1, "a": [true, null, "x\\n", 1.5, -3]}') == '{"a":[true,null,"x\\n",1.5,-3],"b":1}'
    assert (rt('{"name":"CER-after-yl","event":0,"kind":"CudaEventRecord","stream":1}') ==
            '{"event":0,"kind":"CudaEventRecord","name":"CER-after-yl","stream":1}')
    assert rt("[]") == "[]" and rt("{}") == "{}" and rt(" 1e-05 ") == "1e-05"
    assert rt('"\\u00e9\\t"') == '"é\\t"'
    for bad in ('{"a":}', "[1,", '{"a" 1}', "tru"):
        with pytest.raises(RuntimeError):
            rt(bad)


def test_batch_benchmark_interleaves_sequences():
    import time
    g = tz.Graph()
    fast = tz.PyCpuOp("fast", lambda: time.sleep(0.001))
    slow = tz.PyCpuOp("slow", lambda: time.sleep(0.004))
    plat = tz.Platform.make_null(1)
    s_fast = tz.Sequence([g.start(), fast, g.finish()])
    s_slow = tz.Sequence([g.start(), slow, g.finish()])
    opts = tz.BenchOpts(n_iters=6, max_retries=0, target_secs=0.0, check_randomness=False)
    r_fast, r_slow = tz.EmpiricalBenchmarker().benchmark_batch([s_fast, s_slow], plat, opts)
    assert 0.0009 < r_fast.pct50 < 0.003 < r_slow.pct50 < 0.01


def test_op_helpers():
    k = tz.EmptyGpuOp("k")
    plat = tz.Platform.make_null(3)
    var = _C.make_platform_variations(plat, k)
    assert [v.stream() for v in var] == [0, 1, 2]
    n = tz.NoOp("n")
    assert [v.name() for v in _C.make_platform_variations(plat, n)] == ["n"]
    assert _C.unbound_contains(var, k) and not _C.unbound_contains(var, tz.EmptyGpuOp("other"))
    uniq = _C.keep_uniques([var[0], var[1], tz.BoundGpuOp(k, 0), n, tz.NoOp("n")])
    assert [u.desc() for u in uniq] == ["k(s0)", "k(s1)", "n"]
    assert _C.contains([k, n], tz.NoOp("n")) and not _C.contains([k], n)


def test_comm_op_library_on_a_null_platform():
    """Isend / Irecv / Wait / MultiWait / OwningWaitall / Ialltoallv (reference: mpi/ops_mpi.hpp:17-191)
    are host ops the solver can schedule; with a null transport they only exercise the structure:
    posts before waits, names kept, JSON round trip through the broadcast format."""
    tr = _C.P2PTransport(True)
    plat = tz.Platform.make_null(1)
    reqs = [_C.Request() for _ in range(4)]
    isend = _C.Isend(0, 64, 0, 0, reqs[0], tr, "isend")
    irecv = _C.Irecv(0, reqs[1], tr, "irecv")
    a2a = _C.Ialltoallv(0, [0, 64], [64, 32], [0, 0], [0, 0], reqs[2], tr, "a2a")
    wait = _C.Wait(reqs[0], "wait_send")
    multi = _C.MultiWait([reqs[1], reqs[2]], "wait_many")
    owning = _C.OwningWaitall([reqs[3]], "wait_owned")
    g = tz.Graph()
    for post in (isend, irecv, a2a):
        g.start_then(post)
        for w in (wait, multi, owning):
            g.then(post, w)      # all posts before any wait
    for w in (wait, multi, owning):
        g.then_finish(w)
    assert g.validate() == ""
    seqs = tz.dfs.get_all_sequences(g, plat, 1000, True)
    assert len(seqs) == 36           # 3! orders of the posts x 3! orders of the waits
    for seq in seqs:
        names = seq.names()
        assert max(names.index(p) for p in ("isend", "irecv", "a2a")) < \
            min(names.index(w) for w in ("wait_send", "wait_many", "wait_owned"))
        seq.run(plat)                # null transport: nothing to copy, nothing to wait for
    back = _C.sequence_from_json(seqs[0].to_json(g), g)
    assert back.names() == seqs[0].names()
    assert all(op.kind() == "CpuOp" or op.name() in ("Start", "Finish") for op in back)


def _tzrun():
    exe = os.path.join(os.path.dirname(HERE), "build", "bin", "tzrun")
    if not os.path.exists(exe):
        import build
        build.build()
    return exe


def test_tzrun_sets_the_rank_environment_and_returns_the_first_failure():
    exe = _tzrun()
    p = subprocess.run([exe, "-n", "3", "--port", "31999", "sh", "-c",
                        'echo "$RANK/$WORLD_SIZE/$LOCAL_RANK/$MASTER_ADDR/$MASTER_PORT"'],
                       capture_output=True, text=True, timeout=60)
    assert p.returncode == 0
    assert sorted(p.stdout.split()) == [f"{r}/3/{r}/127.0.0.1/31999" for r in range(3)]
    assert subprocess.run([exe], capture_output=True).returncode == 64      # usage


def test_tzrun_does_not_hang_when_a_rank_dies():
    """A dead rank would leave the others waiting in a collective forever (reference: "a rank
    failure hangs the job"); the launcher ends them after the grace period."""
    import time
    exe = _tzrun()
    t0 = time.time()
    p = subprocess.run([exe, "-n", "2", "--grace", "1", "sh", "-c",
                        'if [ "$RANK" = 1 ]; then exit 3; fi; exec sleep 120'],
                       capture_output=True, text=True, timeout=60)
    assert p.returncode == 3
    assert time.time() - t0 < 30
    assert "rank 1 ended with code 3" in p.stderr


def test_tzrun_forwards_signals_to_the_ranks():
    """SIGABRT from the batch system must reach rank 0's handler (partial CSV dump)."""
    import signal
    import time
    exe = _tzrun()
    p = subprocess.Popen([exe, "-n", "2", "sh", "-c",
                          "trap 'echo dumped-$RANK; kill $S; exit 0' ABRT; sleep 120 & S=$!; wait"],
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    time.sleep(1.0)
    p.send_signal(signal.SIGABRT)      # this exact process, not a pattern
    try:
        out, _ = p.communicate(timeout=30)
    finally:
        if p.poll() is None:
            p.kill()
    assert p.returncode == 0
    assert sorted(out.split()) == ["dumped-0", "dumped-1"]


@pytest.mark.timeout(120)
def test_cached_search_stays_in_lock_step_across_ranks():
    """MCTS through a CachingBenchmarker around the collective EmpiricalBenchmarker in a world of
    two: every rank must skip (or take) the same measurements, or the barriers inside the
    benchmarker would dead-lock."""
    out = {}

    def body(rank, comm):
        g = tz.Graph()
        for n in ("a", "b", "c"):
            o = tz.NoOp(n)
            g.start_then(o)
            g.then_finish(o)
        plat = tz.Platform.make_null(1, comm)
        cache = tz.CachingBenchmarker(tz.EmpiricalBenchmarker(), reuse=True)
        opts = _C.mcts.Opts()
        opts.n_iters = 20
        opts.seed = 1
        opts.bench_opts = _C.BenchOpts(n_iters=2, max_retries=0, target_secs=1e-5, check_randomness=False)
        res = _C.mcts.explore(g, plat, cache, "Random", opts)
        out[rank] = (cache.hits(), cache.misses(), [r.path.names() for r in res.sim_results],
                     [r.bench_result.pct10 for r in res.sim_results])
        comm.barrier()

    _C.loopback_run(2, body)
    assert out[0] == out[1]                      # same hits, same misses, same rows, same numbers
    hits, misses, rows, _ = out[0]
    assert misses == len({tuple(r) for r in rows}) <= 6 and hits + misses == len(rows)


@pytest.mark.timeout(300)
def test_native_drivers_execute_on_the_cpu_with_one_process_per_rank():
    """tzrun + SocketComm + host execution: the native search drivers really run both workloads
    (host heap shared between the processes through POSIX shared memory), including the search
    over the ChoiceOp of all implementation styles."""
    tzrun = _tzrun()
    bindir = os.path.dirname(tzrun)
    env = dict(os.environ, TENZING_ACK_NOTICE="1", TENZING_HOST_WAIT_SECS="10")
    port = 30100 + os.getpid() % 300
    runs = [
        (2, ["spmv-min-time", "-i", "10", "-b", "2", "-m", "2000", "--host-exec"], 10),
        # DFS: up to 12 sequences, fewer after equivalent ones are dropped
        (3, ["spmv", "--max-seqs", "12", "-b", "2", "-m", "2000", "--host-exec", "--style", "fused"], None),
        (2, ["halo-mcts-min-time", "-i", "6", "-b", "2", "-n", "8", "--host-streams"], 6),
    ]
    for k, (n, cmd, rows) in enumerate(runs):
        p = subprocess.run([tzrun, "-n", str(n), "--grace", "3", "--port", str(port + k),
                            os.path.join(bindir, cmd[0])] + cmd[1:],
                           capture_output=True, text=True, timeout=200, env=env)
        assert p.returncode == 0, (cmd, p.stderr[-2000:])
        data = [l for l in p.stdout.splitlines() if l[:1].isdigit()]
        assert (len(data) == rows) if rows else (0 < len(data) <= 12), (cmd, p.stdout[-1000:])


@pytest.mark.timeout(60)
def test_socket_rendezvous_times_out_instead_of_waiting_for_ever():
    """a rank that never starts: rank 0 gives up in accept, the others give up connecting"""
    import time
    port = 32200 + os.getpid() % 300
    for rank in (1, 0):
        t0 = time.time()
        with pytest.raises(RuntimeError):
            _C.SocketComm(rank, 2, "127.0.0.1", port + rank, 0.5)
        assert time.time() - t0 < 10
