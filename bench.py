#!/usr/bin/env python
"""Headline benchmark: best-schedule time/iter (us, max over ranks) of the distributed SpMV at
the size the reference's stock driver hard-codes (150000 x 150000 band, ~10 nnz/row; BASELINE.json
config 2), the one workload both arms can run; reported alongside in the same JSON line: the 3-D
halo exchange (config 3: 512^3 fp64 cells per rank, 3 quantities, ghost radius 3, 26 neighbours),
the 1M x 1M fp64 band SpMV (config 2 as named) and the random-sparse SpMV under MCTS with three
streams (config 4).

  python bench.py --gpus N --steps K --warmup W        (N > 1: launched under torch.distributed.run)
  python bench.py --impl reference --gpus N            the unmodified reference, see reference_arm()

A step is one execution of the best schedule the solver found (the search itself is set-up and
is not timed).  Every step is timed on the device with CUDA events, L2 is flushed between
steps, the per-step time is the max over ranks, and one JSON line is printed by rank 0.
"""
import argparse
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("TENZING_ACK_NOTICE", "1")

REF_DIR = os.path.join(ROOT, "baseline", "_ref")
REF_METRIC = "best-schedule distributed SpMV time/iter (us, max over ranks)"


def _ref_job_id():
    """one id per launch, the same on every rank: the MPI stand-in names its shared segment after it"""
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return f"{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}"  # the launcher is the common parent
    return f"solo_{os.getpid()}"


def _ref_ensure_built(local_rank):
    """baseline/_ref/ holds the unmodified reference built by baseline/build_ref.py; it is rebuilt
    here when missing or when it was built for another CPU (the reference compiles with
    -march=native).  One rank builds, the others wait for its marker."""
    import subprocess
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import build_ref
    exe = os.path.join(REF_DIR, "build", "tenzing-mcts", "examples", "spmv-min-time")
    marker = os.path.join(REF_DIR, "BUILD_HOST")

    def ok():
        try:
            return os.path.exists(exe) and open(marker).read().strip() == build_ref.host_id()
        except OSError:
            return False

    if ok():
        return exe, None
    if local_rank == 0:
        env = dict(os.environ, PATH=os.path.dirname(sys.executable) + os.pathsep + os.environ.get("PATH", ""))
        r = subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "build_ref.py")], env=env,
                           capture_output=True, text=True)
        if r.returncode != 0:
            return None, "reference build failed: " + (r.stdout + r.stderr)[-300:].replace("\n", " ")
    else:
        for _ in range(1200):
            if ok():
                break
            time.sleep(1)
    return (exe, None) if ok() else (None, "reference build did not complete")


def _ref_result_rows(lines):
    """[pct01, pct10, pct50, pct90, pct99, stddev] (seconds) of every result row of the reference's
    CSV: `idx|pct01|pct10|pct50|pct90|pct99|stddev|<op json>|...` (tenzing-dfs/src/dfs.cpp:84-104);
    the reproduce header and the `{"dfs__Opts":...}` line are skipped"""
    rows = []
    for line in lines:
        f = line.split("|")
        if len(f) > 7 and f[0].strip().isdigit():
            try:
                rows.append([float(v) for v in f[1:7]])
            except ValueError:
                pass
    return rows


def _ref_parse_e2e(text):
    """the `REF_E2E key=value ...` line of baseline/ref_driver/ref_e2e.cu -> dict of floats, or None"""
    for line in text.splitlines():
        if line.startswith("REF_E2E ") and "median_us=" in line:
            try:
                return {k: float(v) for k, v in (kv.split("=", 1) for kv in line.split()[1:])}
            except ValueError:
                return None
    return None


def _ref_e2e(args, rank, world, job, logdir):
    """End-to-end step of the reference (rank 0 returns the `e2e` record, or None when the driver is
    not there, fails or runs out of time - the arm's main value does not depend on it).
    baseline/ref_driver/ref_e2e.cu is a user program of the reference's public API (the set-up of
    its SpMV example, its DFS solver + EmpiricalBenchmarker over the first --ref-e2e-seqs
    schedules it enumerates, then the best of them run with the H2D of x before and the D2H of y
    after every step), built with the reference as a CMake sub-project."""
    import signal
    import subprocess
    exe = os.path.join(REF_DIR, "build_e2e", "ref-e2e")
    if not os.path.exists(exe) or args.ref_e2e_seconds <= 0:
        return None
    out_path = os.path.join(logdir, f"e2e_r{rank}.out")
    env = dict(os.environ, TENZING_ACK_NOTICE="1", TZMPI_RANK=str(rank), TZMPI_SIZE=str(world), TZMPI_JOB=job + "e")
    t0 = time.time()
    with open(out_path, "w") as fo, open(os.path.join(logdir, f"e2e_r{rank}.err"), "w") as fe:
        child = subprocess.Popen([exe, str(args.steps), str(max(args.warmup, 3)), str(args.ref_e2e_seqs)],
                                 env=env, stdout=fo, stderr=fe, start_new_session=True)
        while child.poll() is None:
            if time.time() - t0 > args.ref_e2e_seconds:
                try:
                    os.killpg(child.pid, signal.SIGKILL)  # exactly the process group started above
                except ProcessLookupError:
                    pass
                break
            time.sleep(0.2)
        child.wait()
    _ref_clean_shm(env["TZMPI_JOB"])
    if rank != 0:
        return None
    got = _ref_parse_e2e(open(out_path, errors="replace").read())
    if not got:
        return None
    return {"value": round(got["median_us"], 3), "unit": "us",
            "h2d_bytes_per_step": int(got.get("h2d_bytes", 0)), "d2h_bytes_per_step": int(got.get("d2h_bytes", 0)),
            "p10_us": round(got.get("p10_us", 0.0), 3), "p90_us": round(got.get("p90_us", 0.0), 3),
            "schedule_pct10_us": round(got.get("schedule_pct10_us", 0.0), 3),
            "wall_seconds": round(time.time() - t0, 1),
            "how": "baseline/ref_driver/ref_e2e.cu: the reference's own API (RowPartSpmv, SpMV, Platform, "
                   f"dfs::explore + EmpiricalBenchmarker over its first {args.ref_e2e_seqs} schedules), then per step "
                   "pinned H2D of this rank's x block, the best schedule op by op, D2H of the y block; host wall "
                   "clock, max over ranks, median"}


def _ref_clean_shm(job):
    """killed ranks can leave message segments of the MPI stand-in behind"""
    import glob
    for f in glob.glob(f"/dev/shm/tzmpi_{os.getuid()}_{job}*"):
        try:
            os.unlink(f)
        except OSError:
            pass


def reference_arm(args):
    """The reference arm: the UNMODIFIED reference (sandialabs/tenzing), built by its own CMake
    into baseline/_ref/ (baseline/build_ref.py), run through its own stock driver
    tenzing-dfs/examples/spmv.cu (`spmv`: exhaustive DFS over its SpMV op graph with two streams,
    its own EmpiricalBenchmarker, 50 measurements per schedule).  The image has no MPI, so the
    binary is linked against the single-node MPI stand-in of baseline/mpi_shim (CUDA-IPC
    rendezvous over NVLink for device buffers); nothing of tenzing_b200 is imported or loaded.

    Why this driver: the reference's MCTS drivers cannot terminate as shipped (Node::parent_ is
    never assigned, so a rollout never reaches Finish - SURVEY D1; confirmed on a B200: the
    process grows without bound), its halo driver throws at graph construction (D5), and its
    SpMV is fp32 only (static_assert, include/tenzing/spmv/ops_spmv.cuh:99-100).  The DFS driver
    hard-codes the problem (m = 150000 rows, 10 nnz/row, band half-width m/ranks) and a budget of
    15000 schedules x 50 measurements x >= 10 ms, i.e. hours.  It is run the way its authors ran
    it under a batch wall-time limit (scripts/perlmutter/spmv.sh: --signal=SIGABRT@10): after
    --ref-seconds rank 0 gets SIGINT, the reference's own handler (src/trap.cpp) prints the
    results collected so far, and the best pct10 of those schedules is reported.

    Its numbers are seconds per iteration of back-to-back runs, host wall clock, max over ranks
    (src/benchmarker.cpp:83-167) - no L2 flush, so the 18 MB fp32 matrix stays in L2, which
    favours the reference."""
    import signal
    import subprocess
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    def unavailable(why):
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": why}))
        return 0

    if not os.path.isdir("/root/reference") and not os.path.isdir(os.path.join(REF_DIR, "src")):
        return unavailable("reference sources not found (/root/reference)")
    exe, err = _ref_ensure_built(local)
    if exe is None:
        return unavailable(err)
    exe = os.path.join(REF_DIR, "build", "tenzing-dfs", "examples", "spmv")
    import tempfile
    logdir = tempfile.mkdtemp(prefix="tz_ref_")  # the reference logs hundreds of MB to stderr
    out_path = os.path.join(logdir, f"ref_n{world}_r{rank}.out")
    err_path = os.path.join(logdir, f"ref_n{world}_r{rank}.err")
    env = dict(os.environ, TENZING_ACK_NOTICE="1", TZMPI_RANK=str(rank), TZMPI_SIZE=str(world),
               TZMPI_JOB=_ref_job_id())
    # first the short end-to-end run (every rank starts it at the same time and it ends by itself),
    # then the stock driver
    try:
        e2e = _ref_e2e(args, rank, world, env["TZMPI_JOB"], logdir) if world > 1 else None
    except Exception:
        e2e = None
    t0 = time.time()
    with open(out_path, "w") as fo, open(err_path, "w") as fe:
        child = subprocess.Popen([exe], env=env, stdout=fo, stderr=fe, start_new_session=True)
        deadline = t0 + args.ref_seconds
        signalled = False
        while child.poll() is None:
            now = time.time()
            if not signalled and now >= deadline:
                if rank == 0:
                    child.send_signal(signal.SIGINT)  # the reference's handler dumps the partial CSV
                signalled = True
            if now >= deadline + args.ref_grace:
                try:
                    os.killpg(child.pid, signal.SIGKILL)  # exactly the process group started above
                except ProcessLookupError:
                    pass
                break
            time.sleep(0.2)
        child.wait()
    wall = time.time() - t0
    _ref_clean_shm(env["TZMPI_JOB"])
    if rank != 0:
        import shutil
        shutil.rmtree(logdir, ignore_errors=True)
        return 0
    rows = _ref_result_rows(open(out_path, errors="replace"))
    with open(err_path, "rb") as fe:
        fe.seek(max(0, os.path.getsize(err_path) - 400))
        tail = fe.read().decode(errors="replace").replace("\n", " ")
    import shutil
    shutil.rmtree(logdir, ignore_errors=True)
    if not rows:
        return unavailable(f"the reference's DFS driver benchmarked no schedule within {args.ref_seconds:.0f} s "
                           f"(exit {child.returncode}); stderr tail: {tail}")
    best = min(rows, key=lambda v: v[1])
    us = best[1] * 1e6
    print(json.dumps({
        **({"e2e": e2e} if e2e else {}),
        "impl": "reference", "metric": REF_METRIC, "value": round(us, 3), "unit": "us", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(us / 1e3, 6),
        "reference_measurements_per_schedule": 50,
        "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "fp32",
        "data": "synthetic",
        "config": {"model": "banded SpMV 150000 x 150000, 10 nnz/row, band half-width n/ranks, fp32 "
                            "(the reference's shipped example, tenzing-dfs/examples/spmv.cu:86-98)",
                   "global_batch": "150000 rows", "seq_len": None, "parallelism": f"row-partition {world}",
                   "driver": "baseline/_ref/build/tenzing-dfs/examples/spmv (stock, no arguments), "
                             f"SIGINT after {args.ref_seconds:.0f} s",
                   "protocol": "reference's EmpiricalBenchmarker decides the sample counts itself (--steps/--warmup "
                               "cannot be passed to its stock driver): 50 measurements of back-to-back runs >= 10 ms "
                               "each, host wall clock, max over ranks, pct10; best of the schedules reached; no L2 flush",
                   "transport": "baseline/mpi_shim (single-node MPI stand-in: CUDA-IPC rendezvous, "
                                "copy-engine pulls over NVLink on 4 streams, event completion)"},
        "schedules_benchmarked": len(rows), "pct50_us": round(best[2] * 1e6, 3),
        "wall_seconds": round(wall, 1)}))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--nx", type=int, default=512)
    ap.add_argument("--nq", type=int, default=3)
    ap.add_argument("--spmv-n", type=int, default=150_000,
                    help="headline SpMV size: the reference's shipped example (its DFS driver hard-codes it)")
    ap.add_argument("--spmv-big-n", type=int, default=1_000_000, help="BASELINE config 2 (1M x 1M band, fp64)")
    ap.add_argument("--random-rows-per-gpu", type=int, default=1_000_000,
                    help="BASELINE config 4: random-sparse SpMV, this many rows per GPU (8 M rows on 8 GPUs)")
    ap.add_argument("--search-iters", type=int, default=30)
    ap.add_argument("--no-search", action="store_true", help="use a fixed fused schedule as is")
    ap.add_argument("--search-budget", type=float, default=30.0, help="wall-clock seconds for the halo search")
    ap.add_argument("--skip", default="", help="comma separated sub-records to skip: halo,spmv_1m,spmv_random,sweep,oracle")
    ap.add_argument("--ref-seconds", type=float, default=150.0,
                    help="reference arm: wall-clock seconds its DFS driver gets before SIGINT")
    ap.add_argument("--ref-grace", type=float, default=25.0,
                    help="reference arm: seconds after the signal before the ranks are killed")
    ap.add_argument("--ref-e2e-seconds", type=float, default=90.0,
                    help="reference arm: wall-clock limit of its end-to-end run (0 = skip it)")
    ap.add_argument("--ref-e2e-seqs", type=int, default=40,
                    help="reference arm: schedules its DFS generates before the end-to-end run picks the best")
    ap.add_argument("--records-budget", type=float, default=1200.0,
                    help="seconds the sub-records after the headline (halo, spmv_1m, spmv_random) get before the line "
                         "is printed with what is complete")
    ap.add_argument("--side-budget", type=float, default=300.0,
                    help="seconds the side records (msg_sweep, oracle_pass) get before the line is printed without them")
    ap.add_argument("--host-dry-run", action="store_true",
                    help="no GPU: the same control flow with the workloads executed on the CPU "
                         "(host memory, gloo); a test of this script, its number is not a result")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    args.warmup = max(args.warmup, 3)
    skip = set(filter(None, args.skip.split(",")))

    import torch
    from tenzing_b200.parallel import init_process_group_from_env, make_comm
    from tenzing_b200.models import HaloExchange, DistributedSpMV, common
    from tenzing_b200 import _C, ops
    from tenzing_b200.utils.clocks import ClockSampler

    host = args.host_dry_run
    rank, world, local = init_process_group_from_env(backend="gloo" if host else None)
    comm = make_comm()
    dev = torch.device("cpu") if host else torch.device("cuda", torch.cuda.current_device())
    flush = torch.empty((1 << 20) if host else (256 << 20), dtype=torch.uint8, device=dev)  # > 126 MB L2

    def device_sync():
        if not host:
            torch.cuda.synchronize()

    def stream_sync():
        if not host:
            torch.cuda.current_stream().synchronize()

    def max_over_ranks(values):
        t = torch.tensor(values, dtype=torch.float64, device=dev)
        if world > 1:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        return t.cpu().tolist()

    def timed(wl, seq, steps, warmup):
        """device time of single steps: CUDA events on the launching stream, L2 flushed before
        every step, ranks lined up by a device-side barrier, max over ranks"""
        wl.platform.provision_events(seq)
        if host:  # wall clock around host-executed steps; same reductions as the device path
            t = []
            for _ in range(warmup + steps):
                comm.barrier()
                t0 = time.perf_counter()
                seq.run(wl.platform)
                wl.platform.synchronize()
                t.append(time.perf_counter() - t0)
            return max_over_ranks(t[warmup:]), 0
        _C.device_time_steps(seq, wl.platform, warmup, flush.data_ptr(), flush.numel())
        n0 = ops.launch_count()
        t = _C.device_time_steps(seq, wl.platform, steps, flush.data_ptr(), flush.numel())
        launches = ops.launch_count() - n0
        return max_over_ranks(t), launches

    def back_to_back(wl, seq, n_meas):
        """the reference's own protocol (src/benchmarker.cpp:83-167): measurements of back-to-back
        runs of at least 10 ms, no flush, host wall clock, max over ranks; returns pct10 in seconds"""
        if host:
            return None
        wl.platform.provision_events(seq)
        bo = common.bench_opts(n_iters=n_meas, target_secs=0.01, max_retries=0, check_randomness=False)
        return _C.EmpiricalBenchmarker().benchmark(seq, wl.platform, bo).pct10

    def graph_replay_of(wl, seq):
        """the schedule captured as a CUDA graph (one launch per step, the kernels read the epoch
        from a device word), or None where it cannot be (host-driven ops), fails to capture on
        some rank, or does not reproduce the oracle"""
        if host:
            return None
        ok, _why = _C.GraphReplay.supported(seq)
        replay = None
        if ok:
            try:
                wl.platform.provision_events(seq)
                replay = _C.GraphReplay(seq, wl.platform)
            except Exception as e:  # noqa: BLE001 - falls back to stream launches
                if rank == 0:
                    print(f"graph capture failed: {e}", file=sys.stderr)
                replay = None
        flag = torch.tensor([1.0 if replay is not None else 0.0], device=dev)
        if world > 1:
            torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN)
        if flag.item() < 0.5:
            return None
        wl.fill()
        comm.barrier()
        replay.launch()
        replay.launch()
        wl.sync()
        comm.barrier()
        good = torch.tensor([1.0 if wl.check() else 0.0], device=dev)
        if world > 1:
            torch.distributed.all_reduce(good, op=torch.distributed.ReduceOp.MIN)
        return replay if good.item() > 0.5 else None

    def timed_graph(wl, replay, steps, warmup):
        _C.device_time_graph_steps(replay, wl.platform, warmup, flush.data_ptr(), flush.numel())
        t = _C.device_time_graph_steps(replay, wl.platform, steps, flush.data_ptr(), flush.numel())
        return max_over_ranks(t), replay.kernel_nodes() * steps

    def measure(wl, seq, steps, warmup):
        """the schedule timed both ways - ops issued one by one into the streams, and the same
        schedule replayed as a CUDA graph; the faster one is the number"""
        t_stream, l_stream = timed(wl, seq, steps, warmup)
        rec = {"us_stream_launches": round(statistics.mean(t_stream) * 1e6, 3)}
        times, launches, mode, replay = t_stream, l_stream, "stream launches", None
        r = graph_replay_of(wl, seq)
        if r is not None:
            t_graph, l_graph = timed_graph(wl, r, steps, warmup)
            rec["us_cuda_graph"] = round(statistics.mean(t_graph) * 1e6, 3)
            if statistics.mean(t_graph) < statistics.mean(t_stream):
                times, launches, mode, replay = t_graph, l_graph, "cuda graph replay", r
        rec["launch"] = mode
        return times, launches, rec, replay

    # implementation styles whose decision space is small enough to enumerate completely
    SMALL = {"fused_all", "fused_staged", "fused_mixed", "fused_pipelined", "fused_pull",
             "fused", "fused_separate"}

    def find_best(wl, method, styles, iters, budget_s, n_streams_note=""):
        """best schedule over the given implementation styles: exhaustive DFS where the space is
        small (a few hundred sequences), MCTS (FastMin) elsewhere.  Results are max-reduced over
        ranks inside the benchmarker, so every rank picks the same winner.  The search is set-up,
        not the timed region, but it is kept inside a wall-clock budget: once rank 0 sees the
        budget spent, the remaining styles are skipped (all ranks agree through a broadcast).
        The five best candidates are then measured again with 20 measurements each, because the
        search itself looks at every schedule only briefly."""
        if args.no_search:
            return common.fixed_schedule(wl.graph(styles[0]), wl.platform)
        bo = common.bench_opts(n_iters=3, target_secs=0.002, max_retries=0, check_randomness=False)
        cands = []  # (pct10, seq)
        t0 = time.time()
        for style in styles:
            spent = torch.tensor([time.time() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                torch.distributed.broadcast(spent, src=0)
            if cands and float(spent.item()) > budget_s:
                break
            left = max(2.0, budget_s - float(spent.item()))
            if style in SMALL or method == "dfs":
                # ~6 ms per evaluated sequence: never enumerate more than the budget allows
                res = wl.search("dfs", style=style, bench=bo, max_seqs=int(min(1000, left / 0.006)))
            else:
                # cache: rollouts that collapse to the same cleaned schedule are measured once
                res = wl.search("mcts", style=style, iters=iters, bench=bo, strategy="FastMin",
                                cache=True)
            top = sorted(res.sim_results, key=lambda r: r.bench_result.pct10)[:3]
            for r in top:
                cands.append((r.bench_result.pct10, r.path if hasattr(r, "path") else r.seq))
        cands.sort(key=lambda c: c[0])
        if host or len(cands) <= 1:
            return cands[0][1]
        bo2 = common.bench_opts(n_iters=20, target_secs=0.002, max_retries=0, check_randomness=False)
        bench = _C.DeviceBenchmarker()
        best_seq, best_t = None, float("inf")
        for _, seq in cands[:5]:
            wl.platform.provision_events(seq)
            t = bench.benchmark(seq, wl.platform, bo2).pct50
            if t < best_t:
                best_seq, best_t = seq, t
        return best_seq

    def stats(us):
        q = statistics.quantiles(us, n=10) if len(us) >= 2 else [us[0]] * 9
        return {"mean": round(statistics.mean(us), 3), "median": round(statistics.median(us), 3),
                "p10": round(q[0], 3), "p90": round(q[-1], 3)}

    def run_and_check(wl, seq):
        wl.fill()
        comm.barrier()
        wl.run(seq)
        wl.sync()
        comm.barrier()
        return wl.check()

    def spmv_e2e(wl, seq, steps, warmup, replay=None):
        """through the public API, per step: H2D of this rank's x block from pinned host memory,
        the schedule, D2H of this rank's y block into pinned host memory; host wall clock"""
        rows = wl.x.numel()
        if host:
            xh = torch.rand(rows, dtype=wl.dtype)
            yh = torch.empty(rows, dtype=wl.dtype)
        else:
            # staging buffers from the framework's own pinned pool: page-locked host memory on the
            # NUMA node next to this rank's GPU (reference counterpart: src/numa.cpp, dead code there)
            from tenzing_b200.utils.tensor import host_tensor
            nbytes = max(rows, 1) * wl.x.element_size()
            pool = _C.PinnedPool(torch.cuda.current_device())
            xh = host_tensor(pool.get(nbytes), (rows,), wl.dtype, owner=pool)
            yh = host_tensor(pool.get(nbytes), (rows,), wl.dtype, owner=pool)
            xh.copy_(torch.rand(rows, dtype=wl.dtype))
        wl.platform.provision_events(seq)
        t = []
        for i in range(warmup + steps):
            if not host:
                _C.kernels.l2_flush(flush.data_ptr(), flush.numel(), i, 0)
            device_sync()
            comm.barrier()
            t0 = time.perf_counter()
            wl.x.copy_(xh, non_blocking=True)
            stream_sync()
            if replay is not None:
                replay.launch()
                wl.platform.synchronize()
            else:
                seq.run(wl.platform)
            yh.copy_(wl.y, non_blocking=True)
            stream_sync()
            t.append(time.perf_counter() - t0)
        return [v * 1e6 for v in max_over_ranks(t[warmup:])], rows * xh.element_size(), rows * yh.element_size()

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0) * 1e9
    nvl = 770e9
    t_setup = time.time()
    ok_all = True

    # ---------------- headline: distributed SpMV at the reference's shipped size ----------------
    SPMV_STYLES = ["fused_all", "fused", "fused_separate"]
    spmv = DistributedSpMV(comm=comm, n=args.spmv_n, nnz_per_row=10, host=host)
    if host:
        spmv.platform.set_host_execution(True)
    spmv.fill()
    seq_s = find_best(spmv, "dfs", SPMV_STYLES, 64, args.search_budget / 2)
    ok_all &= run_and_check(spmv, seq_s)
    # SM clock / throttle reasons: sampled from before the timed region until after the
    # end-to-end loop (NVML thread + nvidia-smi -lms 20; start() returns once a sample exists)
    sampler = ClockSampler(0 if host else torch.cuda.current_device())
    if rank == 0:
        sampler.start()
    comm.barrier()
    times_s, launches_s, rec_s, replay_s = measure(spmv, seq_s, args.steps, args.warmup)
    e2e_s, h2d_s, d2h_s = spmv_e2e(spmv, seq_s, args.steps, args.warmup, replay_s)
    b2b_s = back_to_back(spmv, seq_s, 20)
    clocks = sampler.stop() if rank == 0 else {}
    us_s = statistics.mean(times_s) * 1e6
    roof_s = max((spmv.recv_bytes() if world > 1 else 0) / nvl, spmv.hbm_bytes() / hbm)
    # the same in fp32, the reference's only type: the like-for-like row
    spmv32 = DistributedSpMV(comm=comm, n=args.spmv_n, nnz_per_row=10, dtype=torch.float32, host=host)
    if host:
        spmv32.platform.set_host_execution(True)
    spmv32.fill()
    seq_s32 = find_best(spmv32, "dfs", SPMV_STYLES, 64, args.search_budget / 4)
    ok_all &= run_and_check(spmv32, seq_s32)
    times_s32, _, rec_s32, _ = measure(spmv32, seq_s32, args.steps, args.warmup)
    b2b_s32 = back_to_back(spmv32, seq_s32, 20)

    out = {
        "metric": REF_METRIC, "value": round(us_s, 3), "unit": "us", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(us_s / 1e3, 6), "higher_is_better": False,
        "scaling": "strong", "vs_baseline": None, "dtype": "fp64", "data": "synthetic",
        "config": {"model": f"banded SpMV {args.spmv_n} x {args.spmv_n}, ~10 nnz/row, band half-width n/ranks "
                            "(the reference's shipped example, tenzing-dfs/examples/spmv.cu:86-98; BASELINE "
                            "config 2 at the size its stock driver hard-codes), fp64 where the reference is fp32",
                   "global_batch": f"{args.spmv_n} rows", "seq_len": None, "parallelism": f"row-partition {world}",
                   "l2": "256 MB flush buffer written and read back before every timed step",
                   "timing": "CUDA events, one step per measurement, ranks lined up by a device-side barrier, "
                             "max over ranks, mean of the steps",
                   "launch": rec_s["launch"] + " (the schedule is timed both ways, the faster is the value)",
                   "search": "no search" if args.no_search else
                   "exhaustive DFS inside each fused style, the 5 best re-measured with 20 measurements each"},
        "schedule": seq_s.names(),
        "spmv": {"us": round(us_s, 3), **rec_s, **stats([t * 1e6 for t in times_s]), "n": args.spmv_n,
                 "roofline_us": round(roof_s * 1e6, 3), "roofline_frac_of_measured": round(roof_s * 1e6 / us_s, 4),
                 "reference_protocol_pct10_us": None if b2b_s is None else round(b2b_s * 1e6, 3),
                 "fp32": {"us": round(statistics.mean(times_s32) * 1e6, 3), **rec_s32, "schedule": seq_s32.names(),
                          "reference_protocol_pct10_us": None if b2b_s32 is None else round(b2b_s32 * 1e6, 3)}},
        "e2e": {"value": round(statistics.median(e2e_s), 3), "unit": "us", **stats(e2e_s),
                "h2d_bytes_per_step": h2d_s, "d2h_bytes_per_step": d2h_s},
        "gpu_launches": int(launches_s), "clocks": clocks,
        "impl": "ours-host-dry-run (not a result)" if host else "ours",
    }
    del spmv32, replay_s

    # From here on the headline record is complete.  Everything below adds sub-records and must
    # not be able to take it down: if a section fails (say a device wait that trapped, after
    # which the CUDA context is gone), or the rest has not finished after --records-budget seconds
    # (a rank stuck in a collective behind a failed peer), rank 0 prints the line with what is
    # complete, says so in it, and every rank leaves with exit code 0.
    import threading
    done_lock = threading.Lock()
    state = {"printed": False, "ok_all": ok_all}

    def emit(note=None):
        with done_lock:
            if state["printed"]:
                return False
            state["printed"] = True
        if note:
            out["records_incomplete"] = note
            out["correct"] = bool(state["ok_all"])
        out["setup_and_search_seconds"] = round(time.time() - t_setup, 1)
        if rank == 0:
            print(json.dumps(out), flush=True)
        return True

    def failsafe(what, seconds):
        if emit(f"{what} did not finish within {seconds:.0f} s; the records present are complete"):
            os._exit(0)

    timer = threading.Timer(args.records_budget, failsafe, ("the sub-records", args.records_budget))
    timer.daemon = True
    timer.start()
    try:
        # ---------------- BASELINE config 3: 3-D halo exchange ----------------
        if "halo" not in skip:
            halo = HaloExchange(comm=comm, nX=args.nx, nQ=args.nq, nGhost=3, storage_order="XYZQ",
                                neighbors=26, host=host)
            if host:
                halo.platform.set_host_execution(True)
            halo.fill()
            # the search: exhaustive inside every small fused implementation style, MCTS for
            # fused_groups and over the ChoiceOp of ALL styles, reference structure included;
            # most promising styles first, so a spent budget drops the least promising ones
            fused = ["fused_mixed", "fused_staged", "fused_all", "fused_pipelined", "fused_groups", "fused_pull"]
            seq_h = find_best(halo, "mcts", fused + ["choice"], args.search_iters, args.search_budget)
            ok_h = run_and_check(halo, seq_h)
            ok_all &= ok_h
            times_h, launches_h, rec_h, replay_h = measure(halo, seq_h, args.steps, args.warmup)
            # end to end: H2D of one fresh interior z-plane (all quantities), the exchange, D2H of a ghost row
            a = halo.args
            g = a.nGhost
            plane = torch.randn(a.nQ, a.nY, a.nX, dtype=torch.float64)
            row_out = torch.empty(a.nX, dtype=torch.float64)
            if not host:
                plane, row_out = plane.pin_memory(), row_out.pin_memory()
            cells = halo.cells()
            halo.platform.provision_events(seq_h)
            e2e = []
            for i in range(args.warmup + args.steps):
                if not host:
                    _C.kernels.l2_flush(flush.data_ptr(), flush.numel(), i, 0)
                device_sync()
                comm.barrier()
                t0 = time.perf_counter()
                cells[:, g, g:g + a.nY, g:g + a.nX].copy_(plane, non_blocking=True)
                stream_sync()
                if replay_h is not None:
                    replay_h.launch()
                    halo.platform.synchronize()
                else:
                    seq_h.run(halo.platform)
                row_out.copy_(cells[0, 0, g, g:g + a.nX], non_blocking=True)
                stream_sync()
                e2e.append(time.perf_counter() - t0)
            e2e_h = [v * 1e6 for v in max_over_ranks(e2e[args.warmup:])]
            us_h = statistics.mean(times_h) * 1e6
            bytes_out = halo.bytes_per_exchange()
            remote = halo.remote_bytes_per_exchange()
            # roofline: the slower of (bytes over NVLink / link bw) and (HBM bytes / HBM bw); every halo
            # byte is read once and written once somewhere (locally or on the peer)
            roof_h = max(remote / nvl, (2 * bytes_out - remote) / hbm)
            out["halo"] = {"us": round(us_h, 3), **rec_h, **stats([t * 1e6 for t in times_h]),
                           "config": f"halo {args.nx}^3 fp64/rank nQ={args.nq} radius 3, 26 neighbours, XYZQ, "
                                     "rank grid " + "x".join(map(str, halo.domain.rank_grid())),
                           "schedule": seq_h.names(), "correct": bool(ok_h),
                           "bytes_out_per_rank": bytes_out, "remote_bytes_per_rank": remote,
                           "roofline_us": round(roof_h * 1e6, 3),
                           "roofline_frac_of_measured": round(roof_h * 1e6 / us_h, 4),
                           "gbs_per_rank": round(bytes_out / (us_h * 1e-6) / 1e9, 1),
                           "gpu_launches": int(launches_h),
                           "e2e_us": stats(e2e_h), "e2e_h2d_bytes": plane.numel() * 8, "e2e_d2h_bytes": row_out.numel() * 8}
            del replay_h, halo, cells

        # ---------------- BASELINE config 2: band SpMV 1M x 1M, fp64, DFS ----------------
        if "spmv_1m" not in skip:
            big = DistributedSpMV(comm=comm, n=args.spmv_big_n, nnz_per_row=10, host=host)
            if host:
                big.platform.set_host_execution(True)
            big.fill()
            seq_b = find_best(big, "dfs", SPMV_STYLES, 64, args.search_budget / 2)
            ok_b = run_and_check(big, seq_b)
            ok_all &= ok_b
            times_b, launches_b, rec_b, replay_b = measure(big, seq_b, args.steps, args.warmup)
            us_b = statistics.mean(times_b) * 1e6
            roof_b = max((big.recv_bytes() if world > 1 else 0) / nvl, big.hbm_bytes() / hbm)
            out["spmv_1m"] = {"us": round(us_b, 3), **rec_b, **stats([t * 1e6 for t in times_b]), "n": args.spmv_big_n,
                              "schedule": seq_b.names(), "correct": bool(ok_b), "roofline_us": round(roof_b * 1e6, 3),
                              "roofline_frac_of_measured": round(roof_b * 1e6 / us_b, 4),
                              "gpu_launches": int(launches_b)}
            del replay_b, big

        # ---------------- BASELINE config 4: random-sparse SpMV, MCTS, 3 streams ----------------
        if "spmv_random" not in skip and not host:
            n_r = args.random_rows_per_gpu * world
            rnd = DistributedSpMV(comm=comm, n=n_r, bw=n_r, nnz_per_row=10, n_streams=3)
            rnd.fill()
            if args.no_search:
                seq_r = common.fixed_schedule(rnd.graph("fused_all"), rnd.platform)
            else:
                bo = common.bench_opts(n_iters=3, target_secs=0.002, max_retries=0, check_randomness=False)
                res = rnd.search("mcts", style="choice", iters=args.search_iters, bench=bo, strategy="FastMin", cache=True)
                seq_r, _ = common.best_of(res)
            ok_r = run_and_check(rnd, seq_r)
            ok_all &= ok_r
            times_r, launches_r, rec_r, replay_r = measure(rnd, seq_r, args.steps, args.warmup)
            us_r = statistics.mean(times_r) * 1e6
            roof_r = max((rnd.recv_bytes() if world > 1 else 0) / nvl, rnd.hbm_bytes() / hbm)
            out["spmv_random"] = {"us": round(us_r, 3), **rec_r, **stats([t * 1e6 for t in times_r]), "n": n_r, "streams": 3,
                                  "search": f"MCTS FastMin {args.search_iters} iterations over the ChoiceOp of all styles",
                                  "schedule": seq_r.names(), "correct": bool(ok_r),
                                  "roofline_us": round(roof_r * 1e6, 3),
                                  "roofline_frac_of_measured": round(roof_r * 1e6 / us_r, 4),
                                  "gpu_launches": int(launches_r)}
            del replay_r, rnd

        state["ok_all"] = ok_all
    except Exception as e:  # noqa: BLE001
        timer.cancel()
        if emit(f"a sub-record failed: {str(e)[:200]}"):
            os._exit(0)
    timer.cancel()

    # The two side records get their own, shorter budget.
    timer = threading.Timer(args.side_budget, failsafe, ("msg_sweep / oracle_pass", args.side_budget))
    timer.daemon = True
    timer.start()
    try:
        ok_all = _side_records(args, out, skip, host, world, comm, common, flush, max_over_ranks, dev, ok_all)
        state["ok_all"] = ok_all
        ok = torch.tensor([1.0 if ok_all else 0.0], device=dev)
        if world > 1:
            torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN)
        out["correct"] = bool(ok.item() > 0.5)
    except Exception as e:  # noqa: BLE001 - e.g. a device wait that trapped: the context is gone, the line is not
        timer.cancel()
        if emit(f"msg_sweep / oracle_pass failed: {str(e)[:200]}"):
            os._exit(0)
    timer.cancel()
    emit()
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
    return 0


def _side_records(args, out, skip, host, world, comm, common, flush, max_over_ranks, dev, ok_all):
    """BASELINE config 5 (message sizes) and the per-style oracle pass list; returns ok_all"""
    import torch
    from tenzing_b200.models import HaloExchange, DistributedSpMV
    # ---------------- BASELINE config 5: message sizes, every rank to every other rank ----------------
    if "sweep" not in skip and not host and world > 1:
        from tenzing_b200.utils.sweep import all_pairs_sweep
        try:
            plat_sw = common.make_platform(comm, 1)
            rows = all_pairs_sweep(comm, plat_sw, [64 << 10, 16 << 20, 256 << 20], max(5, args.steps // 4), flush,
                                   max_over_ranks)
            out["msg_sweep"] = {"pattern": f"every rank sends to each of its {world - 1} peers at once and "
                                           "waits for theirs", "rows": rows}
            ok_all &= all(r["correct"] for r in rows)
            del plat_sw
        except Exception as e:  # noqa: BLE001 - a side record must not take the headline down
            out["msg_sweep"] = {"error": str(e)[:200]}

    # ---------------- every implementation style against the oracles, on these ranks ----------------
    if "oracle" not in skip and not args.no_search:
        from tenzing_b200.models.halo import STYLES as HALO_ALL
        from tenzing_b200.models.spmv import STYLES as SPMV_ALL, EXPERIMENTAL_STYLES
        passed = {"halo": [], "spmv": []}
        small_h = HaloExchange(comm=comm, nX=24, nY=12, nZ=16, nQ=2, nGhost=3, host=host)
        small_s = DistributedSpMV(comm=comm, n=40_000, nnz_per_row=10, host=host)
        for wl, styles, key in ((small_h, HALO_ALL, "halo"), (small_s, SPMV_ALL + EXPERIMENTAL_STYLES, "spmv")):
            if host:
                wl.platform.set_host_execution(True)
            for style in styles:
                seq = common.fixed_schedule(wl.graph(style), wl.platform)
                wl.platform.provision_events(seq)
                wl.fill()
                comm.barrier()
                wl.run(seq, 2)          # both staging parities
                wl.sync()
                comm.barrier()
                good = torch.tensor([1.0 if wl.check() else 0.0], device=dev)
                if world > 1:
                    torch.distributed.all_reduce(good, op=torch.distributed.ReduceOp.MIN)
                if good.item() > 0.5:
                    passed[key].append(style)
                else:
                    ok_all = False
        out["oracle_pass"] = {**passed, "of": {"halo": len(HALO_ALL), "spmv": len(SPMV_ALL + EXPERIMENTAL_STYLES)}}
        del small_h, small_s
    return ok_all


if __name__ == "__main__":
    sys.exit(main())
