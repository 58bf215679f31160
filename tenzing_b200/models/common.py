"""Shared search / replay helpers for the workload front ends."""
import os

from .. import _C


def make_platform(comm, n_streams=2, null=False):
    plat = _C.Platform.make_n_streams(n_streams, comm, null)
    # opt-in: line up measurement starts with a GPU-side all-rank barrier (see DeviceBarrier)
    if not null and comm.size() > 1 and os.environ.get("TENZING_DEVICE_BARRIER") == "1":
        plat.enable_device_barrier()
    return plat


def bench_opts(n_iters=50, target_secs=0.01, max_retries=2, check_randomness=True):
    return _C.BenchOpts(n_iters=n_iters, max_retries=max_retries, target_secs=target_secs,
                        check_randomness=check_randomness)


def with_cache(benchmarker, graph, resume_csv=None, cache=False):
    """Wrap `benchmarker` in a CachingBenchmarker when asked to.

    resume_csv: path (or text) of the results CSV of an earlier, interrupted run of the same
    search: with the same seed (and the same `cache` setting) the solver retraces that run from
    the file and continues where it stopped.  cache=True: never measure a schedule twice.
    Every rank must be given the same file."""
    if not resume_csv and not cache:
        return benchmarker
    cb = _C.CachingBenchmarker(benchmarker, reuse=bool(cache))
    if resume_csv:
        if os.path.exists(resume_csv):
            prior = _C.CsvBenchmarker(resume_csv, graph)
        else:
            prior = _C.CsvBenchmarker.from_text(resume_csv, graph)
        cb.preload_csv(prior)
    return cb


def search_mcts(graph, platform, benchmarker=None, iters=50, strategy="FastMin", bench=None,
                seed=0, expand_rollout=True, dump_tree_prefix=None, resume_csv=None, cache=False):
    """MCTS over `graph`; returns _C.mcts.Result (rank 0 holds the tree; every rank gets results)."""
    opts = _C.mcts.Opts()
    opts.n_iters = iters
    opts.seed = seed
    opts.expand_rollout = expand_rollout
    if dump_tree_prefix:
        opts.dump_tree = True
        opts.dump_tree_prefix = dump_tree_prefix
    opts.bench_opts = bench or bench_opts()
    benchmarker = with_cache(benchmarker or _C.DeviceBenchmarker(), graph, resume_csv, cache)
    return _C.mcts.explore(graph, platform, benchmarker, strategy, opts)


def search_dfs(graph, platform, benchmarker=None, max_seqs=-1, bench=None,
               remove_redundant_syncs=True, resume_csv=None, cache=False):
    """Exhaustive search; returns _C.dfs.Result."""
    opts = _C.dfs.Opts()
    opts.max_seqs = max_seqs
    opts.remove_redundant_syncs = remove_redundant_syncs
    opts.bench_opts = bench or bench_opts()
    benchmarker = with_cache(benchmarker or _C.DeviceBenchmarker(), graph, resume_csv, cache)
    return _C.dfs.explore(graph, platform, benchmarker, opts)


def best_of(result):
    """(sequence, BenchResult) of the fastest schedule by pct10, or (None, None)."""
    i = result.best()
    if i < 0:
        return None, None
    sr = result.sim_results[i]
    seq = sr.path if hasattr(sr, "path") else sr.seq
    return seq, sr.bench_result


def fixed_schedule(graph, platform, prefer_stream=None):
    """A deterministic complete sequence: always take the first decision offered.

    Useful as a quick valid schedule (tests, smoke) without running a search."""
    state = _C.State(graph)
    while not state.is_terminal():
        decisions = state.get_decisions(platform)
        if not decisions:
            raise RuntimeError("dead end while building a fixed schedule")
        state = state.apply(decisions[0])
    seq = state.sequence()
    _C.remove_redundant_syncs(seq)
    return seq
