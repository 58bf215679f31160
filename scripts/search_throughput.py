#!/usr/bin/env python
"""Search-engine throughput on the CPU, this framework against the unmodified reference
(BASELINE.json config 1: noop-graph DFS schedule enumeration, no GPU).

Both arms enumerate every schedule of the same graphs through their public DFS entry point
(`dfs::get_all_sequences`): K independent chains of L NoOps between Start and Finish, which have
(K*L)! / (L!)^K schedules.  The reference side is baseline/ref_driver/ref_enum.cu built against
the reference by baseline/build_ref.py; both run on this machine, one thread.

A second section times MCTS playouts on the workload graphs (this framework only).

  python scripts/search_throughput.py [--out profiles/search_throughput.txt]
"""
import argparse
import math
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("TENZING_ACK_NOTICE", "1")

CASES = [(3, 1), (3, 2), (4, 2), (3, 3), (5, 2), (4, 3)]


def ours(K, L, repeats):
    import tenzing_b200 as tz
    g = tz.Graph()
    for k in range(K):
        prev = None
        for l in range(L):
            op = tz.NoOp(f"op{k}_{l}")
            if prev is None:
                g.start_then(op)
            else:
                g.then(prev, op)
            prev = op
        g.then_finish(prev)
    plat = tz.Platform.make_null(1)
    best, n = float("inf"), 0
    for _ in range(repeats):
        t0 = time.perf_counter()
        seqs = tz.dfs.get_all_sequences(g, plat)
        best = min(best, time.perf_counter() - t0)
        n = len(seqs)
    return n, best


def reference(K, L, repeats, limit_s):
    exe = os.path.join(ROOT, "baseline", "_ref", "build_e2e", "ref-enum")
    if not os.path.exists(exe):
        return None, None
    env = dict(os.environ, TZMPI_RANK="0", TZMPI_SIZE="1", TZMPI_JOB=f"enum{os.getpid()}")
    best, n = float("inf"), None
    for _ in range(repeats):
        try:
            r = subprocess.run([exe, str(K), str(L)], env=env, capture_output=True, text=True, timeout=limit_s)
        except subprocess.TimeoutExpired:
            return None, f">{limit_s:.0f}"
        for line in r.stdout.splitlines():
            if line.startswith("REF_ENUM "):
                kv = dict(f.split("=") for f in line.split()[1:])
                n, best = int(kv["sequences"]), min(best, float(kv["seconds"]))
    return n, (best if n is not None else None)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--repeats", type=int, default=3)
    ap.add_argument("--ref-limit", type=float, default=600.0, help="seconds per reference run")
    ap.add_argument("--reuse-reference", default="",
                    help="take the reference's times from this earlier output file instead of running it again "
                         "(its side does not change when this framework does)")
    args = ap.parse_args()
    reused = {}
    if args.reuse_reference:
        for line in open(os.path.join(ROOT, args.reuse_reference)):
            f = line.split()
            if len(f) >= 8 and f[0].isdigit():
                reused[(int(f[0]), int(f[1]))] = (int(f[5]), float(f[6]))
    lines = ["# noop-graph DFS enumeration on the CPU (one thread, best of %d): K chains of L NoOps" % args.repeats,
             "# chains len schedules_expected ours_schedules ours_s reference_schedules reference_s reference/ours"]
    for K, L in CASES:
        want = math.factorial(K * L) // math.factorial(L) ** K
        n_o, t_o = ours(K, L, args.repeats)
        n_r, t_r = reused.get((K, L)) or reference(K, L, 1 if want > 20000 else args.repeats, args.ref_limit)
        ratio = f"{t_r / t_o:.1f}" if isinstance(t_r, float) and t_o > 0 else "-"
        t_r_s = f"{t_r:.6f}" if isinstance(t_r, float) else str(t_r)
        lines.append(f"{K} {L} {want} {n_o} {t_o:.6f} {n_r} {t_r_s} {ratio}")
        print(lines[-1], flush=True)
        assert n_o == want, "this framework's enumeration is wrong"
    # MCTS playouts on the workload graphs (this framework only: the reference's MCTS cannot finish a
    # rollout as shipped).  A constant-cost benchmarker, so the time is the search's own.
    import tenzing_b200 as tz
    from tenzing_b200 import _C
    from tenzing_b200.models import DistributedSpMV, HaloExchange
    # exhaustive enumeration of the workload graphs with GPU ops (two streams; this framework only -
    # the reference needs a GPU to build these graphs; its stock SpMV driver spends about 40 s on the
    # same 2036 schedules before its first benchmark)
    lines.append("# DFS enumeration of workload graphs, 2 streams: workload style distinct_schedules seconds")
    spmv_e = DistributedSpMV(n=4000, nnz_per_row=6, host=True)
    for style in ("reference", "fused", "choice"):
        g = spmv_e.graph(style)
        best, n = float("inf"), 0
        for _ in range(args.repeats):
            t0 = time.perf_counter()
            n = len(tz.dfs.get_all_sequences(g, spmv_e.platform, 200000))
            best = min(best, time.perf_counter() - t0)
        lines.append(f"spmv {style} {n} {best:.4f}")
        print(lines[-1], flush=True)
    lines.append("# MCTS (FastMin, rollouts expanded into the tree), constant-cost benchmarker: workload style ops_per_schedule rollouts ms_per_rollout")
    halo = HaloExchange(nX=16, nQ=1, nGhost=1, n_streams=2, host=True)
    spmv = DistributedSpMV(n=4000, nnz_per_row=6, host=True)
    for name, wl, style in (("halo", halo, "reference"), ("halo", halo, "direct"), ("halo", halo, "choice"),
                            ("spmv", spmv, "reference"), ("spmv", spmv, "choice")):
        g = wl.graph(style)
        opts = _C.mcts.Opts()
        opts.n_iters = 200
        t0 = time.perf_counter()
        res = _C.mcts.explore(g, wl.platform, tz.PyBenchmarker(lambda s: 1.0), "FastMin", opts)
        dt = time.perf_counter() - t0
        n = len(res.sim_results)
        ops = max(len(r.path) for r in res.sim_results)
        lines.append(f"{name} {style} {ops} {n} {dt / n * 1e3:.3f}")
        print(lines[-1], flush=True)
    if args.out:
        with open(os.path.join(ROOT, args.out), "w") as f:
            f.write("\n".join(lines) + "\n")


if __name__ == "__main__":
    main()
