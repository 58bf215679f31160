"""Three kernels k1 -> {k2, k3} on two streams: what the solver enumerates, and how a timeline
simulation ranks the schedules (no GPU needed)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tenzing_b200 as tz  # noqa: E402

k1, k2, k3 = (tz.EmptyGpuOp(n) for n in ("k1", "k2", "k3"))
g = tz.Graph()
g.start_then(k1)
g.then(k1, k2)
g.then(k1, k3)
g.then_finish(k2)
g.then_finish(k3)
plat = tz.Platform.make_null(2)
sim = tz.SimBenchmarker({"k1": 20e-6, "k2": 100e-6, "k3": 100e-6}, launch_cost=3e-6)
res = tz.dfs.explore(g, plat, sim)
rows = sorted(res.sim_results, key=lambda r: r.bench_result.pct10)
print(f"{len(rows)} distinct schedules")
print("fastest:", round(rows[0].bench_result.pct10 * 1e6, 1), "us:", rows[0].seq.desc())
print("slowest:", round(rows[-1].bench_result.pct10 * 1e6, 1), "us:", rows[-1].seq.desc())
