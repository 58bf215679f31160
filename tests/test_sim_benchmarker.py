"""SimBenchmarker: timeline simulation of schedules (streams, events, host syncs)."""
import pytest

import tenzing_b200 as tz
from tenzing_b200 import _C
from tenzing_b200.models import HaloExchange

US = 1e-6


def test_stream_overlap_and_sync_semantics():
    a, b, c = tz.EmptyGpuOp("a"), tz.EmptyGpuOp("b"), tz.EmptyGpuOp("c")
    sim = tz.SimBenchmarker({"a": 100 * US, "b": 100 * US, "c": 50 * US}, launch_cost=1 * US)
    B = tz.BoundGpuOp
    # same stream: serial
    one = tz.Sequence([B(a, 0), B(b, 0), tz.CudaEventRecord(0, 0), tz.CudaEventSync(0)])
    assert sim.simulate(one) == pytest.approx(201 * US, rel=1e-6)
    # two streams: a and b overlap; host waits for both
    two = tz.Sequence([B(a, 0), B(b, 1), tz.CudaEventRecord(0, 0), tz.CudaEventRecord(1, 1),
                       tz.CudaEventSync(0), tz.CudaEventSync(1)])
    assert sim.simulate(two) == pytest.approx(102 * US, rel=1e-6)
    # c on stream 1 must wait for a (stream 0) through the event
    dep = tz.Sequence([B(a, 0), tz.CudaEventRecord(0, 0), tz.CudaStreamWaitEvent(1, 0), B(c, 1),
                       tz.CudaEventRecord(1, 1), tz.CudaEventSync(1)])
    assert sim.simulate(dep) == pytest.approx(151 * US, rel=1e-6)
    # without the stream wait the simulator lets c start early (it models what was written)
    nodep = tz.Sequence([B(a, 0), B(c, 1), tz.CudaEventRecord(1, 1), tz.CudaEventSync(1)])
    assert sim.simulate(nodep) < 60 * US
    # contention stretches overlapping ops
    slow = tz.SimBenchmarker({"a": 100 * US, "b": 100 * US}, launch_cost=1 * US, contention=1.0)
    assert slow.simulate(two) > 190 * US


def test_search_on_the_simulator_finds_the_pipelined_mixed_halo_schedule():
    """with per-op times like the ones measured at 8 GPUs (profiles/), the best fused_mixed
    schedule is the one the real search found: thin pack+put first, then the thick put with the
    thin unpack running beside it on the other stream"""
    h = HaloExchange(nX=8, nQ=1, nGhost=1, host=True)
    g = h.graph("fused_mixed")
    plat = h.platform
    # the unpack can only start once the peers' thin puts have landed: model its flag wait by
    # giving it the thin put's duration as a floor through the stream/event dependency
    sim = tz.SimBenchmarker({"he_packput_thin": 80 * US, "he_put_thick": 125 * US,
                             "he_waitunpack_thin": 70 * US, "he_wait_thick": 3 * US}, launch_cost=3 * US)
    res = tz.dfs.explore(g, plat, sim)
    assert len(res.sim_results) > 100
    best = res.sim_results[res.best()]
    ops = [(o.name(), o.stream()) for o in best.seq if o.kind() == "BoundGpuOp"]
    names = [n for n, _ in ops]
    stream = dict(ops)
    assert names.index("he_packput_thin") < names.index("he_put_thick")
    assert stream["he_waitunpack_thin"] != stream["he_put_thick"]
    # serial lower bound is 80 + 125 + 70; overlap brings it close to 80 + 125
    assert best.bench_result.pct10 < 80 * US + 125 * US + 40 * US
    worst = max(r.bench_result.pct10 for r in res.sim_results)
    assert worst > 80 * US + 125 * US + 70 * US
    # MCTS reaches (nearly) the same optimum in a fraction of the evaluations
    opts = _C.mcts.Opts()
    opts.n_iters = 120
    opts.seed = 2
    m = _C.mcts.explore(g, plat, sim, "FastMin", opts)
    assert min(r.bench_result.pct10 for r in m.sim_results) <= 1.1 * best.bench_result.pct10
