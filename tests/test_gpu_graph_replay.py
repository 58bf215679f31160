"""Sequences replayed as CUDA graphs (tenzing/cuda/graph_replay.hpp): same results as issuing the
ops one by one, epoch after epoch (the flag protocol needs a fresh epoch per replay: the kernels
read it from a device word the graph's first node increments), also when replays and ordinary
runs alternate."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("style", ["fused_all", "fused_staged", "fused_mixed", "fused_pipelined", "staged", "direct",
                                   "fused_groups", "fused_pull"])
def test_halo_exchange_replayed_as_a_graph(style):
    from tenzing_b200 import _C
    from tenzing_b200.models import HaloExchange, common
    h = HaloExchange(nX=40, nY=24, nZ=16, nQ=2, nGhost=3)
    seq = common.fixed_schedule(h.graph(style), h.platform)
    ok, why = _C.GraphReplay.supported(seq)
    assert ok, why
    h.platform.provision_events(seq)
    h.fill()
    h.run(seq)                      # warm-up the ordinary way (one-time set-up inside the launch paths)
    h.sync()
    replay = _C.GraphReplay(seq, h.platform)
    assert replay.kernel_nodes() >= 1
    for round_ in range(2):
        h.fill()
        for _ in range(3):          # both epoch parities
            replay.launch()
        h.sync()
        assert h.check(), (style, round_)
        h.fill()
        h.run(seq)                  # an ordinary run in between: the device epoch must follow
        h.sync()
        assert h.check()


@pytest.mark.parametrize("style", ["fused", "fused_all", "fused_separate", "fused_pull"])
@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
def test_spmv_replayed_as_a_graph(style, dtype):
    from tenzing_b200 import _C
    from tenzing_b200.models import DistributedSpMV, common
    s = DistributedSpMV(n=60_000, nnz_per_row=10, dtype=dtype)
    seq = common.fixed_schedule(s.graph(style), s.platform)
    ok, why = _C.GraphReplay.supported(seq)
    assert ok, why
    s.platform.provision_events(seq)
    s.fill()
    s.run(seq)
    s.sync()
    replay = _C.GraphReplay(seq, s.platform)
    for _ in range(3):
        s.fill()
        replay.launch()
        replay.launch()
        s.sync()
        assert s.check()


def test_host_driven_sequences_are_refused():
    from tenzing_b200 import _C
    from tenzing_b200.models import HaloExchange, common
    h = HaloExchange(nX=16, nQ=1, nGhost=2)
    seq = common.fixed_schedule(h.graph("reference"), h.platform)
    ok, why = _C.GraphReplay.supported(seq)
    assert not ok and "host" in why
    with pytest.raises(RuntimeError):
        _C.GraphReplay(seq, h.platform)
