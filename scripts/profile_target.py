#!/usr/bin/env python
"""Small fixed launch list for ncu: each interesting halo kernel twice (warm-up + measured)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")
import torch  # noqa: E402

from tenzing_b200 import _C  # noqa: E402
from tenzing_b200.models import HaloExchange, DistributedSpMV, common  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "halo"
if which == "halo":
    h = HaloExchange(nX=512, nQ=3, nGhost=3, n_streams=1)
    h.fill()
    names = {"direct": ["he_put_dx0_dy0_dz1", "he_put_dx1_dy0_dz0"],
             "staged": ["he_packput_dx1_dy0_dz0", "he_packput_dx0_dy0_dz1"],
             "fused_all": ["he_put_all"], "fused_staged": ["he_packput_all"]}
    for rep in range(2):
        for style, ops in names.items():
            g = h.graph("flat:" + style)
            for n in ops:
                op = g.find_name(n)
                # a direct put waits for the neighbour's permission, which the style's he_ready op grants
                pre = [_C.BoundGpuOp(g.find_name("he_ready"), 0)] if style == "direct" else []
                seq = _C.Sequence(pre + [_C.BoundGpuOp(op, 0)])
                seq.run(h.platform)
                torch.cuda.synchronize()
        # a full fused_staged exchange (put + wait/unpack)
        seq = common.fixed_schedule(h.graph("fused_staged"), h.platform)
        h.run(seq)
        torch.cuda.synchronize()
else:
    s = DistributedSpMV(n=1_000_000, nnz_per_row=10)
    s.fill()
    for style in ("fused", "fused_all"):
        seq = common.fixed_schedule(s.graph(style), s.platform)
        for rep in range(2):
            s.run(seq)
            torch.cuda.synchronize()
print("done")
