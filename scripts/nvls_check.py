#!/usr/bin/env python
"""Is NVLS multicast available here, and what does the `nvls` SpMV style cost next to the fused
styles?  Random-sparse matrix (every rank needs most of x), run under torch.distributed.run."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")
import torch  # noqa: E402

from tenzing_b200 import _C  # noqa: E402
from tenzing_b200.models import DistributedSpMV, common  # noqa: E402
from tenzing_b200.ops import nvls  # noqa: E402
from tenzing_b200.parallel import init_process_group_from_env, make_comm  # noqa: E402

rank, world, _ = init_process_group_from_env()
comm = make_comm()
ok = nvls.available()
if rank == 0:
    print("NVLS multicast available:", ok, flush=True)
n = 1_000_000 * world
s = DistributedSpMV(comm=comm, n=n, bw=n, nnz_per_row=10)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for style in ("fused", "fused_all") + (("nvls",) if ok else ()):
    seq = common.fixed_schedule(s.graph(style), s.platform)
    s.fill()
    comm.barrier()
    s.run(seq, 2)
    s.sync()
    comm.barrier()
    good = s.check()
    s.platform.provision_events(seq)
    _C.device_time_steps(seq, s.platform, 3, flush.data_ptr(), flush.numel())
    t = torch.tensor(_C.device_time_steps(seq, s.platform, 9, flush.data_ptr(), flush.numel()), device="cuda")
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    if rank == 0:
        print(f"random SpMV {n} rows on {world} GPUs, style {style:10s} {float(t.median()) * 1e6:9.2f} us  ok={good}", flush=True)
torch.distributed.barrier()
torch.distributed.destroy_process_group()
