"""Search the halo-exchange schedule on the GPUs this process group has (run under
torch.distributed.run for more than one), print the best schedule and the rules the
post-processor extracts from the search log."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from tenzing_b200.models import HaloExchange, common
from tenzing_b200.parallel import init_process_group_from_env, make_comm
from tenzing_b200 import postprocess

rank, world, _ = init_process_group_from_env()
halo = HaloExchange(comm=make_comm(), nX=128, nQ=3, nGhost=3)
halo.fill()
res = halo.search("mcts", style="choice", iters=100, strategy="FastMin",
                  bench=common.bench_opts(n_iters=5, target_secs=0.002))
seq, stats = common.best_of(res)
halo.fill()
halo.run(seq)
assert halo.check()
if rank == 0:
    print("best:", round(stats.pct10 * 1e6, 1), "us:", seq.desc())
    print(postprocess.analyse(res.csv(), prominence_pct=90.0)["text"])
if world > 1:
    torch.distributed.destroy_process_group()
