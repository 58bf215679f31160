#!/usr/bin/env python
"""Per-op device times of the halo-exchange kernels on one GPU (periodic self-exchange).

For each implementation style every GPU op of its graph is run alone on one stream and timed
with CUDA events (L2 flushed between runs); prints us, payload bytes and GB/s (read+write).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")
import torch  # noqa: E402

from tenzing_b200 import _C  # noqa: E402
from tenzing_b200.models import HaloExchange  # noqa: E402
from tenzing_b200.parallel import init_process_group_from_env, make_comm  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=512)
    ap.add_argument("--nq", type=int, default=3)
    ap.add_argument("--order", default="XYZQ")
    ap.add_argument("--styles", default="direct,staged,fused_all,fused_staged")
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--l2gran", type=int, default=0)
    ap.add_argument("--rank-grid", default="", help="e.g. 1,1,2 (under torch.distributed.run)")
    ap.add_argument("--only", default="", help="comma separated substrings of op names to time")
    args = ap.parse_args()
    rank, world, _ = init_process_group_from_env()
    comm = make_comm()
    if args.l2gran:
        print("l2 fetch granularity (err, value):", _C.kernels.set_l2_fetch_granularity(args.l2gran))
    rg = [int(v) for v in args.rank_grid.split(",")] if args.rank_grid else None
    h = HaloExchange(comm=comm, nX=args.nx, nQ=args.nq, nGhost=3, storage_order=args.order, n_streams=1,
                     rank_grid=rg)
    quiet = rank != 0
    if quiet:
        import builtins
        builtins.print = lambda *a, **k: None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    # calibration: plain copies of one face / one whole exchange worth of bytes
    for nbytes in (18874368, 114578496):
        a = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        b = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        cp = _C.PyGpuOp("copy", lambda st, a=a, b=b, n=nbytes: _C.kernels.copy_bytes(b.data_ptr(), a.data_ptr(), n, st))
        ts = _C.device_time_steps(_C.Sequence([_C.BoundGpuOp(cp, 0)]), h.platform, args.reps,
                                  flush.data_ptr(), flush.numel())
        us = sorted(ts)[len(ts) // 2] * 1e6
        print(f"{'calibration':13s} {'plain copy ' + str(nbytes >> 20) + ' MB':32s} {us:9.2f} us  {2*nbytes/us/1e3:8.1f} GB/s r+w")
        del a, b
    # calibration: the smallest possible kernel under the same timing harness
    flag = torch.zeros(1, dtype=torch.int64, device="cuda")
    tiny = _C.PyGpuOp("tiny", lambda st: _C.kernels.flags_signal([flag.data_ptr()], 1, st))
    seq = _C.Sequence([_C.BoundGpuOp(tiny, 0)])
    ts = _C.device_time_steps(seq, h.platform, args.reps, flush.data_ptr(), flush.numel())
    print(f"{'calibration':13s} {'1-thread kernel':32s} {sorted(ts)[len(ts)//2]*1e6:9.2f} us")
    ts = _C.device_time_steps(_C.Sequence([]), h.platform, args.reps, flush.data_ptr(), flush.numel())
    print(f"{'calibration':13s} {'no kernel (event pair)':32s} {sorted(ts)[len(ts)//2]*1e6:9.2f} us")
    for style in args.styles.split(","):
        g = h.graph("flat:" + style)
        total = 0.0
        for op in g.vertices():
            if op.kind() != "GpuOp" or "wait" in op.name() or "finish" in op.name():
                continue  # a wait op run alone would spin for a flag nobody publishes
            if args.only and not any(k in op.name() for k in args.only.split(",")):
                continue
            if op.name() == "he_ready":
                continue
            bound = _C.BoundGpuOp(op, 0)
            # a direct put waits for the neighbour's permission, which the style's he_ready op grants
            pre = [_C.BoundGpuOp(g.find_name("he_ready"), 0)] if style == "direct" else []
            seq = _C.Sequence(pre + [bound])
            ts = _C.device_time_steps(seq, h.platform, args.reps, flush.data_ptr(), flush.numel())
            us = sorted(ts)[len(ts) // 2] * 1e6
            d = [int(x[2:]) for x in op.name().split("_")[-3:]] if "_dx" in op.name() else None
            remote = d is not None and h.domain.neighbor(*d) != h.rank
            nbytes = h.domain.bytes_per_exchange()
            if d is not None:
                ext = [3 if v else args.nx for v in d]
                nbytes = ext[0] * ext[1] * ext[2] * args.nq * 8
            rows.append({"style": style, "op": op.name(), "us": round(us, 2), "bytes": nbytes,
                         "gbs_rw": round(2 * nbytes / us / 1e3, 1)})
            total += us
            where = "remote" if (d is not None and remote) else "local "
            print(f"{style:13s} {op.name():32s} {where} {us:9.2f} us  {nbytes/1e6:8.2f} MB  {nbytes/us/1e3:8.1f} GB/s payload")
        print(f"{style:13s} {'SUM of put/pack ops':32s} {total:9.2f} us")
        if args.only:
            continue
        from tenzing_b200.models import common
        seq = common.fixed_schedule(h.graph(style), h.platform)
        h.platform.provision_events(seq)
        ts = _C.device_time_steps(seq, h.platform, args.reps, flush.data_ptr(), flush.numel())
        us = sorted(ts)[len(ts) // 2] * 1e6
        nb = h.domain.bytes_per_exchange()
        rows.append({"style": style, "op": "WHOLE EXCHANGE (1 stream)", "us": round(us, 2), "bytes": nb})
        print(f"{style:13s} {'WHOLE EXCHANGE (1 stream)':32s} {us:9.2f} us  {nb/1e6:8.2f} MB")
    if args.out:
        json.dump(rows, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
