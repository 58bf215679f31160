#!/usr/bin/env python
"""Device time of the CSR product kernel on one GPU (L2 flushed between runs) against its
HBM roofline, for the matrix shapes of the benchmark configs."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("TENZING_ACK_NOTICE", "1")
import torch  # noqa: E402

from tenzing_b200 import ops  # noqa: E402


def make_csr(rows, cols, nnz_per_row, band, dtype, seed=0):
    """~nnz_per_row entries per row (varying 0.5x..1.5x), columns uniform in row +- band"""
    g = torch.Generator(device="cuda").manual_seed(seed)
    lo, hi = max(1, nnz_per_row // 2), nnz_per_row + nnz_per_row // 2
    counts = torch.randint(lo, hi + 1, (rows,), generator=g, device="cuda")
    row_ptr = torch.zeros(rows + 1, dtype=torch.int64, device="cuda")
    row_ptr[1:] = torch.cumsum(counts, 0)
    nnz = int(row_ptr[-1])
    r = torch.repeat_interleave(torch.arange(rows, device="cuda"), counts)
    off = torch.randint(-band, band + 1, (nnz,), generator=g, device="cuda")
    col = ((r * cols) // rows + off).clamp_(0, cols - 1)
    # sort the columns inside each row, as a CSR built from a sorted COO would be
    key = r * cols + col
    col = (torch.sort(key).values - r * cols).to(torch.int32)
    val = torch.rand(nnz, generator=g, device="cuda", dtype=torch.float64).to(dtype)
    return row_ptr.to(torch.int32), col, val


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = peaks.get("hbm_gbs", 6650.0) * 1e9
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows_out = []
    cases = [("band 1M, bw=n (config 2 at 1 GPU)", 1_000_000, 1_000_000, 10, 1_000_000),
             ("band 1M/8 rows, bw=n/8 (config 2 at 8 GPUs, local part)", 125_000, 125_000, 10, 125_000),
             ("random 1M x 8M (config 4 at 8 GPUs, one rank)", 1_000_000, 8_000_000, 10, 8_000_000),
             ("long rows 100k x 100k, 64 nnz/row", 100_000, 100_000, 64, 100_000)]
    for name, rows, cols, npr, band in cases:
        for dtype in (torch.float64, torch.float32):
            rp, ci, v = make_csr(rows, cols, npr, band, dtype)
            x = torch.rand(cols, device="cuda", dtype=torch.float64).to(dtype)
            y = torch.zeros(rows, device="cuda", dtype=dtype)
            es = v.element_size()
            nbytes = v.numel() * (es + 4) + rows * (4 + es) + min(cols, v.numel()) * es
            ts = []
            for i in range(args.reps + 3):
                flush.fill_(i & 0xff)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                ops.spmv_csr(rp, ci, v, x, y=y)
                e1.record()
                torch.cuda.synchronize()
                if i >= 3:
                    ts.append(e0.elapsed_time(e1) * 1e3)
            ts.sort()
            us = ts[len(ts) // 2]
            # check against torch (fp64 accumulate)
            r = torch.repeat_interleave(torch.arange(rows, device="cuda"), (rp[1:] - rp[:-1]).long())
            ref = torch.zeros(rows, device="cuda", dtype=torch.float64)
            ref.index_add_(0, r, v.double() * x.double()[ci.long()])
            err = float((y.double() - ref).abs().max() / ref.abs().max())
            roof = nbytes / hbm * 1e6
            print(f"{name:58s} {str(dtype)[6:]:8s} {us:8.2f} us (min {ts[0]:7.2f})  {nbytes/us/1e3:7.1f} GB/s  "
                  f"roofline {roof:6.2f} us  frac {roof/us:5.2f}  rel.err {err:.1e}")
            rows_out.append({"case": name, "dtype": str(dtype), "us": us, "bytes": nbytes,
                             "roofline_us": roof, "frac": roof / us, "rel_err": err})
            del rp, ci, v, x, y
    if args.out:
        json.dump(rows_out, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
