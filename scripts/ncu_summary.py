#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): per kernel the duration, DRAM bytes and
throughput, cache throughputs, occupancy, registers and the top stall locations."""
import csv
import io
import subprocess
import sys


def run(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main(rep, out):
    raw = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units, rows = raw[0], raw[1], raw[2:]
    want = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "dram__throughput.avg.pct_of_peak_sustained_elapsed",
            "l1tex__throughput.avg.pct_of_peak_sustained_active",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size",
            "launch__registers_per_thread", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio"]
    idx = [(w, hdr.index(w)) for w in want if w in hdr]
    lines = [f"# ncu summary of {rep} (captured with --set full --clock-control none on a B200; "
             "cold-cache, serialised: compare shares, not absolutes)", ""]
    for k, r in enumerate(rows):
        lines.append(f"## launch {k}: {r[hdr.index('Kernel Name')][:110]}")
        for w, i in idx[1:]:
            lines.append(f"  {w:78s} {r[i]:>16s} {units[i]}")
        try:
            rd = float(r[hdr.index('dram__bytes_read.sum')])
            wr = float(r[hdr.index('dram__bytes_write.sum')])
            us = float(r[hdr.index('gpu__time_duration.sum')])
            u1, u2 = units[hdr.index('dram__bytes_read.sum')], units[hdr.index('gpu__time_duration.sum')]
            if u1 == "Mbyte" and u2 == "us":
                lines.append(f"  {'=> DRAM traffic / duration':78s} {(rd + wr) / us:16.3f} TB/s")
        except Exception:
            pass
        src = list(csv.reader(io.StringIO(run(["-i", rep, "--page", "source", "--csv", "--kernel-id",
                                               f"::regex:.*:{k + 1}"]))))
        h = None
        for i, row in enumerate(src):
            if "Source" in row and "# Samples" in row:
                h, start = row, i + 1
                break
        if h:
            si, ki = h.index("Source"), h.index("# Samples")
            seen, data = set(), []
            for row in src[start:]:
                try:
                    if row[0] in seen:
                        continue
                    seen.add(row[0])
                    data.append((float(row[ki]), row[si]))
                except Exception:
                    pass
            tot = sum(d[0] for d in data) or 1
            lines.append("  top stall locations (SASS, % of samples):")
            for v, s in sorted(data, reverse=True)[:6]:
                lines.append(f"    {100 * v / tot:5.1f}%  {s[:100]}")
        lines.append("")
    open(out, "w").write("\n".join(lines))
    print("wrote", out)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
