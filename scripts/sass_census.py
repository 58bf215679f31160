#!/usr/bin/env python
"""SASS census of the kernel library: per kernel, the memory / synchronisation / tensor / TMA
instructions in the sm_100a SASS (cuobjdump -sass).  Writes profiles/sass_census.txt and the
listings of the measured kernels under profiles/sass/.

  python scripts/sass_census.py [path/to/libtenzing_b200_kernels.so]
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tenzing_b200", "libtenzing_b200_kernels.so")
KEEP = re.compile(r"^(LD|ST|ATOM|RED|MEMBAR|ERRBAR|CCTL|BAR|NANOSLEEP|SHFL|UBLKCP|UTMA|SYNCS|HMMA|IMMA|DMMA|"
                  r"UTC|FENCE|LDSM|STSM|LDGSTS|UCGABAR)")
HEADER = """# SASS census of libtenzing_b200_kernels.so (cuobjdump -sass, sm_100a): memory / sync / math
# instructions per kernel (scripts/sass_census.py).  What to look for: ST*.STRONG.SYS +
# MEMBAR.*.SYS + ERRBAR = the release side of the cross-GPU flag protocol; LD*.STRONG.SYS +
# NANOSLEEP = the acquire spin; LDG/ST .128 = 16-byte accesses on aligned sides; UBLKCP.S.G /
# UBLKCP.G.S + SYNCS.*.TRANS64 = the bulk-copy engine (cp.async.bulk + mbarrier) of the
# opt-in box_bulk_kernel; LDG/STG.E.ENL2.256 = 256-bit accesses of the thin-row unpack; no HMMA/UTC*MMA anywhere (bandwidth-bound kernels, no tensor-core
# work).
"""


def demangle(names):
    try:
        out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True, check=True).stdout
        return out.splitlines()
    except Exception:
        return names


def strip_params(name):
    """drop the parameter list: cut at the first '(' outside the template argument brackets"""
    depth = 0
    for i, c in enumerate(name):
        if c == "<":
            depth += 1
        elif c == ">":
            depth -= 1
        elif c == "(" and depth == 0:
            return name[:i]
    return name


# the listings kept under profiles/sass/: file name -> the demangled kernel it holds (fp64 instances)
LISTINGS = {
    "box_copy_kernel_f64": "box_copy_kernel<double, (bool)0>",
    "box_copy_kernel_f64_vec": "box_copy_kernel<double, (bool)1>",
    "box_copy_sync_kernel_f64": "box_copy_sync_kernel<double, (int)4>",
    "box_copy_sync_kernel_f64_ordered": "box_copy_sync_kernel<double, (int)2>",
    "box_bulk_kernel": "box_bulk_kernel",
    "flags_signal_kernel": "flags_signal_kernel",
    "flags_wait_kernel": "flags_wait_kernel",
    "gather_put_kernel_f64": "gather_put_kernel<double>",
    "multicast_put_kernel_f64": "multicast_put_kernel<double>",
    "spmv_csr_pull_kernel_f64_tpr8": "spmv_csr_pull_kernel<double, (int)8>",
    "spmv_csr_stream_kernel_f64_u12": "spmv_csr_stream_kernel<double, (int)12, (bool)0>",
    "spmv_csr_stream_kernel_f64_u12_wait": "spmv_csr_stream_kernel<double, (int)12, (bool)1>",
    "spmv_fused_all_kernel_f64_u12": "spmv_fused_all_kernel<double, (int)12>",
}


def write_listings(sass):
    """profiles/sass/<name>.sass: the instruction stream of each kernel of LISTINGS, encodings stripped"""
    blocks, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            blocks[cur] = []
            continue
        if cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            blocks[cur].append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line))
    names = demangle(list(blocks))
    outdir = os.path.join(ROOT, "profiles", "sass")
    os.makedirs(outdir, exist_ok=True)
    written = 0
    for fname, want in LISTINGS.items():
        for mangled, name in zip(blocks, names):
            short = strip_params(re.sub(r"^void ", "", name))
            if short.endswith("::" + want) or short == want:
                with open(os.path.join(outdir, fname + ".sass"), "w") as f:
                    f.write("// " + name + "\n" + "\n".join(blocks[mangled]) + "\n")
                written += 1
                break
        else:
            print("no kernel matches", want)
    print(f"{written} listings -> {outdir}")


def write_resources():
    """profiles/kernel_resources.txt: registers, stack (spills), static shared memory per kernel"""
    text = subprocess.run(["cuobjdump", "--dump-resource-usage", LIB], capture_output=True, text=True, check=True).stdout
    rows, mangled = [], None
    for line in text.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            mangled = m.group(1)
            continue
        m = re.match(r"\s*REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and mangled:
            rows.append((mangled,) + tuple(int(v) for v in m.groups()))
            mangled = None
    names = demangle([r[0] for r in rows])
    path = os.path.join(ROOT, "profiles", "kernel_resources.txt")
    with open(path, "w") as f:
        f.write("# cuobjdump --dump-resource-usage of libtenzing_b200_kernels.so (sm_100a): registers per thread,\n"
                "# stack bytes per thread (non-zero = spills or local arrays), static shared memory bytes per block\n"
                "# (scripts/sass_census.py).  64 registers x 256 threads = 4 blocks per SM; 32 x 256 = 8.\n"
                "# box_copy_sync_kernel<*, 2> (the ordered plans of fused_groups / fused_pipelined) is held to 64\n"
                "# registers so that its partner kernel is guaranteed SM slots beside it; it pays 120 bytes of stack.\n")
        for (m_, reg, stack, shared, local), name in zip(rows, names):
            short = strip_params(re.sub(r"^void ", "", re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", name)))
            f.write(f"{short}: REG {reg} STACK {stack} SHARED {shared} LOCAL {local}\n")
    print(f"{len(rows)} kernels -> {path}")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    write_listings(sass)
    write_resources()
    kernels, cur = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
        if m and cur and KEEP.match(m.group(1)) and not m.group(1).startswith(("LDC", "LDCU")):
            kernels[cur][m.group(1)] += 1
    names = demangle(list(kernels))
    lines = [HEADER]
    for mangled, name in zip(kernels, names):
        short = re.sub(r"\(anonymous namespace\)::|<unnamed>::", "", name)
        short = strip_params(re.sub(r"^void ", "", short))
        short = short.replace("tzb::", "", 1) if short.startswith("tzb::box") or short.startswith("tzb::spmv") or short.startswith("tzb::gather") else short
        ops = ", ".join(f"{k} x{v}" for k, v in sorted(kernels[mangled].items()))
        lines.append(f"{short}\n    {ops}")
    path = os.path.join(ROOT, "profiles", "sass_census.txt")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    print(f"{len(kernels)} kernels -> {path}")


if __name__ == "__main__":
    main()
