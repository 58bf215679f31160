#!/usr/bin/env python
"""In-tree native build: C++17 host framework + sm_100a CUDA kernels.

Generates build/build.ninja and runs ninja.  Products:
  build/libtenzing_b200.a                      core + solvers + workloads + kernels
  tenzing_b200/_C.<abi>.so                     pybind11 module (what Python imports)
  tenzing_b200/libtenzing_b200_kernels.so      C-ABI kernel library (include/tenzing_b200.h)
  build/bin/<name>                             drivers from src/apps/

nvcc cross-compiles for sm_100a without a GPU, so this runs on the CPU-only
build box; the .so files are git-ignored but travel to the GPU box with the tree.

  python build.py            # incremental build of everything
  python build.py --clean    # remove build/ and the in-tree .so files first
  python build.py -v         # show compiler command lines
"""
import glob
import os
import shutil
import subprocess
import sys
import sysconfig

ROOT = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(ROOT, "build")
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")
ARCH_FLAGS = "-gencode arch=compute_100a,code=sm_100a"


def _git_hash():
    try:
        return subprocess.check_output(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"],
                                       stderr=subprocess.DEVNULL).decode().strip()
    except Exception:
        return "unknown"


def _rel(p):
    return os.path.relpath(p, BUILD)


def generate():
    import pybind11

    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(os.path.join(BUILD, "bin"), exist_ok=True)
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    pyinc = sysconfig.get_paths()["include"]
    inc = f"-I{ROOT}/include -I{CUDA_HOME}/include"
    # the version hash lives in a generated header so a new commit does not
    # invalidate every object file
    vh = os.path.join(BUILD, "tenzing_version_hash.h")
    content = f'#define TENZING_VERSION_HASH "{_git_hash()}"\n'
    if not os.path.exists(vh) or open(vh).read() != content:
        open(vh, "w").write(content)

    # optional debug macros, e.g. TENZING_DEFINES="TENZING_RUNTIME_SANITY_CHECKS TENZING_RUNTIME_CHECK_BOUNDS"
    # (graph invariants after every rewrite / device-side bounds checks in ArrayView)
    defs = " ".join("-D" + d for d in os.environ.get("TENZING_DEFINES", "").split())
    cxxflags = (f"-std=c++17 -O2 -g1 -fPIC -Wall -Wextra -Wno-unused-parameter {inc} -include {vh} "
                f"{defs}").rstrip()
    nvccflags = (f"-std=c++17 -O3 {ARCH_FLAGS} -lineinfo -Xcompiler -fPIC "
                 f"--expt-relaxed-constexpr -Xptxas -v {inc} -include {vh} {defs}").rstrip()
    ldflags = f"-L{CUDA_HOME}/lib64 -lcudart_static -ldl -lrt -lpthread"

    lib_cpp = sorted(glob.glob(f"{ROOT}/src/core/*.cpp") + glob.glob(f"{ROOT}/src/solvers/*.cpp") +
                     glob.glob(f"{ROOT}/src/workloads/*.cpp"))
    lib_cu = sorted(glob.glob(f"{ROOT}/src/cuda/*.cu") + glob.glob(f"{ROOT}/src/workloads/*.cu"))
    bind_cpp = sorted(glob.glob(f"{ROOT}/src/bindings/*.cpp"))
    apps = sorted(glob.glob(f"{ROOT}/src/apps/*.cpp") + glob.glob(f"{ROOT}/src/apps/*.cu") +
                  glob.glob(f"{ROOT}/scripts/probes/*.cu"))  # probes: stand-alone measurement programs

    out = []
    w = out.append
    w("ninja_required_version = 1.3")
    w(f"cxxflags = {cxxflags}")
    w(f"nvccflags = {nvccflags}")
    w(f"ldflags = {ldflags}")
    w("rule cxx\n  command = g++ $cxxflags $extra -MMD -MF $out.d -c $in -o $out\n"
      "  depfile = $out.d\n  deps = gcc\n  description = CXX $in")
    w(f"rule nvcc\n  command = {CUDA_HOME}/bin/nvcc $nvccflags $extra -MD -MF $out.d -c $in -o $out "
      "2> $out.ptxas.log || (cat $out.ptxas.log >&2; exit 1)\n"
      "  depfile = $out.d\n  deps = gcc\n  description = NVCC $in")
    w("rule ar\n  command = rm -f $out && ar rcs $out $in\n  description = AR $out")
    w("rule solink\n  command = g++ -shared -o $out $in $libs $ldflags\n  description = SOLINK $out")
    w("rule link\n  command = g++ -o $out $in $libs $ldflags -rdynamic\n  description = LINK $out")

    def obj(src):
        return os.path.join(BUILD, "obj", os.path.relpath(src, ROOT).replace("/", "_") + ".o")

    lib_objs = []
    for s in lib_cpp:
        w(f"build {obj(s)}: cxx {s}")
        lib_objs.append(obj(s))
    cu_objs = []
    for s in lib_cu:
        w(f"build {obj(s)}: nvcc {s}")
        lib_objs.append(obj(s))
        cu_objs.append(obj(s))
    lib = os.path.join(BUILD, "libtenzing_b200.a")
    w(f"build {lib}: ar {' '.join(lib_objs)}")

    bind_objs = []
    for s in bind_cpp:
        w(f"build {obj(s)}: cxx {s}\n  extra = -I{pybind11.get_include()} -I{pyinc} "
          "-fvisibility=hidden")
        bind_objs.append(obj(s))
    pymod = os.path.join(ROOT, "tenzing_b200", "_C" + ext)
    if bind_objs:
        w(f"build {pymod}: solink {' '.join(bind_objs)} | {lib}\n"
          f"  libs = -Wl,--whole-archive {lib} -Wl,--no-whole-archive")

    # plain C-ABI kernel library (no Python, no C++ framework symbols needed by callers)
    capi = os.path.join(ROOT, "tenzing_b200", "libtenzing_b200_kernels.so")
    if cu_objs:
        w(f"build {capi}: solink {' '.join(cu_objs)}")

    bins = []
    for s in apps:
        name = os.path.splitext(os.path.basename(s))[0]
        rule = "nvcc" if s.endswith(".cu") else "cxx"
        w(f"build {obj(s)}: {rule} {s}")
        b = os.path.join(BUILD, "bin", name)
        w(f"build {b}: link {obj(s)} | {lib}\n  libs = {lib}")
        bins.append(b)

    targets = [lib] + ([pymod] if bind_objs else []) + ([capi] if cu_objs else []) + bins
    w(f"default {' '.join(targets)}")
    text = "\n".join(out) + "\n"
    path = os.path.join(BUILD, "build.ninja")
    if not os.path.exists(path) or open(path).read() != text:
        open(path, "w").write(text)
    return path


def clean():
    shutil.rmtree(BUILD, ignore_errors=True)
    for f in glob.glob(f"{ROOT}/tenzing_b200/*.so"):
        os.remove(f)


def build(verbose=False, jobs=None):
    path = generate()
    cmd = ["ninja", "-f", path]
    if verbose:
        cmd.append("-v")
    if jobs:
        cmd += ["-j", str(jobs)]
    r = subprocess.run(cmd, cwd=BUILD)
    if r.returncode != 0:
        raise RuntimeError("native build failed")


def ptxas_report():
    """registers / spills / smem per kernel, from the logs nvcc -Xptxas -v left behind"""
    rep = []
    for log in sorted(glob.glob(os.path.join(BUILD, "obj", "*.ptxas.log"))):
        rep.append(f"== {os.path.basename(log)}")
        rep.extend(l.rstrip() for l in open(log) if "ptxas info" in l or "warning" in l.lower())
    return "\n".join(rep)


if __name__ == "__main__":
    if "--clean" in sys.argv:
        clean()
    build(verbose="-v" in sys.argv)
    if "--ptxas" in sys.argv:
        print(ptxas_report())
