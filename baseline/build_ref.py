#!/usr/bin/env python
"""Build the UNMODIFIED reference (sandialabs/tenzing, /root/reference) into baseline/_ref/ with
its own CMake build, against the single-node MPI stand-in of baseline/mpi_shim (this image has
no MPI).  Only the build hazards of SURVEY App. C are handled, all outside the reference's
sources:

  * MPI:      baseline/mpi_shim (mpi.h, libmpi.so, an mpicxx wrapper for CMake's FindMPI)
  * numa.h:   a declarations-only stand-in next to mpi.h (the reference never calls libnuma)
  * version:  the copy of the source tree is made a git checkout tagged v0.1.0, because the
              reference generates its version header from `git describe --tags`
  * arch:     -DCMAKE_CUDA_ARCHITECTURES=100a

  python baseline/build_ref.py      ->  baseline/_ref/build/tenzing-dfs/examples/spmv etc.
"""
import hashlib
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("TENZING_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
SHIM = os.path.join(HERE, "mpi_shim")
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def run(cmd, **kw):
    print("+", " ".join(cmd) if isinstance(cmd, list) else cmd, flush=True)
    subprocess.run(cmd, check=True, **kw)


def host_id():
    """CPU model of this machine: the reference adds -march=native, so a build is only valid on
    the same kind of CPU"""
    model = ""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name") or line.startswith("flags"):
                model += line
                if line.startswith("flags"):
                    break
    except OSError:
        pass
    return hashlib.sha1(model.encode()).hexdigest()[:16]


def build_shim():
    inc, lib, bind = (os.path.join(OUT, "mpi", d) for d in ("include", "lib", "bin"))
    for d in (inc, lib, bind):
        os.makedirs(d, exist_ok=True)
    for h in ("mpi.h", "numa.h"):
        shutil.copy(os.path.join(SHIM, "include", h), inc)
    run(["g++", "-std=c++17", "-O2", "-g", "-fPIC", "-shared", f"-I{inc}", f"-I{CUDA_HOME}/include",
         os.path.join(SHIM, "src", "mpi_shim.cpp"), "-o", os.path.join(lib, "libmpi.so"),
         f"-L{CUDA_HOME}/lib64", "-lcudart", "-lrt", "-lpthread", f"-Wl,-rpath,{CUDA_HOME}/lib64"])
    wrapper = os.path.join(bind, "mpicxx")
    with open(wrapper, "w") as f:
        f.write(f"""#!/bin/bash
# compiler wrapper of the MPI stand-in: answers the questions CMake's FindMPI asks and
# otherwise compiles with g++ against the stand-in
INC="{inc}"; LIB="{lib}"
case "$1" in
  -showme:compile|-compile_info|-compile-info) echo "-I$INC"; exit 0;;
  -showme:link|-link_info|-link-info) echo "-L$LIB -Wl,-rpath,$LIB -lmpi"; exit 0;;
  -showme:incdirs) echo "$INC"; exit 0;;
  -showme:libdirs) echo "$LIB"; exit 0;;
  -showme:libs) echo "mpi"; exit 0;;
  -showme:version) echo "tz-mpi-shim 1.0"; exit 0;;
  -showme|-show) echo "g++ -I$INC -L$LIB -Wl,-rpath,$LIB -lmpi"; exit 0;;
esac
exec g++ -I"$INC" "$@" -L"$LIB" -Wl,-rpath,"$LIB" -lmpi
""")
    os.chmod(wrapper, 0o755)
    return inc, lib, wrapper


def tag_checkout(dst, message):
    git = ["git", "-c", "user.name=baseline", "-c", "user.email=baseline@localhost", "-c", "init.defaultBranch=main"]
    run(git + ["init", "-q"], cwd=dst)
    run(git + ["add", "-A"], cwd=dst)
    run(git + ["commit", "-q", "-m", message], cwd=dst)
    run(git + ["tag", "v0.1.0"], cwd=dst)


def checkout(dst):
    """copy of the reference sources as a tagged git checkout (its CMake derives the version
    header from `git describe --tags`)"""
    if not os.path.isdir(REF_SRC):
        # on a GPU box only the copy that travelled with the repository exists; if it travelled
        # without its .git directory it is made a tagged checkout again
        if os.path.isdir(os.path.join(dst, ".git")):
            return
        if os.path.isfile(os.path.join(dst, "CMakeLists.txt")):
            tag_checkout(dst, "reference snapshot")
            return
        raise SystemExit(f"reference sources not found: neither {REF_SRC} nor {dst}")
    if os.path.exists(dst):
        shutil.rmtree(dst)
    shutil.copytree(REF_SRC, dst, symlinks=True, ignore=shutil.ignore_patterns(".git"))
    run(["chmod", "-R", "u+w", dst])
    tag_checkout(dst, "reference snapshot")


def cmake_build(src, bld, wrapper, targets):
    os.makedirs(bld, exist_ok=True)
    run(["cmake", "-S", src, "-B", bld, "-G", "Ninja",
         "-DCMAKE_CUDA_ARCHITECTURES=100a", f"-DCMAKE_CUDA_COMPILER={CUDA_HOME}/bin/nvcc",
         f"-DMPI_CXX_COMPILER={wrapper}", "-DMPI_CXX_SKIP_MPICXX=ON",
         "-DTENZING_ENABLE_TESTS=OFF", "-DCMAKE_BUILD_TYPE=Release"])
    run(["ninja", "-C", bld] + list(targets))


def main():
    os.makedirs(OUT, exist_ok=True)
    inc, lib, wrapper = build_shim()
    src = os.path.join(OUT, "src")
    checkout(src)
    shutil.rmtree(os.path.join(OUT, "build"), ignore_errors=True)  # paths of another machine in its cache
    cmake_build(src, os.path.join(OUT, "build"), wrapper, ["spmv-min-time", "spmv", "halo-mcts-min-time"])
    # the end-to-end driver: a user program of the reference's API, built with the reference as a
    # CMake sub-project so that the reference's own files decide every flag.  The reference's
    # git module only accepts a .git directory below the top-level source directory, hence the
    # tagged checkout is copied below the driver's sources.
    try:
        drv, bld = os.path.join(OUT, "e2e_src"), os.path.join(OUT, "build_e2e")
        for d in (drv, bld):
            shutil.rmtree(d, ignore_errors=True)
        shutil.copytree(os.path.join(HERE, "ref_driver"), drv)
        shutil.copytree(src, os.path.join(drv, "tenzing"), symlinks=True)
        src = os.path.join(drv, "tenzing")
        os.makedirs(bld, exist_ok=True)
        run(["cmake", "-S", drv, "-B", bld, "-G", "Ninja", f"-DTENZING_REF_SRC={src}",
             "-DCMAKE_CUDA_ARCHITECTURES=100a", f"-DCMAKE_CUDA_COMPILER={CUDA_HOME}/bin/nvcc",
             f"-DMPI_CXX_COMPILER={wrapper}", "-DMPI_CXX_SKIP_MPICXX=ON", "-DTENZING_ENABLE_TESTS=OFF",
             "-DCMAKE_BUILD_TYPE=Release"])
        run(["ninja", "-C", bld, "ref-e2e", "ref-enum"])
    except subprocess.CalledProcessError as e:
        print("the end-to-end driver did not build (the stock drivers are unaffected):", e)
    with open(os.path.join(OUT, "BUILD_HOST"), "w") as f:
        f.write(host_id() + "\n")
    print("reference built in", OUT)


if __name__ == "__main__":
    main()
