#!/bin/bash
# Host-side sanitizer runs of the native test cases (no GPU needed):
#   scripts/sanitize.sh            AddressSanitizer + UndefinedBehaviorSanitizer, then ThreadSanitizer
#   scripts/sanitize.sh asan|tsan  one of them
# The library sources and src/apps/tenzing-cpu-tests.cpp are compiled with -fsanitize=... into
# build/<kind>/ and linked against the regular CUDA objects (python build.py first).  The
# threaded cases (ranks as threads of one process, LoopbackWorld) are what TSan is for.
# The reference integrates no sanitizer.
set -euo pipefail
cd "$(dirname "$0")/.."
ROOT=$PWD
CUDA_HOME=${CUDA_HOME:-/usr/local/cuda}
[ -f build/libtenzing_b200.a ] || python build.py
kinds=${1:-"asan tsan"}
for kind in $kinds; do
  case $kind in
    asan) flags="-fsanitize=address,undefined" ;;
    tsan) flags="-fsanitize=thread" ;;
    *) echo "unknown sanitizer '$kind'" >&2; exit 64 ;;
  esac
  out=build/$kind
  mkdir -p "$out"
  for f in src/core/*.cpp src/solvers/*.cpp src/workloads/*.cpp src/apps/tenzing-cpu-tests.cpp; do
    echo "g++ -std=c++17 -O1 -g $flags -fno-omit-frame-pointer -I$ROOT/include -I$CUDA_HOME/include \
      -include $ROOT/build/tenzing_version_hash.h -c $ROOT/$f -o $ROOT/$out/$(basename "$f").o"
  done | xargs -P "$(nproc)" -I{} sh -c "{}"
  g++ $flags -o "$out/tests" "$out"/*.cpp.o build/obj/src_cuda_*.o \
    -L"$CUDA_HOME/lib64" -lcudart_static -ldl -lrt -lpthread
  echo "== $kind =="
  # protect_shadow_gap=0: the CUDA runtime reserves address space ASan also wants
  ASAN_OPTIONS=protect_shadow_gap=0:detect_leaks=1 TENZING_ACK_NOTICE=1 "$out/tests" "-tc=*[cpu]*" \
    2> "$out/report.txt" | tail -1
  if grep -q "SUMMARY: \|runtime error:" "$out/report.txt"; then
    grep "SUMMARY: \|runtime error:" "$out/report.txt" | sort | uniq -c
    exit 1
  fi
  if [ "$kind" = tsan ]; then
    # negative control: a graph that lacks an edge between two writers of one buffer MUST be
    # reported, or the clean result above would mean nothing
    TENZING_ACK_NOTICE=1 "$out/tests" "-tc=*[tsan-must-report]*" > /dev/null 2> "$out/control.txt" || true
    if grep -q "ThreadSanitizer: data race" "$out/control.txt"; then
      echo "negative control: the racy graph is reported, as it must be"
      # second control: flags published without release semantics (fault injection) must make
      # the host-executed halo exchange racy
      TENZING_FAULT=relaxed_flags TENZING_ACK_NOTICE=1 "$out/tests" "-tc=halo exchange executes on the host*" \
        > /dev/null 2> "$out/control2.txt" || true
      if grep -q "ThreadSanitizer: data race" "$out/control2.txt"; then
        echo "negative control: an exchange whose flags lack release semantics is reported"
      else
        echo "negative control FAILED: relaxed flag stores went unnoticed" >&2
        exit 1
      fi
    else
      echo "negative control FAILED: ThreadSanitizer did not report the racy graph" >&2
      exit 1
    fi
  fi
done
echo "sanitizers: clean"
