"""include/tenzing_b200.h is a C header: a C99 program must compile against it with -pedantic,
link against libtenzing_b200_kernels.so alone (no Python, no C++ framework) and be able to call
the host-side entry points without a GPU (plan builder + host interpreter of the plan)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "tenzing_b200")

PROGRAM = r"""
#include <stdio.h>
#include <stdlib.h>
#include "tenzing_b200.h"

int main(void) {
  /* pack a 3 x 4 x 70 region of a padded [3][6][80] grid into a dense buffer */
  enum { N3 = 3, N1 = 4, RUN = 70, E1 = 6, ROW = 80 };
  static double grid[N3][E1][ROW], dense[N3][N1][RUN];
  TzbBox box;
  TzbBoxPlanInfo info;
  void *image;
  int i3, i1, x, bad = 0;
  for (i3 = 0; i3 < N3; ++i3) for (i1 = 0; i1 < E1; ++i1) for (x = 0; x < ROW; ++x)
    grid[i3][i1][x] = 10000.0 * i3 + 100.0 * i1 + x;
  box.src = &grid[0][1][5]; box.dst = dense;
  box.s1 = ROW; box.s2 = 0; box.s3 = (int64_t)E1 * ROW;
  box.d1 = RUN; box.d2 = 0; box.d3 = (int64_t)N1 * RUN;
  box.n1 = N1; box.n2 = 1; box.n3 = N3; box.run = RUN;
  box.signal = NULL; box.wait = NULL; box.dstSlackLo = 0; box.dstSlackHi = 0; box.dstRemote = 0; box.reserved0 = 0; box.ready = NULL;
  if (tzb_boxcopy_plan(&box, 1, 8, &info, NULL) != 0) return 2;
  image = malloc(info.planBytes);
  if (tzb_boxcopy_plan(&box, 1, 8, &info, image) != 0) return 3;
  if (tzb_boxcopy_host_execute(image, info.planBytes) != info.nItems) return 4;
  for (i3 = 0; i3 < N3; ++i3) for (i1 = 0; i1 < N1; ++i1) for (x = 0; x < RUN; ++x)
    bad += dense[i3][i1][x] != grid[i3][1 + i1][5 + x];
  printf("items %d payload %lu launches %lu bad %d\n", info.nItems, (unsigned long)info.payloadBytes,
         (unsigned long)tzb_launch_count(), bad);
  free(image);
  return bad != 0;
}
"""


@pytest.mark.skipif(shutil.which("gcc") is None, reason="no C compiler")
def test_c99_program_builds_against_the_header_and_runs_without_a_gpu(tmp_path):
    if not os.path.exists(os.path.join(LIBDIR, "libtenzing_b200_kernels.so")):
        import build
        build.build()
    src = tmp_path / "user.c"
    src.write_text(PROGRAM)
    exe = tmp_path / "user"
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"),
           str(src), "-o", str(exe), "-L", LIBDIR, "-ltenzing_b200_kernels", f"-Wl,-rpath,{LIBDIR}"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "payload 6720" in r.stdout and "bad 0" in r.stdout      # 3 * 4 * 70 doubles
