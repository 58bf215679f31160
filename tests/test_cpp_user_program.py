"""INTEGRATION.md, C++ section: a user program includes tenzing/tenzing.hpp, links
build/libtenzing_b200.a and the CUDA runtime, and drives the documented flow
(init -> reproduce header -> ops -> Graph -> Platform -> State::get_decisions / apply -> solver)."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")

PROGRAM = r"""
#include <iostream>
#include "tenzing/tenzing.hpp"
using namespace tenzing;

int main(int argc, char **argv) {
  tenzing::init(argc, argv);
  reproduce::dump_with_cli(argc, argv);
  auto k1 = std::make_shared<LambdaGpuOp>("k1");
  auto k2 = std::make_shared<LambdaGpuOp>("k2");
  auto host = std::make_shared<NoOp>("host");
  Graph<OpBase> g;
  g.start_then(k1);
  g.then(k1, k2);
  g.then(k1, host);
  g.then_finish(k2);
  g.then_finish(host);
  Platform plat = Platform::make_null(2);
  SDP::State st(g);
  int steps = 0;
  while (!st.is_terminal()) {
    auto ds = st.get_decisions(plat);
    if (ds.empty()) return 2;
    st = st.apply(*ds.front());
    ++steps;
  }
  ModelBenchmarker bench([](const Sequence<BoundOp> &s) { return 1e-6 * double(s.size()); });
  dfs::Result all = dfs::explore(g, plat, bench);
  mcts::Opts opts;
  opts.nIters = 10;
  mcts::Result some = mcts::explore<mcts::FastMin>(g, plat, bench, opts);
  std::cout << "steps " << steps << " dfs " << all.simResults.size() << " mcts " << some.simResults.size()
            << " best " << all.simResults[size_t(all.best())].seq.desc() << std::endl;
  return 0;
}
"""


@pytest.mark.skipif(shutil.which("g++") is None, reason="no C++ compiler")
def test_documented_cpp_flow_compiles_links_and_runs(tmp_path):
    lib = os.path.join(ROOT, "build", "libtenzing_b200.a")
    if not os.path.exists(lib):
        import build
        build.build()
    src = tmp_path / "user.cpp"
    src.write_text(PROGRAM)
    exe = tmp_path / "user"
    cmd = ["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(CUDA, "include"),
           str(src), lib, "-o", str(exe), "-L", os.path.join(CUDA, "lib64"), "-lcudart_static", "-ldl", "-lrt",
           "-lpthread"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-3000:]
    r = subprocess.run([str(exe), "--flag"], capture_output=True, text=True, timeout=60,
                       env=dict(os.environ, TENZING_ACK_NOTICE="1"))
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.strip().splitlines()
    assert lines[0].startswith('{"args":[') and '"--flag"' in lines[0]
    assert "dfs " in lines[-1] and " mcts 10 " in lines[-1] or " mcts " in lines[-1]
