"""The examples that need no GPU must keep running."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.timeout(300)
@pytest.mark.parametrize("script,expect", [
    ("01_noop_graph.py", None),
    ("02_streams_and_syncs.py", None),
    ("04_run_without_gpu.py", "results match the oracles on all 4 ranks"),
    ("05_resume_a_search.py", "replayed from the file"),
])
def test_example_runs(script, expect):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", TENZING_ACK_NOTICE="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "examples", script)], capture_output=True,
                       text=True, timeout=280, env=env, cwd=ROOT)
    assert p.returncode == 0, (p.stdout + p.stderr)[-2000:]
    if expect:
        assert expect in p.stdout
