#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
echo "== graph replay tests (1 GPU part)"; timeout 300 python -m pytest tests/test_gpu_graph_replay.py -x -q > gpurun_out/pytest_graph.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_graph.log | cut -c1-300
echo "== multi gpu worker"; timeout 300 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/pytest_multi.log | cut -c1-300
