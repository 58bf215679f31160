#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
echo "== gpu tests (2 GPUs)"; TZ_WORKER_VERBOSE=1 timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_gpu2.log | cut -c1-250
