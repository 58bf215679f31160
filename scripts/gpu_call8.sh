#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29741"
echo "== gpu tests (2 GPUs, experimental on)"; TENZING_TEST_EXPERIMENTAL=1 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/pytest_gpu2.log
echo "== bench n2 ours"; (time timeout 500 $TR bench.py --gpus 2 --steps 20 --warmup 5) > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "rc=$?"; grep '^{' gpurun_out/bench_n2.json; grep real gpurun_out/bench_n2.err
