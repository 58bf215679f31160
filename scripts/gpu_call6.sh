#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731"
echo "== reference N=2 (DFS driver, 100 s)"; timeout 200 $TR bench.py --impl reference --gpus 2 --ref-seconds 100 > gpurun_out/ref_n2.json 2> gpurun_out/ref_n2.err; echo "rc=$?"; tail -2 gpurun_out/ref_n2.json; tail -4 gpurun_out/ref_n2.err | cut -c1-300
