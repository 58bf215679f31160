#!/bin/bash
# reference arm on 1 and 2 GPUs (the MPI stand-in's device path runs here for the first time)
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721"
echo "== reference N=1"; timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/ref_n1.json 2> gpurun_out/ref_n1.err; echo "rc=$?"; cat gpurun_out/ref_n1.json; tail -5 gpurun_out/ref_n1.err
echo "== reference N=2"; timeout 900 $TR bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > gpurun_out/ref_n2.json 2> gpurun_out/ref_n2.err; echo "rc=$?"; tail -2 gpurun_out/ref_n2.json; tail -5 gpurun_out/ref_n2.err
ls baseline/_ref/build/tenzing-mcts/examples/ | head
