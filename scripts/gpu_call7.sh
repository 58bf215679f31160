#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
echo "== bench n1"; (time timeout 500 python bench.py --steps 20 --warmup 5) > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"; cat gpurun_out/bench_n1.json; tail -6 gpurun_out/bench_n1.err
echo "== spmv microbench"; timeout 200 python scripts/microbench_spmv.py > gpurun_out/microbench_spmv.log 2>&1; cat gpurun_out/microbench_spmv.log
