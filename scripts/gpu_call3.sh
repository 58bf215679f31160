#!/bin/bash
# 2 GPUs: multi-rank correctness (all styles incl. the ones that had never run), synchronised
# message sweep, NVLink put of thick rows (LSU vs bulk), bench and style report at N = 2
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711"
echo "== multi-gpu tests (experimental styles on)"; TENZING_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/pytest_multi.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_multi.log
echo "== msg sweep n2"; timeout 300 ./build/bin/tzrun -n 2 ./build/bin/halo-msg-sweep --reps 10 > gpurun_out/msg_sweep_n2.csv 2> gpurun_out/msg_sweep_n2.err; echo "rc=$?"; cat gpurun_out/msg_sweep_n2.csv
echo "== msg sweep n2 bulk"; TZB_BULK=1 timeout 300 ./build/bin/tzrun -n 2 ./build/bin/halo-msg-sweep --reps 10 > gpurun_out/msg_sweep_n2_bulk.csv 2> gpurun_out/msg_sweep_n2_bulk.err; echo "rc=$?"; cat gpurun_out/msg_sweep_n2_bulk.csv
echo "== thick put over NVLink (y faces remote), LSU"; timeout 300 $TR scripts/microbench_halo.py --rank-grid 1,2,1 --styles fused_mixed,fused_staged --reps 9 > gpurun_out/mb_n2_lsu.log 2>&1; echo "rc=$?"; grep -v "_dx" gpurun_out/mb_n2_lsu.log | tail -12
echo "== thick put over NVLink, bulk"; TZB_BULK=1 timeout 300 $TR scripts/microbench_halo.py --rank-grid 1,2,1 --styles fused_mixed --reps 9 > gpurun_out/mb_n2_bulk.log 2>&1; echo "rc=$?"; grep -v "_dx" gpurun_out/mb_n2_bulk.log | tail -8
echo "== bench n2"; timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "rc=$?"; tail -1 gpurun_out/bench_n2.json; tail -3 gpurun_out/bench_n2.err
echo "== style report n2"; timeout 600 $TR scripts/style_report.py --reps 9 --halo-styles reference,dist,staged,fused_staged,fused_mixed,fused_all --out gpurun_out/style_report_n2.json > gpurun_out/style_report_n2.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/style_report_n2.log
