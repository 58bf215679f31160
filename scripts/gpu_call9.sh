#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29751"
echo "== bench n8 ours"; (time timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 5) > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "rc=$?"; grep '^{' gpurun_out/bench_n8.json; grep real gpurun_out/bench_n8.err
echo "== reference n8"; timeout 200 $TR bench.py --impl reference --gpus 8 --ref-seconds 100 > gpurun_out/ref_n8.json 2> gpurun_out/ref_n8.err; echo "rc=$?"; grep '^{' gpurun_out/ref_n8.json
echo "== multi-gpu test 8"; TENZING_TEST_EXPERIMENTAL=1 timeout 250 python -m pytest tests/test_gpu_multi.py -x -q -k eight > gpurun_out/pytest_multi8.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/pytest_multi8.log
echo "== style report n8"; timeout 200 $TR scripts/style_report.py --reps 9 --no-spmv --halo-styles fused_staged,fused_mixed,fused_all,fused_pipelined --out gpurun_out/style_report_n8.json > gpurun_out/style_report_n8.log 2>&1; echo "rc=$?"; grep "halo\|spmv" gpurun_out/style_report_n8.log
