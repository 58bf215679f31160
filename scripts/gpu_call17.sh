#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29801"
echo "== bench n4 ours"; (time timeout 300 $TR bench.py --gpus 4 --steps 20 --warmup 5) > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err; echo "rc=$?"; grep '^{' gpurun_out/bench_n4.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('spmv', d['value'], 'e2e', d['e2e']['value'], 'halo', d['halo']['us'], d['halo']['schedule'], 'spmv_1m', d['spmv_1m']['us'], 'random', d['spmv_random']['us'], d['correct'])"; grep real gpurun_out/bench_n4.err
echo "== nvls check (4 GPUs)"; timeout 120 $TR scripts/nvls_check.py 2>&1 | grep -E "NVLS|nvls|fused|error|Error" | head
