#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29791"
for v in "TZB_LINK_BLOCKS_PER_SM=2" "TZB_LINK_BLOCKS_PER_SM=1"; do
  echo "== $v"
  env $v timeout 150 $TR scripts/style_report.py --reps 9 --no-spmv --halo-styles fused_mixed --mixed-hand 2>&1 | grep "halo "
done
echo "== bench n8 ours"; (time timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 5) > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "rc=$?"; grep '^{' gpurun_out/bench_n8.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('spmv', d['value'], 'e2e', d['e2e']['value'], 'halo', d['halo']['us'], d['halo']['schedule'], 'spmv_1m', d['spmv_1m']['us'], 'random', d['spmv_random']['us'], d['correct'])"
