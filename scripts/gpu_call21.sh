#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29831"
(time timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 5) > gpurun_out/bench_nX.json 2> gpurun_out/bench_nX.err; echo "rc=$?"; grep real gpurun_out/bench_nX.err; tail -3 gpurun_out/bench_nX.err | cut -c1-300
grep '^{' gpurun_out/bench_nX.json | python -c "
import json,sys; d=json.loads(sys.stdin.read())
print('spmv', d['value'], d['spmv']['launch'], 'halo', d['halo']['us'], 'correct', d['correct'])
print(json.dumps(d.get('msg_sweep')))
print(json.dumps(d.get('oracle_pass')))"
