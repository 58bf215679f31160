#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $NGPU --master-addr 127.0.0.1 --master-port 29821"; echo "== bench n$NGPU"; (time timeout 400 $TR bench.py --gpus $NGPU --steps 20 --warmup 5) > gpurun_out/bench_nX.json 2> gpurun_out/bench_nX.err; echo "rc=$?"; grep real gpurun_out/bench_nX.err
grep '^{' gpurun_out/bench_nX.json | python -c "
import json,sys; d=json.loads(sys.stdin.read())
print('spmv', d['value'], d['spmv'].get('us_stream_launches'), d['spmv'].get('us_cuda_graph'), d['spmv']['launch'], 'e2e', d['e2e']['value'], 'launches', d['gpu_launches'])
for k in ('halo','spmv_1m','spmv_random'):
    print(k, d[k]['us'], d[k].get('us_stream_launches'), d[k].get('us_cuda_graph'), d[k]['launch'], d[k]['correct'])
print(d['correct'])"
