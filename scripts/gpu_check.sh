#!/bin/bash
# One-GPU acceptance run on a B200 box: the GPU test tier, the smoke test, one bench line.
#   gpurun --timeout 900 -- bash scripts/gpu_check.sh
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
echo "== pytest gpu"; timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/pytest_gpu.log | cut -c1-250
echo "== smoke"; timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
echo "== bench n1"; timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?"; tail -2 gpurun_out/bench_n1.err | cut -c1-300
grep '^{' gpurun_out/bench_n1.json | python -c "
import json,sys; d=json.loads(sys.stdin.read())
print('spmv', d['value'], d['spmv'].get('us_stream_launches'), d['spmv'].get('us_cuda_graph'), 'e2e', d['e2e']['value'], 'refproto', d['spmv']['reference_protocol_pct10_us'])
for k in ('halo','spmv_1m','spmv_random'):
    print(k, d[k]['us'], d[k].get('us_stream_launches'), d[k].get('us_cuda_graph'), d[k]['correct'])
print(d['correct'], d.get('oracle_pass'))"
