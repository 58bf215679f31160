#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29811"
for v in "TZB_WIDE256=0" "TZB_WIDE256=1"; do
  echo "== $v"
  env $v timeout 200 $TR scripts/microbench_halo.py --rank-grid 1,2,1 --styles fused_mixed --only he_put_thick --reps 15 2>&1 | grep -E "he_put_thick|plain copy 109"
  env $v timeout 200 $TR scripts/style_report.py --reps 9 --no-spmv --rank-grid 1,2,1 --halo-styles fused_mixed --mixed-hand 2>&1 | grep "halo "
done
env TZB_WIDE256=1 timeout 300 python -m pytest tests/test_gpu_workloads.py tests/test_gpu_boxcopy_thin.py -x -q 2>&1 | tail -2
