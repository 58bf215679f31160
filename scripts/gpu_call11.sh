#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29771"
for v in "default" "TZB_NO_READY=1"; do
  echo "== $v"
  env $( [ "$v" = default ] || echo $v ) timeout 150 $TR scripts/style_report.py --reps 9 --no-spmv --halo-styles fused_mixed --mixed-hand 2>&1 | grep "halo "
done
