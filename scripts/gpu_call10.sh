#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29761"
for v in "default" "TZB_THIN_SECTOR=0" "TZB_NO_READY=1" "TZB_THIN_SECTOR=0 TZB_NO_READY=1"; do
  echo "== $v"
  env $( [ "$v" = default ] || echo $v ) timeout 200 $TR scripts/style_report.py --reps 9 --no-spmv --rank-grid 1,2,1 --halo-styles fused_mixed,fused_staged --mixed-hand 2>&1 | grep "halo "
done
