#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29781"
for v in "TZB_LINK_BLOCKS_PER_SM=1" "TZB_LINK_BLOCKS_PER_SM=2" "TZB_LINK_BLOCKS_PER_SM=4"; do
  echo "== $v"
  env $v timeout 200 $TR scripts/style_report.py --reps 9 --no-spmv --rank-grid 1,2,1 --halo-styles fused_mixed --mixed-hand 2>&1 | grep "halo "
done
echo "== multi tests"; TENZING_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_gpu_multi.py tests/test_gpu_workloads.py -x -q 2>&1 | tail -3
