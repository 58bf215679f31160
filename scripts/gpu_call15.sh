#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
echo "== halo profile target plain"; timeout 200 python scripts/profile_target.py halo > gpurun_out/pt_halo.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:box_copy -s 7 -c 7 -o gpurun_out/prof_halo python scripts/profile_target.py halo > gpurun_out/ncu_halo_full.log 2>&1; echo "rc=$?"
echo "== spmv profile target plain"; timeout 200 python scripts/profile_target.py spmv > gpurun_out/pt_spmv.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmv -s 3 -c 3 -o gpurun_out/prof_spmv python scripts/profile_target.py spmv > gpurun_out/ncu_spmv_full.log 2>&1; echo "rc=$?"
ls -la gpurun_out/*.ncu-rep
