#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
echo "== halo profile target plain"; timeout 100 python scripts/profile_target.py halo > gpurun_out/pt_halo.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:box_copy -s 8 -c 8 -o gpurun_out/prof_halo python scripts/profile_target.py halo > gpurun_out/ncu_halo_full.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pt_halo.log
ls -la gpurun_out/*.ncu-rep
