#!/bin/bash
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
echo "== pytest gpu"; timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
echo "== halo microbench sector"; timeout 300 python scripts/microbench_halo.py --styles fused_all,fused_staged,fused_mixed --reps 9 > gpurun_out/mb_sector.log 2>&1; echo "rc=$?"; grep -v "_dx" gpurun_out/mb_sector.log
echo "== halo microbench element (TZB_THIN_SECTOR=0)"; TZB_THIN_SECTOR=0 timeout 300 python scripts/microbench_halo.py --styles fused_all,fused_staged,fused_mixed --reps 9 > gpurun_out/mb_elem.log 2>&1; echo "rc=$?"; grep -v "_dx" gpurun_out/mb_elem.log
echo "== per-direction x faces"; timeout 300 python scripts/microbench_halo.py --styles direct,staged --only dx-1_dy0_dz0,dx1_dy0_dz0,dx0_dy1_dz0 --reps 9 > gpurun_out/mb_dirs.log 2>&1; cat gpurun_out/mb_dirs.log | grep -v calib
echo "== bench"; timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_n1.json; tail -3 gpurun_out/bench_n1.err
echo "== ncu he_put_all / packput / waitunpack"
timeout 200 python scripts/microbench_halo.py --styles fused_all,fused_staged --reps 1 > gpurun_out/plain_for_ncu.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:box_copy --csv --log-file gpurun_out/ncu_halo_bytes.csv python scripts/microbench_halo.py --styles fused_all,fused_staged --reps 1 > gpurun_out/ncu_halo.log 2>&1; echo "ncu rc=$?"
