#!/bin/bash
# first GPU call of round 2: sanity of everything that changed on the CPU since the last GPU run,
# the experimental kernels, the access-pattern probe, baseline micro-benchmarks
mkdir -p gpurun_out
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
export TENZING_ACK_NOTICE=1
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > gpurun_out/clocks.csv &
SMI=$!
echo "== probe"; timeout 300 build/bin/thin_probe 7 > gpurun_out/probe.log 2>&1; echo "probe rc=$?"; tail -40 gpurun_out/probe.log
echo "== pytest gpu"; timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
echo "== experimental"; TENZING_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_gpu_workloads.py -q -k experimental > gpurun_out/pytest_exp.log 2>&1; echo "exp rc=$?"; tail -5 gpurun_out/pytest_exp.log
echo "== spmv microbench"; timeout 300 python scripts/microbench_spmv.py --out gpurun_out/microbench_spmv.json > gpurun_out/microbench_spmv.log 2>&1; echo "rc=$?"; cat gpurun_out/microbench_spmv.log
echo "== halo microbench base"; timeout 300 python scripts/microbench_halo.py --styles fused_all,fused_staged,fused_mixed --reps 7 > gpurun_out/mb_base.log 2>&1; echo "rc=$?"; grep -v "_dx" gpurun_out/mb_base.log
echo "== halo microbench bulk"; TZB_BULK=1 timeout 300 python scripts/microbench_halo.py --styles fused_all,fused_mixed --reps 7 > gpurun_out/mb_bulk.log 2>&1; echo "rc=$?"; grep -v "_dx" gpurun_out/mb_bulk.log
for h in 64 128; do echo "== halo microbench hint $h"; TZB_THIN_HINT=$h timeout 300 python scripts/microbench_halo.py --styles fused_all,fused_staged --reps 7 > gpurun_out/mb_hint$h.log 2>&1; echo "rc=$?"; grep -v "_dx" gpurun_out/mb_hint$h.log; done
echo "== bench"; timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; cat gpurun_out/bench_n1.json; tail -3 gpurun_out/bench_n1.err
echo "== ncu dram bytes of the probe"
timeout 120 build/bin/thin_probe 1 > gpurun_out/probe1.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/ncu_probe.csv build/bin/thin_probe 1 > gpurun_out/ncu_probe.log 2>&1; echo "ncu rc=$?"
kill $SMI
