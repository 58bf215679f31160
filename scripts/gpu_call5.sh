#!/bin/bash
# cheap diagnosis of the reference drivers on ONE GPU, everything under short timeouts
mkdir -p gpurun_out
export TENZING_ACK_NOTICE=1
python -c "import sys; sys.path.insert(0,'baseline'); import build_ref; print('host', build_ref.host_id())"; cat baseline/_ref/BUILD_HOST
echo "== build on this box"; (time timeout 300 python baseline/build_ref.py) > gpurun_out/build_ref.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/build_ref.log
echo "== MCTS driver, 25 s"; TZMPI_RANK=0 TZMPI_SIZE=1 TZMPI_JOB=a$$ timeout 25 baseline/_ref/build/tenzing-mcts/examples/spmv-min-time -m 20000 -i 2 -b 3 > gpurun_out/ref_mcts.out 2> gpurun_out/ref_mcts.err; echo "rc=$?"; wc -c gpurun_out/ref_mcts.err; tail -c 600 gpurun_out/ref_mcts.err | tr '\n' ' ' | cut -c1-600; echo
echo "== DFS driver via bench arm, 60 s"; timeout 150 python bench.py --impl reference --gpus 1 --ref-seconds 60 > gpurun_out/ref_n1.json 2> gpurun_out/ref_n1.err; echo "rc=$?"; cat gpurun_out/ref_n1.json; tail -3 gpurun_out/ref_n1.err
wc -c gpurun_out/ref_n1_r0.err; grep -c "benchmark" gpurun_out/ref_n1_r0.err; grep -n "benchmark .* sequences\|remove equivalent" gpurun_out/ref_n1_r0.err | head -5; head -c 400 gpurun_out/ref_n1_r0.out
truncate -s 2000000 gpurun_out/ref_n1_r0.err gpurun_out/ref_mcts.err 2>/dev/null
