#!/bin/bash
# Batch job for a Slurm cluster with B200 nodes: one node, one rank per GPU, MCTS over the
# distributed SpMV (the reference's scripts/perlmutter/spmv.sh: 4 ranks x 1 GPU).
# SIGABRT ten seconds before the wall-time limit makes rank 0 dump the results collected so far
# (tenzing::register_handler), so a job that runs out of time still leaves a CSV.
#SBATCH --job-name=tenzing-spmv
#SBATCH --nodes=1
#SBATCH --ntasks-per-node=1
#SBATCH --gpus-per-node=8
#SBATCH --cpus-per-task=32
#SBATCH --time=00:30:00
#SBATCH --signal=B:SIGABRT@10
#SBATCH --output=tenzing-spmv-%j.out
set -euo pipefail
cd "${SLURM_SUBMIT_DIR:-$(dirname "$0")/../..}"
source load-env.sh
[ -x build/bin/tzrun ] || python build.py
NGPU=${SLURM_GPUS_ON_NODE:-8}
OUT=spmv-min-time-${SLURM_JOB_ID:-local}.csv
# tzrun forks one rank per GPU on this node (TCP rendezvous on 127.0.0.1) and forwards signals
./build/bin/tzrun -n "$NGPU" ./build/bin/spmv-min-time -i 300 -b 50 > "$OUT" &
pid=$!
trap 'kill -ABRT $pid' ABRT TERM
wait $pid || true
wait $pid || true
echo "results in $OUT"
