#!/bin/bash
# One node, one rank per GPU (the reference's batch script: scripts/perlmutter/spmv.sh, 4 ranks x 1
# GPU under srun with --signal=SIGABRT@10 so a partial CSV is dumped before the wall-time limit).
# Usage: scripts/run_8gpu.sh [N] [driver and args...]   default: 8 ranks of halo-mcts-min-time
set -euo pipefail
cd "$(dirname "$0")/.."
N=${1:-8}; shift || true
export TENZING_ACK_NOTICE=1
[ -x build/bin/tzrun ] || python build.py
if [ $# -eq 0 ]; then set -- ./build/bin/halo-mcts-min-time -i 300 -b 50; fi
# SIGABRT/SIGINT/SIGTERM make rank 0 dump the results collected so far (tenzing::register_handler)
exec ./build/bin/tzrun -n "$N" "$@"
