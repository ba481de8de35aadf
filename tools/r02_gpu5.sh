#!/bin/bash
# GPU box, round 2, run 5: bisect of the tensor-kernel slowdown between runs 3 and 4 (variants), multi-GPU tests file
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q > $O/r02_pytest5.log 2>&1; tail -3 $O/r02_pytest5.log
ROUNDS=2 bash tools/ab_bench.sh > $O/r02_ab5.log 2>&1; grep "^round" $O/r02_ab5.log
