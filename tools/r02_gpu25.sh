#!/bin/bash
# GPU box, round 2, run 25: suite on the final build (default and 8-entry work list), the no-allocation test alone
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
RTIOW_LIB=build/variants/coopovf/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest25_coopovf.log 2>&1; tail -2 $O/r02_pytest25_coopovf.log
python -m pytest tests -m gpu -x -q > $O/r02_pytest25.log 2>&1; tail -2 $O/r02_pytest25.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -1
