#!/bin/bash
# GPU box, round 2, run 10: the no-allocation test alone and after the suite; path-cost tail of the 1M-sphere frame
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r02_pytest10_noalloc_alone.log 2>&1; tail -3 $O/r02_pytest10_noalloc_alone.log
python -m pytest tests -m gpu -x -q > $O/r02_pytest10.log 2>&1; tail -3 $O/r02_pytest10.log
RTIOW_LIB=build/variants/pathstat/librtiow_b200.so timeout 600 python tools/pathstat.py 256 > $O/r02_pathstat.log 2>&1; head -50 $O/r02_pathstat.log
