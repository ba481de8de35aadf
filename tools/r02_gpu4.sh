#!/bin/bash
# GPU box, round 2, run 4: tests (default + RT_DEBUG build), variant A/B incl. the round-1 library, bench
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > $O/r02_pytest4.log 2>&1; tail -4 $O/r02_pytest4.log
RTIOW_LIB=build/variants/debug/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest4_debug.log 2>&1; tail -3 $O/r02_pytest4_debug.log
ROUNDS=2 bash tools/ab_bench.sh > $O/r02_ab4.log 2>&1; grep "^round" $O/r02_ab4.log
python bench.py --steps 5 --warmup 3 > $O/r02_bench4.json 2> $O/r02_bench4.err; tail -3 $O/r02_bench4.err; cut -c1-250 $O/r02_bench4.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/r02_bench4_ref.json 2> $O/r02_bench4_ref.err; cut -c1-200 $O/r02_bench4_ref.json
