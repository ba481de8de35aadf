#!/bin/bash
# GPU box, round 2, run 3: tests (default + RT_DEBUG build), variant A/B, skip-rate counters, bench, ncu captures
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > $O/r02_pytest3.log 2>&1; tail -4 $O/r02_pytest3.log
RTIOW_LIB=build/variants/debug/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest3_debug.log 2>&1; tail -3 $O/r02_pytest3_debug.log
ROUNDS=2 bash tools/ab_bench.sh > $O/r02_ab3.log 2>&1; grep "^round" $O/r02_ab3.log
RTIOW_LIB=build/variants/skipstat/librtiow_b200.so python tools/run_config.py --scene huge --spp 32 --reps 1 > $O/r02_skipstat.log 2>&1; tail -1 $O/r02_skipstat.log
python bench.py --steps 5 --warmup 3 > $O/r02_bench3.json 2> $O/r02_bench3.err; tail -3 $O/r02_bench3.err; cut -c1-250 $O/r02_bench3.json
python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/r02_plain_launch.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/r02_ncu_launch.log 2>&1
python tools/profile_render.py huge 16 > $O/r02_plain_prof.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render_tc -s 1 -c 1 -o $O/r02_tc python tools/profile_render.py huge 16 > $O/r02_ncu_tc.log 2>&1
RTIOW_LIB=build/variants/regroup/librtiow_b200.so python tools/profile_render.py huge 16 > $O/r02_plain_regroup.log 2>&1 && \
RTIOW_LIB=build/variants/regroup/librtiow_b200.so ncu --set full --clock-control none --import-source on -k regex:k_render_tc -s 1 -c 1 -o $O/r02_tc_regroup python tools/profile_render.py huge 16 > $O/r02_ncu_tc_regroup.log 2>&1
tail -2 $O/r02_plain_prof.log $O/r02_plain_regroup.log
ls -la $O | tail -20
