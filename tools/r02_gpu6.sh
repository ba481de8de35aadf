#!/bin/bash
# GPU box, round 2, run 6: tests (default + RT_DEBUG build), variant A/B, bench (both arms), ncu captures
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > $O/r02_pytest6.log 2>&1; tail -4 $O/r02_pytest6.log
RTIOW_LIB=build/variants/debug/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest6_debug.log 2>&1; tail -3 $O/r02_pytest6_debug.log
ROUNDS=2 bash tools/ab_bench.sh > $O/r02_ab6.log 2>&1; grep "^round" $O/r02_ab6.log
python bench.py --steps 5 --warmup 3 > $O/r02_bench6.json 2> $O/r02_bench6.err; tail -3 $O/r02_bench6.err; cut -c1-250 $O/r02_bench6.json
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke6.log 2>&1; tail -2 $O/r02_smoke6.log
python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/r02_plain_launch.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/r02_ncu_launch.log 2>&1
python tools/profile_render.py huge 64 > $O/r02_plain_prof.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render_tc -s 1 -c 1 -o $O/r02_tc python tools/profile_render.py huge 64 > $O/r02_ncu_tc.log 2>&1
RTIOW_LIB=build/variants/regroup/librtiow_b200.so python tools/profile_render.py huge 64 > $O/r02_plain_regroup.log 2>&1 && \
RTIOW_LIB=build/variants/regroup/librtiow_b200.so ncu --set full --clock-control none --import-source on -k regex:k_render_tc -s 1 -c 1 -o $O/r02_tc_regroup python tools/profile_render.py huge 64 > $O/r02_ncu_tc_regroup.log 2>&1
ncu --set full --clock-control none -k regex:k_tf32_probe -c 1 -o $O/r02_tf32probe python -c "
import sys; sys.path.insert(0,'.')
from raytracinginoneweekend_b200.backend import Backend
b=Backend(0); print(b.tf32_peak_probe())" > $O/r02_ncu_probe.log 2>&1
tail -n 2 $O/r02_plain_prof.log $O/r02_plain_regroup.log
