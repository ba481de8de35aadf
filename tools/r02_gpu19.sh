#!/bin/bash
# GPU box, round 2, run 19: one shading copy + parked state + per-warp node counter: suite (default; work list of 64 and of 8 entries),
# timing against the two-copy build of run 17 and the build without the straggler mode
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
for v in coopmid coopovf; do
  RTIOW_LIB=build/variants/$v/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest19_$v.log 2>&1; echo $v; tail -2 $O/r02_pytest19_$v.log
done
python -m pytest tests -m gpu -x -q > $O/r02_pytest19.log 2>&1; tail -2 $O/r02_pytest19.log
for rep in 1 2; do
for v in default nopark twocopiesend nocoop; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  echo "== $v"; RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 | cut -c75-200
done
done > $O/r02_coop_ab7.log 2>&1
cat $O/r02_coop_ab7.log
RTIOW_LIB=raytracinginoneweekend_b200/librtiow_b200.so python tools/run_config.py --scene huge --width 1280 --height 720 --spp 100 --reps 2 --lbvh | cut -c75-200
RTIOW_LIB=build/variants/nocoop/librtiow_b200.so python tools/run_config.py --scene huge --width 1280 --height 720 --spp 100 --reps 2 --lbvh | cut -c75-200
