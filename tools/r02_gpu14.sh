#!/bin/bash
# GPU box, round 2, run 14: straggler mode with the mid-frame hand-over -- correctness (default; coopmid: every query older than one
# shading pass goes to the whole warp; coopovf: the same with an 8-entry work list, so every such query overflows), then timing
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
for v in coopmid coopovf; do
  RTIOW_LIB=build/variants/$v/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest14_$v.log 2>&1; echo $v; tail -3 $O/r02_pytest14_$v.log
done
python -m pytest tests -m gpu -x -q > $O/r02_pytest14.log 2>&1; tail -3 $O/r02_pytest14.log
for rep in 1 2; do
for v in default coopend nocoop; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  echo "== $v"; RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 3 | cut -c1-140
done
done > $O/r02_coop_ab2.log 2>&1
cat $O/r02_coop_ab2.log
