#!/bin/bash
# GPU box, round 2, run 11: straggler mode of k_render_bvh (bvh_coop_query) -- correctness with the mode forced on every frame end
# (coopall) and with a work list of 8 entries (coopovf: every query overflows into the sequential fallback), then timing
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
for v in coopall coopovf; do
  RTIOW_LIB=build/variants/$v/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest11_$v.log 2>&1; echo $v; tail -3 $O/r02_pytest11_$v.log
done
python -m pytest tests -m gpu -x -q > $O/r02_pytest11.log 2>&1; tail -3 $O/r02_pytest11.log
for rep in 1 2; do
for v in default nocoop coopall; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  echo "== $v plain"; RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 3 | cut -c1-140
  echo "== $v queue"; RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 3 --queue | cut -c1-140
done
done > $O/r02_coop_ab.log 2>&1
cat $O/r02_coop_ab.log
