#!/bin/bash
# GPU box, round 2, run 16: one-GPU timing of the mid-frame hand-over build against the frame-end-only build
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
for rep in 1 2; do
for v in default coopend; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  echo "== $v"; RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 | cut -c75-330
done
done > $O/r02_coop_ab4.log 2>&1
cat $O/r02_coop_ab4.log
