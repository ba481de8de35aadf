#!/bin/bash
# GPU box, round 2, run 22: single-copy straggler build, LBVH frames on one GPU through the tile queue (default) or as one sample-major tile (plainq)
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
for v in default plainq nocoop; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  echo "== $v"; RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 | cut -c75-130
  RTIOW_LIB=$lib python tools/run_config.py --scene huge --width 1280 --height 720 --spp 100 --reps 2 --lbvh | cut -c75-130
  RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 32 --reps 2 | cut -c75-130
done > $O/r02_coop_ab9.log 2>&1
cat $O/r02_coop_ab9.log
