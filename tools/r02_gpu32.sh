#!/bin/bash
# GPU box, round 2, run 32: when to enter the straggler mode -- at most 32 / 8 / 4 / 2 lanes still traversing (after 1 / 16 / 64 / 256 loop
# iterations) -- on frames whose end phase weighs differently: 1M spheres at 32 and 256 spp, 488 spheres through the LBVH
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
for v in default c8 c4 c2 nocoop; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  echo "== $v"
  for spp in 32 256; do RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp $spp --reps 2 | cut -c75-130; done
  RTIOW_LIB=$lib python tools/run_config.py --scene huge --width 1280 --height 720 --spp 100 --reps 2 --lbvh | cut -c75-130
done > $O/r02_coop_ab11.log 2>&1
cat $O/r02_coop_ab11.log
