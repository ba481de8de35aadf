#!/bin/bash
# GPU box, round 2, run 23: throughput LBVH kernel in its own unit without FMA contraction (rtb) -- suite (default; 8-entry work list),
# the overlap test ten times, timing against the same unit WITH contraction (bvhfma) and the build without the straggler mode
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
RTIOW_LIB=build/variants/coopovf/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest23_coopovf.log 2>&1; tail -2 $O/r02_pytest23_coopovf.log
python -m pytest tests -m gpu -x -q > $O/r02_pytest23.log 2>&1; tail -2 $O/r02_pytest23.log
for i in 1 2 3 4 5 6 7 8 9 10; do python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "overlapped_chunks" 2>&1 | tail -1; done
for v in default bvhfma nocoop; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  echo "== $v"; RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 | cut -c75-130
  RTIOW_LIB=$lib python tools/run_config.py --scene huge --width 1280 --height 720 --spp 100 --reps 2 --lbvh | cut -c75-130
  RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 32 --reps 2 | cut -c75-130
done > $O/r02_coop_ab10.log 2>&1
cat $O/r02_coop_ab10.log
