#!/bin/bash
# GPU box, round 2, run 7: tiles-open-at-once (feed interleave) for the LBVH kernel in queue mode, 256 spp
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
one() { python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])"; }
for round in 1 2; do
for d in build/variants/*/; do
  n=$(basename $d)
  mq=$(RTIOW_LIB=$d/librtiow_b200.so timeout 120 python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 --queue 2>&1 | tail -1 | one)
  mp=$(RTIOW_LIB=$d/librtiow_b200.so timeout 120 python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 2>&1 | tail -1 | one)
  hq=$(RTIOW_LIB=$d/librtiow_b200.so timeout 60 python tools/run_config.py --scene huge --spp 100 --reps 2 --queue 2>&1 | tail -1 | one)
  echo "round $round $n: synthetic1m-256spp queue $mq plain $mp | huge queue $hq Msamples/s"
done; done > $O/r02_ab7.log 2>&1
cat $O/r02_ab7.log
RTIOW_LIB=build/variants/il8/librtiow_b200.so python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q 2>&1 | tail -2
