#!/bin/bash
# GPU box: time the LBVH render path of every variant under build/variants (1M-sphere scene and the huge scene via --lbvh)
for d in build/variants/*/; do
  n=$(basename $d)
  echo "== $n"
  RTIOW_LIB=$d/librtiow_b200.so timeout 120 python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp ${SPP:-8} --reps 2 2>&1 | tail -1 | cut -c1-200
  RTIOW_LIB=$d/librtiow_b200.so timeout 60 python tools/run_config.py --scene huge --lbvh --spp 20 --reps 2 2>&1 | tail -1 | cut -c1-200
done
