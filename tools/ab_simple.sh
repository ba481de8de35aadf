#!/bin/bash
# GPU box: A/B timing of the variants under build/variants on the 5-sphere scene (interleaved rounds)
for round in $(seq 1 ${ROUNDS:-2}); do
  for d in build/variants/*/; do
    n=$(basename $d)
    s=$(RTIOW_LIB=$d/librtiow_b200.so timeout 60 python tools/run_config.py --scene simple --spp ${SPP:-64} --reps 3 2>&1 | tail -1 | python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])")
    echo "round $round $n: simple $s Msamples/s"
  done
done
