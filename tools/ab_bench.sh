#!/bin/bash
# GPU box: A/B timing of the variants under build/variants on the three scenes, interleaved (ROUNDS rounds) so that
# clock/thermal drift hits all variants alike
for round in $(seq 1 ${ROUNDS:-2}); do
  for d in build/variants/*/; do
    n=$(basename $d)
    case $n in r01*) export RTIOW_LIB_OLD_ABI=1;; *) unset RTIOW_LIB_OLD_ABI;; esac   # libraries built from the round-1 tree lack the newer entry points
    h=$(RTIOW_LIB=$d/librtiow_b200.so timeout 60 python tools/run_config.py --scene huge --spp 100 --reps 2 2>&1 | tail -1 | python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])")
    s=$(RTIOW_LIB=$d/librtiow_b200.so timeout 60 python tools/run_config.py --scene simple --spp 64 --reps 2 2>&1 | tail -1 | python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])")
    m=$(RTIOW_LIB=$d/librtiow_b200.so timeout 90 python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 8 --reps 2 2>&1 | tail -1 | python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])")
    x=$(RTIOW_LIB=$d/librtiow_b200.so timeout 60 python tools/run_config.py --scene huge --spp 100 --reps 2 --exact 2>&1 | tail -1 | python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])")
    hq=$(RTIOW_LIB=$d/librtiow_b200.so timeout 60 python tools/run_config.py --scene huge --spp 100 --reps 2 --queue 2>&1 | tail -1 | python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])")
    mq=$(RTIOW_LIB=$d/librtiow_b200.so timeout 90 python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 8 --reps 2 --queue 2>&1 | tail -1 | python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])")
    echo "round $round $n: huge $h simple $s synthetic1m $m huge-parity $x | queue mode: huge $hq synthetic1m $mq Msamples/s"
  done
done
