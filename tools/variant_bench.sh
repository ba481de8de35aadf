#!/bin/bash
# GPU box: time every variant under build/variants with the huge scene (tools/run_config.py) and the filter self-check
for d in build/variants/*/; do
  n=$(basename $d)
  echo "== $n"
  RTIOW_LIB=$d/librtiow_b200.so timeout 40 python tools/run_config.py --scene huge --spp ${SPP:-100} --reps 2 2>&1 | tail -1 | cut -c1-120
  RTIOW_LIB=$d/librtiow_b200.so timeout 40 python tools/tc_check.py 2>&1 | grep filter_check | head -1
done
