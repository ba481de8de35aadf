#!/bin/bash
# GPU box: which kernel for 13..60 spheres?  variants: default (FMA-filter kernel), k_render_small up to 47, tensor kernel from 13
for ns in 4 5 6 7 8; do
  for d in build/variants/*/; do
    v=$(basename $d)
    r=$(RTIOW_LIB=$d/librtiow_b200.so timeout 120 python tools/run_config.py --scene synthetic:$ns --spp 64 --reps 2 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('%d spheres %.1f Msamples/s' % (d['n_spheres'], d['Msamples_per_s']))")
    echo "$v: $r"
  done
done
