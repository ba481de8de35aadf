#!/bin/bash
# GPU box: brute force (tensor filter up to 1024 spheres, FMA filter above) against the LBVH on synthetic grids of n_side^2 + 1 spheres
for ns in 8 12 16 22 27 32 45 64 90; do
  b=$(timeout 120 python tools/run_config.py --scene synthetic:$ns --spp 32 --reps 2 --brute 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('%d spheres brute %.1f' % (d['n_spheres'], d['Msamples_per_s']))")
  l=$(timeout 120 python tools/run_config.py --scene synthetic:$ns --spp 32 --reps 2 --lbvh 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('lbvh %.1f Msamples/s' % d['Msamples_per_s'])")
  echo "$b $l"
done
