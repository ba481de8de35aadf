for ns in 12 16 19 22 24 27; do
  b=$(timeout 120 python tools/run_config.py --scene synthetic:$ns --spp 32 --reps 2 --brute 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('%d spheres brute %.1f' % (d['n_spheres'], d['Msamples_per_s']))")
  l=$(timeout 120 python tools/run_config.py --scene synthetic:$ns --spp 32 --reps 2 --lbvh 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('lbvh %.1f Msamples/s' % d['Msamples_per_s'])")
  echo "$b $l"
done
b=$(timeout 120 python tools/run_config.py --scene huge --spp 64 --reps 2 2>&1 | tail -1 | cut -c1-120); echo "huge default: $b"
l=$(timeout 120 python tools/run_config.py --scene huge --spp 64 --reps 2 --lbvh 2>&1 | tail -1 | cut -c1-120); echo "huge lbvh: $l"
