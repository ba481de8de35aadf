#!/bin/bash
# GPU box, round 2, run 31: final build with the spp gate of the LBVH tile-queue path: suite, smoke, bench, LBVH timing at 32 / 256 spp
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > $O/r02_pytest31.log 2>&1; tail -2 $O/r02_pytest31.log
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -1
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke31.log 2>&1; tail -1 $O/r02_smoke31.log
python bench.py --steps 5 --warmup 3 > $O/r02_bench31.json 2> $O/r02_bench31.err; tail -2 $O/r02_bench31.err; cut -c1-200 $O/r02_bench31.json
for spp in 32 256; do python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp $spp --reps 2 | cut -c75-130; done
python tools/run_config.py --scene huge --width 1280 --height 720 --spp 100 --reps 2 --lbvh | cut -c75-130
