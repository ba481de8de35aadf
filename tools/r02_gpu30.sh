#!/bin/bash
# GPU box, round 2, run 30 (final): suite, smoke, bench (both arms), launch list of the bench, ncu captures of the LBVH and tensor kernels
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > $O/r02_pytest30.log 2>&1; tail -2 $O/r02_pytest30.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke30.log 2>&1; tail -1 $O/r02_smoke30.log
python bench.py --impl reference --steps 2 --warmup 1 > $O/r02_bench30_ref.json 2> $O/r02_bench30_ref.err; cut -c1-200 $O/r02_bench30_ref.json
python bench.py --steps 5 --warmup 3 > $O/r02_bench30.json 2> $O/r02_bench30.err; tail -2 $O/r02_bench30.err; cut -c1-300 $O/r02_bench30.json
python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/r02_plain_launch30.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02_launches_bench30.csv python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline > $O/r02_ncu_launch30.log 2>&1
python tools/profile_render.py synthetic1m 16 > $O/r02_plain_prof_bvh30.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_render_bvh -s 1 -c 1 -o $O/r02_bvh30 python tools/profile_render.py synthetic1m 16 > $O/r02_ncu_bvh30.log 2>&1
tail -n 2 $O/r02_plain_prof_bvh30.log; ls -la $O/*.ncu-rep
