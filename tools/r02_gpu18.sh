#!/bin/bash
# N-GPU box, round 2, run 18: straggler mode with ONE copy of the shading (coop call inside the kernel's loop, state through memory)
N=${1:-2}
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
for v in coopmid coopovf; do
  RTIOW_LIB=build/variants/$v/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest18_$v.log 2>&1; echo $v; tail -2 $O/r02_pytest18_$v.log
done
python -m pytest tests -m gpu -x -q > $O/r02_pytest18.log 2>&1; tail -2 $O/r02_pytest18.log
for v in default coopend twocopiesend nocoop; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  echo "== $v"; RTIOW_LIB=$lib python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 3 | cut -c75-200
done > $O/r02_coop_ab6.log 2>&1
cat $O/r02_coop_ab6.log
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3 "${@:3}" > $O/$2.json 2> $O/$2.err; tail -1 $O/$2.json | cut -c1-220; }
p=29730
for v in default coopend twocopiesend; do
  lib=build/variants/$v/librtiow_b200.so; [ $v = default ] && lib=raytracinginoneweekend_b200/librtiow_b200.so
  p=$((p+1)); echo "== $v"; RTIOW_LIB=$lib run $p r02_bench18_${N}gpu_c5_$v --workload synthetic1m-1080p-256spp --no-secondary
done
