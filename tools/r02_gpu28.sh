#!/bin/bash
# N-GPU box, round 2, run 28: configs[4] with the cooperative query taking two nodes per lane and round (default) against one (wide1)
N=${1:-8}
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3 "${@:3}" > $O/$2.json 2> $O/$2.err; tail -1 $O/$2.json | cut -c1-220; }
run 29751 r02_bench28_${N}gpu_c5 --workload synthetic1m-1080p-256spp --no-secondary
RTIOW_LIB=build/variants/wide1/librtiow_b200.so run 29752 r02_bench28_${N}gpu_c5_wide1 --workload synthetic1m-1080p-256spp --no-secondary
