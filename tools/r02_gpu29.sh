#!/bin/bash
# 4-GPU box, round 2, run 29: headline and configs[4] bench lines on 4 GPUs (final build)
N=${1:-4}
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3 "${@:3}" > $O/$2.json 2> $O/$2.err; tail -1 $O/$2.json | cut -c1-220; }
run 29761 r02_bench29_${N}gpu
run 29762 r02_bench29_${N}gpu_c5 --workload synthetic1m-1080p-256spp --no-secondary
