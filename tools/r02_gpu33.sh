#!/bin/bash
# 8-GPU box, round 2, run 33: configs[4] with the final straggler-mode entry policy (4 lanes / 64 iterations) + trace
N=${1:-8}
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3 "${@:3}" > $O/$2.json 2> $O/$2.err; tail -1 $O/$2.json | cut -c1-220; }
run 29771 r02_bench33_${N}gpu_c5 --workload synthetic1m-1080p-256spp --no-secondary
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29772 tools/queue_trace.py > $O/r02_trace33_${N}gpu.log 2>&1; tail -$N $O/r02_trace33_${N}gpu.log | cut -c1-120
