#!/bin/bash
# 2-GPU box, round 2, run 35 (final build): multi-GPU tests, configs[4] bench line
N=2
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q > $O/r02_pytest35_2gpu.log 2>&1; tail -2 $O/r02_pytest35_2gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29781 bench.py --gpus $N --steps 5 --warmup 3 --workload synthetic1m-1080p-256spp --no-secondary > $O/r02_bench35_2gpu_c5.json 2> $O/r02_bench35_2gpu_c5.err; tail -1 $O/r02_bench35_2gpu_c5.json | cut -c1-220
