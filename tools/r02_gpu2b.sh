#!/bin/bash
# 2-GPU box: LBVH with 32-item batches -- configs[4] on 2 GPUs (device queue / host chunks), 1-GPU rates, multi-GPU tests
N=2
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
one() { python -c "import sys,json; print('%.1f' % json.loads(sys.stdin.readline())['Msamples_per_s'])"; }
echo "1 GPU synthetic1m 256spp: plain $(python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 2>&1 | tail -1 | one) queue $(python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 --queue 2>&1 | tail -1 | one) | huge $(python tools/run_config.py --scene huge --spp 100 --reps 2 2>&1 | tail -1 | one)" | tee $O/r02_gpu2b_1gpu.log
python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q > $O/r02_pytest_2gpu_b.log 2>&1; tail -2 $O/r02_pytest_2gpu_b.log
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 3 --warmup 3 "${@:3}" > $O/$2.json 2> $O/$2.err; tail -1 $O/$2.json | cut -c1-200; }
run 29813 r02_bench_2gpu_c5_b --workload synthetic1m-1080p-256spp --no-secondary
run 29814 r02_bench_2gpu_c5_host_b --workload synthetic1m-1080p-256spp --mgpu host --no-secondary
