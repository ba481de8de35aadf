#!/bin/bash
# 8-GPU box, round 2: bench lines (device-side queue: headline + configs[4] + parity; round-1 host path for comparison) + C++ demo
N=${1:-8}
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi topo -m > $O/r02_topo_${N}gpu.txt 2>&1
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3 "${@:3}" > $O/$2.json 2> $O/$2.err; tail -1 $O/$2.json | cut -c1-220; }
run 29711 r02_bench_${N}gpu
run 29712 r02_bench_${N}gpu_host --mgpu host --no-secondary
run 29713 r02_bench_${N}gpu_c5 --workload synthetic1m-1080p-256spp --no-secondary
run 29714 r02_bench_${N}gpu_c5_host --workload synthetic1m-1080p-256spp --mgpu host --no-secondary
( time raytracinginoneweekend_b200/rtiow_demo --scene huge --width 3840 --height 2160 --spp 1024 --depth 50 --gpus $N --out /tmp/demoN.ppm ) > $O/r02_demo_${N}gpu.log 2>&1
( time RT_MULTI_NO_NCCL=1 raytracinginoneweekend_b200/rtiow_demo --scene huge --width 3840 --height 2160 --spp 1024 --depth 50 --gpus $N --out /tmp/demoN2.ppm ) >> $O/r02_demo_${N}gpu.log 2>&1
sha256sum /tmp/demoN.ppm /tmp/demoN2.ppm >> $O/r02_demo_${N}gpu.log; tail -14 $O/r02_demo_${N}gpu.log
