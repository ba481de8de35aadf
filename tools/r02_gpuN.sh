#!/bin/bash
# GPU box with N GPUs, round 2: multi-GPU tests + bench lines (device-side queue and the round-1 host path) + the C++ demo
# usage: r02_gpuN.sh N
N=${1:-2}
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
nvidia-smi topo -m > $O/r02_topo_${N}gpu.txt 2>&1
python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q > $O/r02_pytest_${N}gpu.log 2>&1; tail -4 $O/r02_pytest_${N}gpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 3 --warmup 3 > $O/r02_bench_${N}gpu.json 2> $O/r02_bench_${N}gpu.err; tail -3 $O/r02_bench_${N}gpu.err; tail -1 $O/r02_bench_${N}gpu.json | cut -c1-300
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus $N --steps 3 --warmup 3 --mgpu host --no-secondary > $O/r02_bench_${N}gpu_host.json 2> $O/r02_bench_${N}gpu_host.err; tail -1 $O/r02_bench_${N}gpu_host.json | cut -c1-200
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus $N --steps 3 --warmup 3 --workload synthetic1m-1080p-256spp --no-secondary > $O/r02_bench_${N}gpu_c5.json 2> $O/r02_bench_${N}gpu_c5.err; tail -1 $O/r02_bench_${N}gpu_c5.json | cut -c1-200
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29614 bench.py --gpus $N --steps 3 --warmup 3 --workload synthetic1m-1080p-256spp --mgpu host --no-secondary > $O/r02_bench_${N}gpu_c5_host.json 2> $O/r02_bench_${N}gpu_c5_host.err; tail -1 $O/r02_bench_${N}gpu_c5_host.json | cut -c1-200
( time raytracinginoneweekend_b200/rtiow_demo --scene huge --width 3840 --height 2160 --spp 256 --depth 50 --gpus $N --out /tmp/demoN.ppm ) > $O/r02_demo_${N}gpu.log 2>&1
( time raytracinginoneweekend_b200/rtiow_demo --scene huge --width 3840 --height 2160 --spp 256 --depth 50 --out /tmp/demo1.ppm ) >> $O/r02_demo_${N}gpu.log 2>&1
sha256sum /tmp/demoN.ppm /tmp/demo1.ppm >> $O/r02_demo_${N}gpu.log; tail -12 $O/r02_demo_${N}gpu.log
