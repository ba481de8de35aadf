#!/bin/bash
# N-GPU box, round 2, run 24: suite on the final straggler build (and with an 8-entry work list), configs[4] + headline bench lines, trace
N=${1:-2}
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
RTIOW_LIB=build/variants/coopovf/librtiow_b200.so python -m pytest tests -m gpu -x -q > $O/r02_pytest24_coopovf.log 2>&1; tail -2 $O/r02_pytest24_coopovf.log
python -m pytest tests -m gpu -x -q > $O/r02_pytest24.log 2>&1; tail -2 $O/r02_pytest24.log
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 5 --warmup 3 "${@:3}" > $O/$2.json 2> $O/$2.err; tail -1 $O/$2.json | cut -c1-220; }
run 29741 r02_bench24_${N}gpu_c5 --workload synthetic1m-1080p-256spp --no-secondary
run 29742 r02_bench24_${N}gpu
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29743 tools/queue_trace.py > $O/r02_trace24_${N}gpu.log 2>&1; tail -$N $O/r02_trace24_${N}gpu.log | cut -c1-200
