#!/bin/bash
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29911 tools/queue_trace.py synthetic1m-1080p-256spp > $O/r02_trace_c5.log 2>&1; grep '^{' $O/r02_trace_c5.log | cut -c1-900
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29912 tools/queue_trace.py huge-4k-1024spp > $O/r02_trace_huge.log 2>&1; grep '^{' $O/r02_trace_huge.log | cut -c1-400
python -m pytest tests/test_multigpu_gpu.py -m gpu -x -q 2>&1 | tail -2
