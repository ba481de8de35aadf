#!/bin/bash
# GPU box, round 2, run 20: is the two-copy build (shading also in a straggler loop of its own) deterministic?  suite x3, overlap test x10
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
export RTIOW_LIB=build/variants/twocopiesend/librtiow_b200.so
for i in 1 2 3; do python -m pytest tests -m gpu -x -q 2>&1 | tail -1; done
for i in 1 2 3 4 5 6 7 8 9 10; do python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "overlapped_chunks" 2>&1 | tail -1; done
