#!/bin/bash
# GPU box, round 2, run 36: suite with the straggler-mode test (RT_FLAG_TEST_STRAGGLERS), 1M-sphere timing with the runtime entry parameters
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests -m gpu -x -q > $O/r02_pytest36.log 2>&1; tail -3 $O/r02_pytest36.log
python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 256 --reps 2 | cut -c75-130
python tools/run_config.py --scene synthetic1m --width 1920 --height 1080 --spp 32 --reps 2 | cut -c75-130
