#!/bin/bash
# GPU box, round 2, run 37: bench.py after the last edits -- the LBVH roofline branch and the default line
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python bench.py --workload synthetic1m-1080p-256spp --no-secondary --no-cpu-baseline --steps 2 --warmup 3 > $O/r02_bench37_c5.json 2> $O/r02_bench37_c5.err; echo "rc=$?"; tail -2 $O/r02_bench37_c5.err
python - <<'P'
import json,os
d=json.loads(open(os.environ['GRAFT_REPO_ROOT']+'/gpurun_out/r02_bench37_c5.json').read().strip().splitlines()[-1])
print(d['value'], json.dumps(d['roofline'])[:900])
P
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/r02_bench37.json 2> $O/r02_bench37.err; echo "rc=$?"; cut -c1-160 $O/r02_bench37.json
