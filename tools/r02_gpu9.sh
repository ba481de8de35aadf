#!/bin/bash
# GPU box, round 2, run 9: the no-allocation entry point alone and after the suite, smoke, bench (both arms), memcheck of the new test
O=$GRAFT_REPO_ROOT/gpurun_out
cd "$(dirname "$0")/.."
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r02_pytest9_noalloc_alone.log 2>&1; tail -4 $O/r02_pytest9_noalloc_alone.log
python -m pytest tests -m gpu -x -q > $O/r02_pytest9.log 2>&1; tail -4 $O/r02_pytest9.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r02_smoke9.log 2>&1; tail -2 $O/r02_smoke9.log
python bench.py --steps 5 --warmup 3 > $O/r02_bench9.json 2> $O/r02_bench9.err; tail -3 $O/r02_bench9.err; cut -c1-300 $O/r02_bench9.json
python bench.py --impl reference --steps 2 --warmup 1 > $O/r02_bench9_ref.json 2> $O/r02_bench9_ref.err; cut -c1-300 $O/r02_bench9_ref.json
timeout 900 compute-sanitizer --tool memcheck python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r02_memcheck9_noalloc.log 2>&1; tail -6 $O/r02_memcheck9_noalloc.log
