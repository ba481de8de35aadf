#!/bin/bash
# tools/build_variant.sh <name> <extra nvcc -D flags...>: builds an experimental variant of the backend into
# build/variants/<name>/librtiow_b200.so (select it with RTIOW_LIB=<path>); not part of the product build.
set -e
HERE="$(cd "$(dirname "$0")/.." && pwd)"
name="$1"; shift
out="$HERE/build/variants/$name"; mkdir -p "$out"
SRC="$HERE/raytracinginoneweekend_b200/csrc"
COMMON="-std=c++17 -O3 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=hidden $*"
nvcc $COMMON --fmad=false -c $SRC/rt_kernels_exact.cu -o $out/exact.o &
nvcc $COMMON --fmad=true -c $SRC/rt_kernels_fast.cu -o $out/fast.o -Xptxas -v 2> $out/fast.ptxas.log &
nvcc $COMMON --fmad=${BVH_FMAD:-false} -c $SRC/rt_kernels_bvh.cu -o $out/bvh.o -Xptxas -v 2> $out/bvh.ptxas.log &
nvcc $COMMON -c $SRC/rt_common.cu -o $out/common.o &
nvcc $COMMON -c $SRC/rt_lbvh.cu -o $out/lbvh.o &
nvcc $COMMON -c $SRC/rt_api.cu -o $out/api.o &
nvcc $COMMON -c $SRC/rt_multi.cu -o $out/multi.o &
nvcc $COMMON -c $SRC/rt_static.cu -o $out/static.o &
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/librtiow_b200.so $out/exact.o $out/fast.o $out/bvh.o $out/common.o $out/lbvh.o $out/api.o $out/multi.o $out/static.o -cudart shared -lrt -ldl
grep -A3 k_render_tc $out/fast.ptxas.log | grep -E "registers|spill" | tr '\n' ' '; echo " -> $out/librtiow_b200.so"
