#!/bin/bash
# usage: build_variant.sh <name> <git-rev or WORK> [extra nvcc flags]   -> scratch/variants/lib<name>.so (gemm.cu differs, rest = current objs)
set -e
name=$1; rev=$2; shift 2
root=/root/repo; out=$root/scratch/variants; tmp=/tmp/t/var_$name
mkdir -p $out $tmp/guda_b200/csrc $tmp/include
if [ "$rev" = WORK ]; then cp $root/guda_b200/csrc/*.cu* $tmp/guda_b200/csrc/; cp $root/include/*.h $tmp/include/
else for f in gemm.cu gemm_simt.cuh common.cuh; do git -C $root show $rev:guda_b200/csrc/$f > $tmp/guda_b200/csrc/$f; done; git -C $root show $rev:include/guda_b200.h > $tmp/include/guda_b200.h; fi
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC "$@" -c $tmp/guda_b200/csrc/gemm.cu -o $tmp/gemm.o
objs=$(ls $root/guda_b200/lib/obj/*.o | grep -v "/gemm.o" | grep -v "${SKIP_OBJ:-NONE}")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out/lib$name.so $tmp/gemm.o $objs -lcuda -ldl
echo built $out/lib$name.so
