#!/bin/bash
# usage: scripts/build_variant.sh NAME file.cu "-DFOO=1 ..." [file2.cu "flags2"]   -> _variants/NAME.so
set -e
cd "$(dirname "$0")/../agents.jl_b200/csrc"
name=$1; shift
out=../../_variants/$name
mkdir -p $out
objs=""
declare -A over
while [ $# -gt 0 ]; do
  f=$1; fl=$2; shift 2
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo --fmad=false -Xcompiler -fPIC -Xcudafe --diag_suppress=177 $fl -c $f -o $out/${f%.cu}.o
  over[${f%.cu}.o]=1
done
for o in ab_core.o ab_resort.o ab_flock.o ab_query.o ab_pairforce.o ab_grid.o ab_comm.o ab_collect.o ab_collide.o ab_field.o; do
  if [ "${over[$o]}" = "1" ]; then objs="$objs $out/$o"; else objs="$objs $o"; fi
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../_variants/$name.so $objs -ldl
echo built _variants/$name.so
