#!/bin/bash
# per-kernel times of the base library and of the variants in _variants/ (points: step 5 and step 100)
mkdir -p gpurun_out
tag=${1:-t}
run() { echo "== $1"; shift; env "$@" python scripts/kernel_times.py approximate 5,100 2>&1 | tail -2; }
{
run base X=1
for v in _variants/*.so; do run $(basename $v .so) AGENTS_B200_LIB=$PWD/$v; done
for t in 50 200 400; do run rs_target$t AB_RS_TARGET=$t; done
} > gpurun_out/${tag}_tune.log 2>&1
cat gpurun_out/${tag}_tune.log
