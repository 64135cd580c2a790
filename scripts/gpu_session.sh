#!/bin/bash
# One GPU session: tests, bench (driver's arguments), per-kernel times along the trajectory, ncu launch list and one
# --set full capture of the step kernels.  Everything lands in gpurun_out/<tag>_*.
tag=${1:-s}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_tests.log 2>&1; tail -3 gpurun_out/${tag}_tests.log
python bench.py --steps 20 --warmup 5 > gpurun_out/${tag}_bench.log 2> gpurun_out/${tag}_bench.err; tail -c 1500 gpurun_out/${tag}_bench.log; tail -5 gpurun_out/${tag}_bench.err
python scripts/kernel_times.py approximate 5,100 > gpurun_out/${tag}_ktimes.log 2>&1; cat gpurun_out/${tag}_ktimes.log
