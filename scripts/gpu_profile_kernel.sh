#!/bin/bash
# usage: gpu_profile_kernel.sh <kernel-regex> <out-name> [bench args...]
# GPU tests, plain bench, then one ncu --set full capture of the named kernel (B200_PROFILING.md rules).
set -u
K=$1; OUT=$2; shift 2
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest_exit=$?" >> gpurun_out/pytest_gpu.log
tail -n 2 gpurun_out/pytest_gpu.log
python bench.py --no-cpu-baseline "$@" > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp_err.log
echo "bench_exit=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline $*"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$K -s 6 -c 1 -f -o gpurun_out/$OUT $CMD > gpurun_out/ncu_full.log 2>&1
echo "full_capture_exit=$?"
