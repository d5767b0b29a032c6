#!/bin/bash
# usage: gpu_profile_two.sh [kernel-regex] [count] [out-name]
# plain bench, then ncu --set full of one launch each of the named kernels (default: mix, decorrelate
# and bus_fdl) after a clean plain run of the same command.
set -u
K=${1:-'mix_kernel|bus_fdl_kernel|decorrelate_kernel'}; C=${2:-3}; OUT=${3:-prof_mix_bus}
mkdir -p gpurun_out
python bench.py --no-cpu-baseline ${BENCH_EXTRA:-} > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp_err.log
echo "bench_exit=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline ${BENCH_EXTRA:-}"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$K" -s $((2 * C)) -c $C -f -o gpurun_out/$OUT $CMD > gpurun_out/ncu_full.log 2>&1
echo "full_capture_exit=$?"
