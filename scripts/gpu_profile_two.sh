#!/bin/bash
# plain bench, then ncu --set full of one launch each of mix_kernel, decorrelate_kernel and
# bus_fdl_kernel (after a clean plain run of the same command).
set -u
mkdir -p gpurun_out
python bench.py --no-cpu-baseline > gpurun_out/bench_tmp.json 2> gpurun_out/bench_tmp_err.log
echo "bench_exit=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'mix_kernel|bus_fdl_kernel|decorrelate_kernel' -s 6 -c 3 -f -o gpurun_out/prof_mix_bus $CMD > gpurun_out/ncu_full.log 2>&1
echo "full_capture_exit=$?"
