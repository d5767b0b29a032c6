#!/bin/bash
# ncu --set full of one launch each of mix_kernel and bus_fdl_kernel (after a clean plain run).
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'mix_kernel|bus_fdl_kernel' -s 4 -c 2 -f -o gpurun_out/prof_mix_bus $CMD > gpurun_out/ncu_full.log 2>&1
echo "full_capture_exit=$?"
