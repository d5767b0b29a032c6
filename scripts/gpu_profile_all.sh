#!/bin/bash
# Run on the GPU box (via gpurun).  Evidence for profiles/: GPU tests, the default bench line, the
# reference arm, then the ncu launch list and one full capture of one launch of every kernel of the
# step -- each ncu pass only after the same command exited 0 without ncu (B200_PROFILING.md).
# usage: gpu_profile_all.sh <tag>      (files land in gpurun_out/<tag>_*)
set -u
TAG=${1:-final}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1
echo "pytest_exit=$?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -n 2 gpurun_out/${TAG}_pytest_gpu.log
python bench.py > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_err.log
echo "bench_exit=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference_n1.json 2>> gpurun_out/${TAG}_bench_err.log
echo "reference_exit=$?"
# the headline step alone (1024 objects, period 256): 11 kernels per step
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --batch-files 0"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch_list_exit=$?"
K='validate_views_kernel|mix_prep_kernel|mix_order_kernel|mix_kernel|mix_dense_kernel|decorrelate_kernel|bus_fdl_kernel|view_group_kernel|cmac_kernel|ifft_out_kernel|carry_kernel'
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$K" -s 22 -c 11 -f -o gpurun_out/${TAG}_step $CMD > gpurun_out/ncu_full.log 2>&1
echo "full_capture_exit=$?"
