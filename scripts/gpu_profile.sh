#!/bin/bash
# Run on the GPU box (via gpurun): GPU tests, then the ncu launch list and one full capture of the
# roofline kernel, each only after the same command exited 0 without ncu (B200_PROFILING.md).
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
echo "pytest_exit=$?" >> gpurun_out/pytest_gpu.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch_list_exit=$?"
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:cmac_kernel -s 3 -c 2 -f -o gpurun_out/prof_cmac $CMD > gpurun_out/ncu_full.log 2>&1
echo "full_capture_exit=$?"
tail -n 3 gpurun_out/pytest_gpu.log
