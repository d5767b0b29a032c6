#!/bin/bash
# times the device-resident headline step with differently tuned builds of the library (bear_b200/variants/*.so)
for lib in bear_b200/variants/*.so; do
  for splits in 0; do
  BEAR_CUDA_LIB=$PWD/$lib python bench.py --object-splits $splits --steps 10 --warmup 3 --no-cpu-baseline --no-extras --batch-files 0 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$lib splits=$splits', 'step %.3f ms' % d['ms_per_step'], {k: round(v, 3) for k, v in d['kernel_ms_per_step'].items()})"
  done
done
