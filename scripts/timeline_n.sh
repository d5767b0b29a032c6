#!/bin/bash
# Development aid: timeline of a few pipelined steps at N GPUs (BEAR_CUDA_TIMELINE), with and without a
# high-priority convolution stream.  Output: gpurun_out/<tag>_tl<variant>.rank<r>, bench lines next to them.
N=${1:-8}; TAG=${2:-tl}
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N --steps 6 --warmup 3 --no-extras --batch-files 0; }
for v in ${VARIANTS:-1}; do
  rm -f gpurun_out/${TAG}_tl$v.rank*
  BEAR_CUDA_CONV_PRIORITY=$v BEAR_CUDA_TIMELINE=gpurun_out/${TAG}_tl$v run 2954$v > gpurun_out/${TAG}_tl$v.json 2> gpurun_out/${TAG}_tl$v.err
  if [ -n "${PLAIN:-}" ]; then BEAR_CUDA_CONV_PRIORITY=$v run 2955$v > gpurun_out/${TAG}_prio$v.json 2>> gpurun_out/${TAG}_tl$v.err; fi
  python - <<PY
import json
for f in ("gpurun_out/${TAG}_tl$v.json", "gpurun_out/${TAG}_prio$v.json"):
    try:
        d = json.load(open(f)); print(f, "step %.3f ms (unpipelined %.3f)" % (d["ms_per_step"], d["exchange"]["ms_per_step_unpipelined"]))
    except Exception as e:
        print(f, "ERR", e)
PY
done
# keep only ranks 0 and the last (the files are a few hundred lines each)
ls gpurun_out/${TAG}_tl*.rank* | grep -v "rank0$\|rank$((N-1))$" | xargs -r rm -f
