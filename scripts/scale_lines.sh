#!/bin/bash
# bench.py at N GPUs: the default line (-> gpurun_out/<tag>_bench_n<N>.json) and, with ALLREDUCE=1 in the
# environment, a second line with the stereo partial mixes all-reduced instead of reduced to rank 0
N=${1:-2}; TAG=${2:-r2}
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
run 29531 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
if [ -n "${ALLREDUCE:-}" ]; then
  BEAR_CUDA_COMM_ALLREDUCE=1 run 29532 --steps 10 --no-extras --batch-files 0 > gpurun_out/${TAG}_bench_n${N}_allreduce.json 2>> gpurun_out/${TAG}_bench_n$N.err
fi
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/${TAG}_bench_n${N}*.json")):
    try:
        d = json.load(open(f))
        print(f, "step %.3f ms value %.0f e2e %.2f ms" % (d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"]), {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d.get("batch_files", {}).get("files_per_s"))
    except Exception as e:
        print(f, "ERR", e)
PY
