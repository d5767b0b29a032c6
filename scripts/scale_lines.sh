#!/bin/bash
# bench.py at N GPUs: the default line (-> gpurun_out/<tag>_bench_n<N>.json); with PYTEST=1 in the
# environment the multi-GPU parity test runs first (-> gpurun_out/<tag>_pytest_n<N>.log)
N=${1:-2}; TAG=${2:-r2}
run() { timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
if [ -n "${PYTEST:-}" ]; then
  timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/${TAG}_pytest_n$N.log 2>&1; tail -15 gpurun_out/${TAG}_pytest_n$N.log
fi
run 29531 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
python - <<PY
import json, glob
for f in sorted(glob.glob("gpurun_out/${TAG}_bench_n${N}*.json")):
    try:
        d = json.load(open(f))
        print(f, "step %.3f ms value %.0f e2e %.2f ms" % (d["ms_per_step"], d["value"], d["e2e"]["ms_per_step"]), {k: round(v, 3) for k, v in d["kernel_ms_per_step"].items()}, d.get("exchange", {}).get("ms_per_step_unpipelined"), d.get("batch_files", {}).get("files_per_s"))
    except Exception as e:
        print(f, "ERR", e)
PY
