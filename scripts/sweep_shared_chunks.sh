#!/bin/bash
# bench.py at N GPUs with the shared-scene call cut into 1, 2, 3, 4, 6 chunks (device-resident value only)
N=${1:-2}
for c in ${CHUNKS:-1 2 3 4 6}; do
  BEAR_CUDA_SHARED_CHUNKS=$c timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520 + c)) \
    bench.py --gpus $N --steps 10 --warmup 3 --no-extras --batch-files 0 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('chunks $c: step %.3f ms (with per-kernel events, one stream: %.3f)' % (d['ms_per_step'], d['ms_per_step_with_kernel_events']), {k: round(v, 3) for k, v in d['kernel_ms_per_step'].items()}, 'e2e %.2f' % d['e2e']['ms_per_step'])"
done
