#!/bin/bash
# BASELINE.json configs[4] at full size: xs::count over 8 x 50 GB = 400 GB sharded across 8 B200 (torchrun, NCCL),
# a subset of the pattern lengths (every run is charged 8x).  Output: gpurun_out/sweep8.jsonl
mkdir -p gpurun_out
: > gpurun_out/sweep8.jsonl
n=${1:-8}
for m in 1 4 8 64; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + m)) \
    bench.py --gpus $n --workload count/sweep/50GB --pattern-len $m --steps 10 --warmup 3 --no-e2e --no-cpu \
    2> gpurun_out/sweep8_$m.err | tail -1 >> gpurun_out/sweep8.jsonl
  echo "m=$m rc=$?"
done
python - <<'PY'
import json
for l in open('gpurun_out/sweep8.jsonl'):
    d=json.loads(l); print(len(d['config']['pattern']), d['n_gpus'], d['value'], d['ms_per_step'], d['config']['global_results'])
PY
