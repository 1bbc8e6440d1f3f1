#!/bin/bash
# BASELINE.json configs[4] on one GPU: xs::count over a 50 GB shard, pattern lengths 1..64 (bench.py --workload count/sweep/50GB)
mkdir -p gpurun_out
: > gpurun_out/sweep.jsonl
for m in 1 2 3 4 8 16 32 64; do
  timeout 600 python bench.py --workload count/sweep/50GB --pattern-len $m --steps 10 --warmup 3 --no-e2e --no-cpu 2>gpurun_out/sweep_$m.err | tail -1 >> gpurun_out/sweep.jsonl
  echo "m=$m rc=$?"
done
python - <<'PY'
import json
for l in open('gpurun_out/sweep.jsonl'):
    d=json.loads(l); print(d['config']['pattern'][:20].ljust(20), len(d['config']['pattern']), d['value'], d['ms_per_step'], d['roofline']['frac'], d['config']['global_results'])
PY
