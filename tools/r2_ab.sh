#!/bin/bash
# same-box A/B of two builds of libxsg.so: tools/ab/libxsg_base.so (XSG_LIB) against the in-tree one.
# usage: tools/r2_ab.sh [bench.py args]      (results: gpurun_out/ab_*.json)
mkdir -p gpurun_out
for rep in 1 2; do
  for side in base new; do
    if [ $side = base ]; then export XSG_LIB=$PWD/tools/ab/libxsg_base.so; else unset XSG_LIB; fi
    timeout 300 python bench.py --no-cpu --no-e2e --steps 40 --warmup 5 "$@" > gpurun_out/ab_${side}_${rep}.json 2> gpurun_out/ab_${side}_${rep}.err
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/ab_${side}_${rep}.json"))
    print("${side} ${rep}: %.1f GB/s  %.4f ms/step  kernel %.4f ms" % (d["value"], d["ms_per_step"], d["roofline"]["kernel_ms"]))
except Exception as e:
    print("${side} ${rep}: failed", e)
PY
  done
done
