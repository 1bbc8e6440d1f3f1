#!/bin/bash
# same-box A/B of builds of libxsg.so: every tools/ab/libxsg_<name>.so (copies of earlier builds, git-ignored; loaded
# through XSG_LIB) and the in-tree one ("new"),
# interleaved, AB_REPS rounds.  usage: tools/r2_ab.sh [bench.py args]      (results: gpurun_out/ab_*.json)
mkdir -p gpurun_out
sides="$(ls tools/ab/libxsg_*.so 2>/dev/null | sed 's/.*libxsg_\(.*\)\.so/\1/') new"
for rep in $(seq 1 ${AB_REPS:-2}); do
  for side in $sides; do
    if [ $side = new ]; then unset XSG_LIB; else export XSG_LIB=$PWD/tools/ab/libxsg_$side.so; fi
    timeout 300 python bench.py --no-cpu --no-e2e --steps 40 --warmup 5 "$@" > gpurun_out/ab_${side}_${rep}.json 2> gpurun_out/ab_${side}_${rep}.err
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/ab_${side}_${rep}.json"))
    k = d["roofline"]["kernel_ms"]
    print("${side} ${rep}: %.1f GB/s  %.4f ms/step  kernel %.4f ms  rest %.1f us" % (d["value"], d["ms_per_step"], k, (d["ms_per_step"] - k) * 1e3))
except Exception as e:
    print("${side} ${rep}: failed", e)
PY
  done
done
