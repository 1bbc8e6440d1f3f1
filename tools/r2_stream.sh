#!/bin/bash
# file e2e through the C++ API + BASELINE configs[3] (100 GB streamed) at N = 1
mkdir -p gpurun_out
O=gpurun_out
bash tools/file_e2e_run.sh 8 > /dev/null 2>&1; cat $O/file_e2e.txt
timeout 900 python bench.py --workload line_byte_offsets+lines/needle32/100GB-streamed --steps 10 --warmup 3 --e2e-steps 3 --cpu-passes 1 \
  > $O/r2_cfg3_n1.json 2> $O/r2_cfg3_n1.err; echo "cfg3 rc=$?"; tail -3 $O/r2_cfg3_n1.err; cat $O/r2_cfg3_n1.json
