#!/bin/bash
# development round trip: parity tests of the scan kernels, the per-variant micro-benchmark, the dense bench line
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $O/dev_parity.log 2>&1; echo "parity rc=$?"; tail -4 $O/dev_parity.log
timeout 600 python tools/dev_bench.py 8 > $O/dev_bench.txt 2>&1; echo "dev_bench rc=$?"; cat $O/dev_bench.txt
timeout 300 python bench.py --workload match_byte_offsets/the/10GB --steps 20 --no-e2e > $O/dev_dense.json 2> $O/dev_dense.err; echo "dense rc=$?"; cut -c1-400 $O/dev_dense.json; tail -2 $O/dev_dense.err
if [ -n "$1" ]; then ./tools/hostreg_probe 4 > $O/box_probe2.txt 2>&1; grep -v "^D O_DIRECT\|FAIL" $O/box_probe2.txt; fi
