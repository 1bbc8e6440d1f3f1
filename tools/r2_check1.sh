#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" 
tail -5 $O/pytest_gpu.log
python bench.py --impl reference --steps 5 --warmup 2 > $O/r2_bench_reference.json 2> $O/bench_ref.err; echo "ref rc=$?"
python bench.py > $O/r2_bench.json 2> $O/bench.err; echo "bench rc=$?"; tail -3 $O/bench.err
python bench.py --workload match_byte_offsets/the/10GB --steps 20 > $O/r2_bench_dense.json 2> $O/bench_dense.err; echo "dense rc=$?"; tail -3 $O/bench_dense.err
python bench.py --workload count/Sherlock/100MB --steps 20 > $O/r2_bench_100mb.json 2> $O/bench_100mb.err; echo "100mb rc=$?"; tail -3 $O/bench_100mb.err
for f in r2_bench_reference r2_bench r2_bench_dense r2_bench_100mb; do echo "== $f"; cat $O/$f.json; done
