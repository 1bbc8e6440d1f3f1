#!/bin/bash
# evidence on the final build of round 2 (after the post-scan stages were shortened): GPU suite, smoke, the headline,
# dense and lines32 bench lines, the headline's ncu launch list
mkdir -p gpurun_out
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > $O/r2_bench.json 2> $O/bench.err; echo "bench rc=$?"
python bench.py --workload match_byte_offsets/the/10GB > $O/r2_bench_dense.json 2> $O/bench_dense.err; echo "dense rc=$?"
python bench.py --workload line_byte_offsets+lines/needle32/10GB > $O/r2_bench_lines32.json 2> $O/bench_lines32.err; echo "lines32 rc=$?"
SHORT="--steps 3 --warmup 3 --no-e2e --no-cpu"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_default.csv \
    python bench.py $SHORT > $O/ncu_launches_default.log 2>&1
for f in r2_bench r2_bench_dense r2_bench_lines32; do cut -c1-260 $O/$f.json; done
