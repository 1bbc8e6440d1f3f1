#!/bin/bash
# last evidence of the round on the final build: smoke, the two headline lines, their ncu launch lists, emit capture
mkdir -p gpurun_out
O=gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --impl reference --steps 10 --warmup 3 > $O/r2_bench_reference.json 2> $O/bench_ref.err
python bench.py > $O/r2_bench.json 2> $O/bench.err
python bench.py --workload match_byte_offsets/the/10GB > $O/r2_bench_dense.json 2> $O/bench_dense.err
SHORT="--steps 3 --warmup 3 --no-e2e --no-cpu"
python bench.py $SHORT > $O/short_default.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_default.csv \
    python bench.py $SHORT > $O/ncu_launches_default.log 2>&1
python bench.py --workload match_byte_offsets/the/10GB $SHORT > $O/short_dense.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_dense.csv \
    python bench.py --workload match_byte_offsets/the/10GB $SHORT > $O/ncu_launches_dense.log 2>&1
bash tools/r2_ncu.sh "emit_kernel" match_byte_offsets/the/10GB emit 8 2 1
for f in r2_bench r2_bench_dense r2_bench_reference; do cut -c1-200 $O/$f.json; done
