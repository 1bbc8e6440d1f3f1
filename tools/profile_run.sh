#!/bin/bash
# Round evidence in one gpurun call: the contract bench lines, then (each only after the same command has exited 0
# without ncu) the ncu launch lists and one --set full capture of the dominant kernels.  Outputs: gpurun_out/.
# Summaries for profiles/ are made afterwards with tools/summarize_ncu.py (ncu -i runs on the CPU box).
mkdir -p gpurun_out
O=gpurun_out
python bench.py --impl reference --steps 10 --warmup 3 > $O/r1_bench_reference.json 2> $O/bench_ref.err
python bench.py > $O/r1_bench.json 2> $O/bench.err
python bench.py --workload match_byte_offsets/the/10GB --no-cpu > $O/r1_bench_dense.json 2> $O/bench_dense.err
python bench.py --workload line_byte_offsets+lines/needle32/10GB --no-cpu > $O/r1_bench_lines32.json 2> $O/bench_l32.err
bash tools/sweep_run.sh > $O/sweep.log 2>&1
cp $O/sweep.jsonl $O/r1_sweep_50gb.jsonl

SHORT="--steps 3 --warmup 3 --no-e2e --no-cpu"
python bench.py $SHORT > $O/short_default.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_default.csv \
    python bench.py $SHORT > $O/ncu_launches_default.log 2>&1
python bench.py --workload match_byte_offsets/the/10GB $SHORT > $O/short_dense.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_dense.csv \
    python bench.py --workload match_byte_offsets/the/10GB $SHORT > $O/ncu_launches_dense.log 2>&1
# full captures: the third launch of each kernel (warm)
python bench.py $SHORT > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:count_kernel -s 2 -c 1 -f -o $O/full_count_kernel \
    python bench.py $SHORT > $O/ncu_full_count.log 2>&1
python bench.py --workload match_byte_offsets/the/10GB $SHORT > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"count_kernel|emit_kernel" -s 4 -c 2 -f -o $O/full_dense \
    python bench.py --workload match_byte_offsets/the/10GB $SHORT > $O/ncu_full_dense.log 2>&1
ls -la $O | tail -30
for f in r1_bench r1_bench_dense r1_bench_lines32 r1_bench_reference; do cut -c1-260 $O/$f.json; done
tail -9 $O/sweep.log
