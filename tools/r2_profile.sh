#!/bin/bash
# Round-2 evidence in one gpurun call: the contract bench lines, then (each only after the same command has exited 0
# without ncu) the ncu launch lists and --set full captures of the dominant kernels.  Outputs: gpurun_out/.
# Summaries for profiles/ are made afterwards on the CPU box (tools/summarize_ncu.py, tools/ncu_hot.py).
mkdir -p gpurun_out
O=gpurun_out
python bench.py --impl reference --steps 10 --warmup 3 > $O/r2_bench_reference.json 2> $O/bench_ref.err
python bench.py > $O/r2_bench.json 2> $O/bench.err
python bench.py --workload match_byte_offsets/the/10GB > $O/r2_bench_dense.json 2> $O/bench_dense.err
python bench.py --workload line_byte_offsets+lines/needle32/10GB > $O/r2_bench_lines32.json 2> $O/bench_l32.err
python bench.py --workload count/Sherlock/100MB > $O/r2_bench_100mb.json 2> $O/bench_100mb.err
python bench.py --impl reference --workload count/Sherlock/100MB --steps 20 --warmup 3 > $O/r2_bench_100mb_reference.json 2>> $O/bench_ref.err
: > $O/r2_sweep_50gb.jsonl
for m in 1 2 3 4 8 16 32 64; do
  timeout 600 python bench.py --workload count/sweep/50GB --pattern-len $m --steps 10 --warmup 3 --no-e2e --cpu-passes 1 2>$O/sweep_$m.err | tail -1 >> $O/r2_sweep_50gb.jsonl
done
SHORT="--steps 3 --warmup 3 --no-e2e --no-cpu"
python bench.py $SHORT > $O/short_default.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_default.csv \
    python bench.py $SHORT > $O/ncu_launches_default.log 2>&1
python bench.py --workload match_byte_offsets/the/10GB $SHORT > $O/short_dense.json 2>/dev/null && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_dense.csv \
    python bench.py --workload match_byte_offsets/the/10GB $SHORT > $O/ncu_launches_dense.log 2>&1
bash tools/r2_ncu.sh "count_kernel" count_lines+line_indices/Sherlock/10GB count 8 2 1
bash tools/r2_ncu.sh "count_kernel|emit_kernel" match_byte_offsets/the/10GB dense 8 4 2
bash tools/r2_ncu.sh "count_kernel" count/sweep/50GB that 4 2 1
ls -la $O | tail -40
for f in r2_bench r2_bench_dense r2_bench_lines32 r2_bench_reference r2_bench_100mb; do cut -c1-260 $O/$f.json; done
cut -c1-200 $O/r2_sweep_50gb.jsonl
