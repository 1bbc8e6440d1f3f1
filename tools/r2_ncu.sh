#!/bin/bash
# ncu --set full captures of the kernels named in $1 (regex) on workload $2 [pattern-len $3]; summaries made on the CPU box
mkdir -p gpurun_out
O=gpurun_out
K=${1:-count_kernel}; W=${2:-count_lines+line_indices/Sherlock/10GB}; N=${3:-out}; PL=${4:-8}; SKIP=${5:-4}; CNT=${6:-2}
SHORT="--steps 3 --warmup 3 --no-e2e --no-cpu --workload $W --pattern-len $PL"
python bench.py $SHORT > $O/short_$N.json 2>$O/short_$N.err && \
ncu --set full --clock-control none --import-source on -k regex:"$K" -s $SKIP -c $CNT -f -o $O/full_$N \
    python bench.py $SHORT > $O/ncu_full_$N.log 2>&1
ls -la $O/full_$N.ncu-rep; tail -3 $O/ncu_full_$N.log
