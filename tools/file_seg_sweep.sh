#!/bin/bash
# file search through xs::extern_search: segment size x reader count (8 GiB in tmpfs)
mkdir -p gpurun_out
F=/dev/shm/xsg_file_e2e.txt
O=gpurun_out/file_seg_sweep.txt
python - "$F" <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import bench
tags, pat, needle, rate, seed, nbytes, scaling = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
bench.host_corpus(8 << 30, needle, rate, seed, os.cpu_count() or 8).tofile(sys.argv[1])
PY
: > $O
for seg in 67108864 134217728 268435456 536870912; do
  for r in 12 16 24; do
    for rep in 1 2; do
      ./xsearch_b200/cpp/xsgrep --tag count --readers $r --segment-bytes $seg Sherlock $F 2> $O.err > /dev/null
      echo "segment=$((seg >> 20))MiB readers=$r | $(tail -1 $O.err)" >> $O
    done
  done
done
rm -f $F
cat $O
