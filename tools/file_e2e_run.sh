#!/bin/bash
# What a user of xs::extern_search sees: a FILE (tmpfs, i.e. page cache) searched through the C++ layer
# (reader threads -> pinned segments -> xsg_stream).  Writes gpurun_out/file_e2e.txt.
#   bash tools/file_e2e_run.sh [GiB]
mkdir -p gpurun_out
G=${1:-8}
F=/dev/shm/xsg_file_e2e.txt
O=gpurun_out/file_e2e.txt
python - "$G" "$F" <<'EOF'
import os, sys
sys.path.insert(0, os.getcwd())
import bench, xsearch_b200 as xs
gib, path = int(sys.argv[1]), sys.argv[2]
# the default workload's corpus (needle, rate and seed of bench.py), generated on the host cores
tags, pat, needle, rate, seed, nbytes, scaling = bench.WORKLOADS[bench.DEFAULT_WORKLOAD]
bench.host_corpus(gib << 30, needle, rate, seed, os.cpu_count() or 8).tofile(path)
EOF
ls -l $F > $O
echo "host cores: $(nproc)" >> $O
: > $O.err
for tag in count lines; do
  for r in 1 4 8 16; do
    for rep in 1 2; do
      s=$(date +%s.%N)
      n=$(./xsearch_b200/cpp/xsgrep --tag $tag --readers $r Sherlock $F 2>> $O.err | wc -l)
      e=$(date +%s.%N)
      echo "tag=$tag readers=$r rep=$rep out_lines=$n wall_s=$(python -c "print('%.3f' % ($e - $s))") | $(tail -1 $O.err)" >> $O
    done
  done
done
# reader mode A/B (xs::Options::reader_mode): memcpy out of a mapping instead of pread, DESIGN.md 8
for mode in pread mmap; do
  for r in 1 4 8 16; do
    ./xsearch_b200/cpp/xsgrep --tag count --readers $r --reader $mode Sherlock $F 2>> $O.err > /dev/null
    echo "tag=count reader=$mode readers=$r | $(tail -1 $O.err)" >> $O
  done
done
rm -f $F
cat $O
