#!/bin/bash
# What a caller of xs::extern_search sees for a regular expression next to the plain literal: xsgrep on a synthetic
# file in tmpfs (page cache), pipeline time as reported by the library.  Output: gpurun_out/regex_e2e.txt
#   tools/regex_e2e_run.sh [GiB]
GIB=${1:-1}
O=gpurun_out
mkdir -p $O
F=/dev/shm/xs_regex_corpus.txt
python - <<PY
import xsearch_b200 as xs
p = xs.SynthParams.make(9, b"Sherlock", 30.0 / 8192)
xs.synth_host(int($GIB * (1 << 30)), p).tofile("$F")
PY
X=xsearch_b200/cpp/xsgrep
{
  for spec in "-F|Sherlock" "-E|She[r ]lock" "-E|(?:Sherlock|which) [a-z]+" "-E|wh(ich|ere)|Sherlock" "-E|th(e|is|at)\\b"; do
    mode=${spec%%|*}; pat=${spec#*|}
    for rep in 1 2; do
      echo "== xsgrep $mode --tag count_lines --readers 8 -j 8 '$pat' (run $rep)"
      $X $mode --tag count_lines --readers 8 -j 8 "$pat" $F 2>&1
    done
  done
} > $O/regex_e2e.txt 2>&1
rm -f $F
cat $O/regex_e2e.txt
