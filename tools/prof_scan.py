#!/usr/bin/env python3
"""Small fixed workload for ncu: a few scans of a 1 GiB synthetic corpus.  argv[1] = config name."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import xsearch_b200 as xs

CONFIGS = {
    "lines": (b"Sherlock", xs.COUNT_LINES | xs.LINE_INDICES),
    "count": (b"Sherlock", xs.COUNT),
    "dense": (b"the", xs.MATCH_BYTE_OFFSETS),
}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "lines"
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    pat, kinds = CONFIGS[name]
    ctx = xs.Context(0)
    corpus = ctx.corpus(1 << 30)
    corpus.fill_synth(xs.SynthParams.make(2, b"Sherlock", 30.0 / 65536))
    corpus.chunk()
    for _ in range(reps):
        res = ctx.scan(corpus, pat, kinds)
        t = res.timings()
        res.free()
    print(name, "main %.3f ms scan %.3f ms" % (t.main_kernel_ms, t.scan_ms))


if __name__ == "__main__":
    main()
