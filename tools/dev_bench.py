#!/usr/bin/env python3
"""Developer micro-benchmark: device-resident scan throughput per kernel variant (not the contract bench)."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import xsearch_b200 as xs


def main():
    gb = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
    n = int(gb * (1 << 30)) // 4096 * 4096
    ctx = xs.Context(0)
    print(ctx.device_info())
    p = xs.SynthParams.make(2, b"Sherlock", 30.0 / 65536)
    corpus = ctx.corpus(n)
    t0 = time.time()
    corpus.fill_synth(p)
    print("fill %.2fs" % (time.time() - t0))
    t0 = time.time()
    corpus.chunk()
    print("chunk %.3fs, %d chunks" % (time.time() - t0, corpus.bounds().size))
    K = xs
    configs = [
        ("count m=8 Sherlock", b"Sherlock", K.COUNT),
        ("count_lines+line_indices m=8", b"Sherlock", K.COUNT_LINES | K.LINE_INDICES),
        ("count_lines m=8", b"Sherlock", K.COUNT_LINES),
        ("match_byte_offsets m=3 the", b"the", K.MATCH_BYTE_OFFSETS),
        ("count m=3 the", b"the", K.COUNT),
        ("count m=1 e", b"e", K.COUNT),
        ("count m=2 th", b"th", K.COUNT),
        ("count m=4 that", b"that", K.COUNT),
        ("count m=4 then", b"then", K.COUNT),
        ("match_byte_offsets m=4 that", b"that", K.MATCH_BYTE_OFFSETS),
        ("match_byte_offsets m=1 e", b"e", K.MATCH_BYTE_OFFSETS),
        ("all kinds m=8", b"Sherlock", K.COUNT | K.COUNT_LINES | K.MATCH_BYTE_OFFSETS | K.LINE_BYTE_OFFSETS | K.LINE_INDICES | K.LINES),
        ("count m=16", b"inspector said t", K.COUNT),
        ("count m=32", b"needle-needle-needle-needle-0032", K.COUNT),
        ("line_byte_offsets+lines m=8", b"Sherlock", K.LINE_BYTE_OFFSETS | K.LINES),
        ("icase count m=8", b"sherlock", K.COUNT, True),
        ("icase count_lines+line_indices m=8", b"sherlock", K.COUNT_LINES | K.LINE_INDICES, True),
        ("icase count m=3 the", b"the", K.COUNT, True),
        ("icase count m=16", b"inspector said t", K.COUNT, True),
        ("count m=9 bordered Sherlock S", b"Sherlock S", K.COUNT),
    ]
    only = sys.argv[2] if len(sys.argv) > 2 else ""
    for cfg in configs:
        name, pat, kinds = cfg[:3]
        icase = len(cfg) > 3 and cfg[3]
        if only and only not in name:
            continue
        best = None
        for it in range(6):
            res = ctx.scan(corpus, pat, kinds, ignore_case=icase)
            t = res.timings()
            cnt = [res.count(k) for k in (K.COUNT, K.COUNT_LINES, K.MATCH_BYTE_OFFSETS, K.LINE_INDICES) if kinds & k]
            raw = res.raw_occurrences
            res.free()
            if it >= 2 and (best is None or t.scan_ms < best[0]):
                best = (t.scan_ms, t.main_kernel_ms, t.post_ms, t.kernel_launches)
        print("%-34s scan %8.3f ms  main %8.3f ms  post %7.3f ms  launches %2d  -> %7.1f GB/s scan, %7.1f GB/s main  "
              "counts=%s raw=%d" % (name, best[0], best[1], best[2], best[3], n / best[0] / 1e6, n / best[1] / 1e6,
                                    cnt, raw), flush=True)
    corpus.free()
    ctx.close()


if __name__ == "__main__":
    main()
