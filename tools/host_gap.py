#!/usr/bin/env python3
"""Where does the time between two synchronous scans go?  Wall clock per step against the device time of the scan."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import xsearch_b200 as xs


def main():
    gb = float(sys.argv[1]) if len(sys.argv) > 1 else 8.0
    n = int(gb * (1 << 30)) // 4096 * 4096
    ctx = xs.Context(0)
    corpus = ctx.corpus(n)
    corpus.fill_synth(xs.SynthParams.make(2, b"Sherlock", 30.0 / 65536))
    corpus.chunk()
    which = sys.argv[2] if len(sys.argv) > 2 else ""
    for name, kinds in (("count", xs.COUNT), ("count_lines+line_indices", xs.COUNT_LINES | xs.LINE_INDICES)):
        if which and which != name:
            continue
        for it in range(5):
            ctx.scan(corpus, b"Sherlock", kinds).free()
        steps = 200
        dev = call = 0.0
        scans, mains, posts = [], [], []
        t0 = time.perf_counter()
        for it in range(steps):
            c0 = time.perf_counter()
            res = ctx.scan(corpus, b"Sherlock", kinds)
            call += time.perf_counter() - c0
            t = res.timings()
            dev += t.scan_ms
            scans.append(t.scan_ms), mains.append(t.main_kernel_ms), posts.append(t.post_ms)
            res.free()
        wall = (time.perf_counter() - t0) / steps * 1e3
        print("%-28s wall %.4f ms/step  scan() call %.4f  device scan_ms %.4f  -> outside the device window %.1f us, "
              "python outside scan() %.1f us" % (name, wall, call / steps * 1e3, dev / steps,
                                                 (wall - dev / steps) * 1e3, (wall - call / steps * 1e3) * 1e3))
        for nm, v in (("scan_ms", scans), ("main_ms", mains), ("post_ms", posts)):
            v = sorted(v)
            print("    %-8s min %.4f  p10 %.4f  median %.4f  p90 %.4f  max %.4f  mean %.4f" %
                  (nm, v[0], v[len(v) // 10], v[len(v) // 2], v[len(v) * 9 // 10], v[-1], sum(v) / len(v)))


if __name__ == "__main__":
    main()
