#!/usr/bin/env python3
"""Dynamic instruction mix of one kernel from an ncu report's source page (needs -lineinfo / --import-source on):
   python tools/ncu_hot.py report.ncu-rep [kernel-substring] [--top N]
prints executed warp instructions per opcode (and per pipe), per granule if --granules G is given, and the N hottest
SASS instructions with their stall samples."""
import collections
import csv
import subprocess
import sys

sys.path.insert(0, __file__.rsplit("/", 1)[0])
from sass_hist import pipe  # noqa: E402


def main():
    rep = sys.argv[1]
    sub = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else ""
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 0
    gran = float(sys.argv[sys.argv.index("--granules") + 1]) if "--granules" in sys.argv else 0
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    blocks = txt.split('"Kernel Name",')
    for b in blocks[1:]:
        lines = b.splitlines()
        name = lines[0].strip('",')
        if sub and sub not in name:
            continue
        rows = list(csv.reader(lines[1:]))
        hdr = rows[0]
        iS, iE, iW = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
        ops, pipes, hot = collections.Counter(), collections.Counter(), []
        total = 0
        for r in rows[1:]:
            if len(r) <= iE or not r[iE]:
                continue
            n = int(r[iE])
            toks = r[iS].split()
            op = toks[1] if toks and toks[0].startswith("@") and len(toks) > 1 else (toks[0] if toks else "?")
            ops[op.split(".")[0]] += n
            pipes[pipe(op)] += n
            total += n
            hot.append((n, int(r[iW] or 0), r[iS][:90]))
        print("## %s\n" % name.replace("xsg::<unnamed>::", "")[:120])
        print("%d warp instructions executed%s: %s\n" % (total, " (%.1f per granule)" % (total / gran) if gran else "",
                                                       ", ".join("%s %.1f%%" % (k, 100.0 * v / total) for k, v in pipes.most_common())))
        print("| opcode | executed | share |%s\n|---|---:|---:|%s" % (" per granule |" if gran else "", "---:|" if gran else ""))
        for op, n in ops.most_common(24):
            print("| %s | %d | %.1f%% |%s" % (op, n, 100.0 * n / total, " %.1f |" % (n / gran) if gran else ""))
        if top:
            print("\nhottest instructions (executed, stall samples):")
            for n, w, src in sorted(hot, key=lambda t: -t[1])[:top]:
                print("  %12d %7d  %s" % (n, w, src))
        print()


if __name__ == "__main__":
    main()
