#!/usr/bin/env python3
"""Turns ncu output brought back in gpurun_out/ into the small text summaries committed under profiles/.

  summarize_ncu.py launches <launches.csv> <out.md> [title]      per-kernel launch counts, mean time, share
  summarize_ncu.py raw <report.ncu-rep> <out.md> [title]         the roofline-relevant counters per launch
"""
import collections
import csv
import subprocess
import sys

RAW_METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "launch__occupancy_limit_registers", "launch__grid_size", "launch__block_size",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
]


def launches(path, out, title):
    lines = [l for l in open(path) if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    agg = collections.OrderedDict()
    for r in rows:
        agg.setdefault(r["Kernel Name"], []).append(float(r["Metric Value"].replace(",", "")))
    # the corpus fill / chunker run once before the steps; shares are over the step kernels only
    setup = ("synth_kernel", "chunker_kernel")
    step_total = sum(sum(v) for k, v in agg.items() if not any(s in k for s in setup))
    with open(out, "w") as f:
        f.write("# %s\n\n" % title)
        f.write("Source: `ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised launches:\n"
                "compare shares, not absolutes).  Shares are over the kernels of the timed steps (set-up kernels listed apart).\n\n")
        f.write("| kernel | launches | mean us | share of step |\n|---|---:|---:|---:|\n")
        for k, v in agg.items():
            name = k.replace("unnamed>::", "").replace("void ", "")
            name = name.split("(")[0]
            share = "set-up" if any(s in k for s in setup) else "%.3f" % (sum(v) / step_total)
            f.write("| `%s` | %d | %.1f | %s |\n" % (name, len(v), sum(v) / len(v) / 1e3, share))


def raw(path, out, title):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(out, "w") as f:
        f.write("# %s\n\nSource: `ncu --set full --clock-control none --import-source on`, `ncu -i ... --page raw --csv`.\n\n" % title)
        for r in rows[2:]:
            f.write("## %s (launch id %s)\n\n| metric | value | unit |\n|---|---:|---|\n" % (
                r[hdr.index("Kernel Name")].replace("unnamed>::", ""), r[hdr.index("ID")]))
            for m in RAW_METRICS:
                if m in hdr:
                    f.write("| %s | %s | %s |\n" % (m, r[hdr.index(m)], units[hdr.index(m)]))
            f.write("\n")


if __name__ == "__main__":
    mode, src, dst = sys.argv[1:4]
    title = sys.argv[4] if len(sys.argv) > 4 else src
    (launches if mode == "launches" else raw)(src, dst, title)
