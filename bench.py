#!/usr/bin/env python3
"""bench.py -- the measurement contract of xsearch-b200 (one JSON line on stdout, rank 0).

A "step" is one pass of the scan-and-match hot path over one batch of synthetic corpus:

  python bench.py --gpus N --steps K --warmup W            # our CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code (oracle/_ref)

Default workload = BASELINE.json configs[1]: xs::count_lines + xs::line_indices of the 8-byte literal
"Sherlock" over 10 GB of synthetic text, device-resident, per GPU (weak scaling: every rank scans its own
contiguous 10 GB shard of one logical file; byte offsets carry the shard base, line indices are fixed up
after one all-gather of the shards' newline counts).

Keys beyond the base contract:
  value      GB/s scanned, whole job, corpus already resident in HBM, every kernel of the step included
  e2e        the same search through the C ABI on a pinned HOST buffer (xsg_search_host: H2D of every byte
             and D2H of the result vector inside the timed region)
  roofline   the byte-scanning kernel (count_kernel) against the measured HBM copy bandwidth
  cpu_baseline  the reference's unmodified matchers (oracle/_ref/libxsref.so) on the box's host cores,
             bounded sample of the same corpus, rank 0 / N == 1 only -- a reported baseline, not a target

Nothing here reads /root/reference at run time.  oracle/ is used only for the timed CPU legs.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

GB = 10 ** 9
METRIC = "GB/s scanned (device-resident and host-streamed) at 1/2/4/8 B200 vs host-core ref"  # BASELINE.json
NEEDLE32 = b"needle-needle-needle-needle-0032"

# name -> (tags, pattern, needle injected into the corpus, per-line injection rate, seed, bytes per GPU)
WORKLOADS = {
    # configs[1] -- the configuration the metric is quoted on
    "count_lines+line_indices/Sherlock/10GB": (("count_lines", "line_indices"), b"Sherlock", b"Sherlock",
                                               30.0 / 65536, 2, 10 * GB),
    # configs[2] -- dense output compaction (the 3-byte literal occurs naturally, ~1.8 % of positions)
    "match_byte_offsets/the/10GB": (("match_byte_offsets",), b"the", b"", 0.0, 3, 10 * GB),
    # configs[3] -- 32-byte literal, line_byte_offsets + lines; meant to be read through "e2e" (streamed)
    "line_byte_offsets+lines/needle32/10GB": (("line_byte_offsets", "lines"), NEEDLE32, NEEDLE32,
                                              30.0 / (1 << 20), 4, 10 * GB),
    # configs[4] -- xs::count, one pattern length per run (--pattern-len), 50 GB per GPU
    "count/sweep/50GB": (("count",), None, NEEDLE32 + NEEDLE32, 30.0 / (1 << 20), 5, 50 * GB),
}
DEFAULT_WORKLOAD = "count_lines+line_indices/Sherlock/10GB"
SWEEP_PATTERNS = {1: b"e", 2: b"th", 3: b"the", 4: b"that", 8: b"Sherlock", 16: (NEEDLE32 + NEEDLE32)[:16],
                  32: NEEDLE32, 64: NEEDLE32 + NEEDLE32}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default=DEFAULT_WORKLOAD)
    ap.add_argument("--pattern-len", type=int, default=8, help="count/sweep only")
    ap.add_argument("--gb", type=float, default=None, help="override bytes per GPU (in GB; invalidates the config)")
    ap.add_argument("--e2e-steps", type=int, default=None, help="default min(steps, 6)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--exchange", choices=["p2p", "nccl"], default="p2p",
                    help="N > 1: shard summaries over NVLink peer memory (own kernels) or an NCCL all-gather")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--clock-interval-ms", type=float, default=10.0,
                    help="pause between two NVML clock samples during the timed region")
    ap.add_argument("--cpu-sample-gb", type=float, default=2.0)
    return ap.parse_args()


def load_peaks():
    """Measured roofline denominators (driver-written); fallback per /opt/skills/guides/B200_PROFILING.md."""
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def load_traffic(workload: str):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture (or None)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f).get(workload)
        return t
    except Exception:
        return None


def workload_spec(args):
    tags, pat, needle, rate, seed, nbytes = WORKLOADS[args.workload]
    if pat is None:
        pat = SWEEP_PATTERNS[args.pattern_len]
    if args.gb is not None:
        nbytes = int(args.gb * GB)
    nbytes = (nbytes + 4095) // 4096 * 4096  # whole synthetic blocks (each ends with a newline)
    return tags, pat, needle, rate, seed, nbytes


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    def __init__(self, index: int, interval_ms: float = 10.0):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self.interval = max(0.001, interval_ms * 1e-3)
        self._stop_evt = threading.Event()
        self.ok = False
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception:
            pass

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop_evt.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                try:
                    r = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(self.interval)

    def stop(self):
        self._stop_evt.set()
        if self.is_alive():
            self.join()
        med = int(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---------------------------------------------------------------------------------------------------
# CPU legs (reference arm and cpu_baseline): the ONLY place where bench.py executes oracle/
# ---------------------------------------------------------------------------------------------------
_PU64 = C.POINTER(C.c_uint64)


class CpuChecker:
    """oracle/_ref/libxsref.so (the reference's unmodified matchers, kind "reference") or, when it is
    absent, oracle/libxsoracle.so (our C restatement, kind "port")."""

    def __init__(self):
        ref = os.path.join(ROOT, "oracle", "_ref", "libxsref.so")
        port = os.path.join(ROOT, "oracle", "libxsoracle.so")
        if os.path.exists(ref):
            self.lib, self.prefix, self.kind = C.CDLL(ref), "xsr_", "reference"
        elif os.path.exists(port):
            self.lib, self.prefix, self.kind = C.CDLL(port), "xso_", "port"
        else:
            raise RuntimeError("neither oracle/_ref/libxsref.so nor oracle/libxsoracle.so is built")
        vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
        self.count = getattr(self.lib, self.prefix + "count")
        self.count.restype, self.count.argtypes = u64, [vp, u64, C.c_char_p, u32, _PU64, u64, i32, i32]
        self.vec = {}
        for name in ("match_byte_offsets", "line_byte_offsets", "line_indices"):
            fn = getattr(self.lib, self.prefix + name)
            fn.restype, fn.argtypes = u64, [vp, u64, C.c_char_p, u32, _PU64, u64, _PU64, u64, i32]
            self.vec[name] = fn
        self.lines = getattr(self.lib, self.prefix + "lines")
        self.lines.restype = u64
        self.lines.argtypes = [vp, u64, C.c_char_p, u32, _PU64, u64, _PU64, _PU64, u64, i32]

    def one_pass(self, tags, addr, n, pat, bounds, nthreads, out_a, out_b):
        """Every requested tag once over buf[0..n); returns the result count of the last tag."""
        bp, nb, cap = bounds.ctypes.data_as(_PU64), int(bounds.size), int(out_a.size)
        pa, pb = out_a.ctypes.data_as(_PU64), out_b.ctypes.data_as(_PU64)
        r = 0
        for tag in tags:
            if tag == "count":
                r = self.count(addr, n, pat, len(pat), bp, nb, 0, nthreads)
            elif tag == "count_lines":
                r = self.count(addr, n, pat, len(pat), bp, nb, 1, nthreads)
            elif tag == "lines":
                r = self.lines(addr, n, pat, len(pat), bp, nb, pa, pb, cap, nthreads)
            else:
                r = self.vec[tag](addr, n, pat, len(pat), bp, nb, pa, cap, nthreads)
        return int(r)


def host_sample(xs, nbytes, needle, rate, seed, nthreads):
    """The first nbytes of the workload's corpus, generated on the host cores (same generator as the device)."""
    p = xs.SynthParams.make(seed, needle, rate)
    out = np.empty(nbytes, dtype=np.uint8)
    nblk = nbytes // xs.SYNTH_BLOCK
    per = (nblk + nthreads - 1) // nthreads

    def work(t):
        b0, b1 = t * per, min(nblk, (t + 1) * per)
        if b0 < b1:
            xs.lib.xsg_synth_host(out.ctypes.data + b0 * xs.SYNTH_BLOCK, b0, b1 - b0, C.byref(p))

    th = [threading.Thread(target=work, args=(t,)) for t in range(nthreads)]
    [t.start() for t in th]
    [t.join() for t in th]
    return out


def time_cpu(chk, tags, sample, pat, bounds, nthreads, passes):
    cap = max(1 << 20, sample.size // 32)
    out_a, out_b = np.empty(cap, dtype=np.uint64), np.empty(cap, dtype=np.uint64)
    best, r = None, 0
    for _ in range(passes):
        t0 = time.perf_counter()
        r = chk.one_pass(tags, sample.ctypes.data, int(sample.size), pat, bounds, nthreads, out_a, out_b)
        dt = time.perf_counter() - t0
        best = dt if best is None or dt < best else best
    return best, r


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # under torchrun only rank 0 measures; the other ranks exit 0 without work
    import xsearch_b200 as xs  # host helpers only (synthetic generator, chunker); no GPU call below
    tags, pat, needle, rate, seed, nbytes = workload_spec(args)
    cores = os.cpu_count() or 1
    sample_bytes = min(nbytes, int(args.cpu_sample_gb * GB)) // 4096 * 4096
    sample = host_sample(xs, sample_bytes, needle, rate, seed, cores)
    bounds = xs.chunk_bounds_host(sample)
    chk = CpuChecker()
    cap = max(1 << 20, sample.size // 32)
    out_a, out_b = np.empty(cap, dtype=np.uint64), np.empty(cap, dtype=np.uint64)
    for _ in range(args.warmup):
        chk.one_pass(tags, sample.ctypes.data, sample_bytes, pat, bounds, cores, out_a, out_b)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = chk.one_pass(tags, sample.ctypes.data, sample_bytes, pat, bounds, cores, out_a, out_b)
    dt = time.perf_counter() - t0
    value = sample_bytes * args.steps / dt / 1e9
    sample_desc = "first %.2f GB of the workload's corpus per step, in host RAM, v1 chunks (>=16 MiB), %d threads" % (
        sample_bytes / 1e9, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": args.workload, "pattern": pat.decode("latin-1"), "tags": list(tags),
                   "bytes_per_step": sample_bytes, "results_last_tag": r},
        "cpu_baseline": {"value": round(value, 3), "unit": "GB/s", "cores": cores, "kind": chk.kind,
                         "sample": sample_desc},
        "e2e": {"value": round(value, 3), "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# our arm: the CUDA path through the C ABI
# ---------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    import xsearch_b200 as xs

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        # stdout carries exactly one JSON line: NCCL's "NCCL version ..." banner (NCCL_DEBUG=VERSION on some boxes)
        # would go there too
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # whatever level is set: not on stdout
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tags, pat, needle, rate, seed, nbytes = workload_spec(args)
    kinds = 0
    for t in tags:
        kinds |= xs.TAGS[t]
    vec_kind = xs.TAGS[tags[-1]] if xs.TAGS[tags[-1]] not in (xs.COUNT, xs.COUNT_LINES) else 0

    ctx = xs.Context(local)
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    # this rank's shard of the logical file: blocks [rank * nblk, (rank + 1) * nblk)
    nblk = nbytes // xs.SYNTH_BLOCK
    params = xs.SynthParams.make(seed, needle, rate)
    corpus = ctx.corpus(nbytes)
    corpus.fill_synth(params, first_block=rank * nblk)
    corpus.chunk()
    bounds = corpus.bounds()
    shard_base = rank * nbytes

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def finish(res):
        """Multi-GPU fix-up: one all-gather of the shards' newline / result counts, then the line-index
        base is added on the device; returns the global number of results of the last tag."""
        cnt = res.count(xs.TAGS[tags[-1]])
        if world == 1:
            return cnt
        line_base, _, totals = xs.shard_exchange(res.newlines, [cnt])
        res.rebase(0, line_base)
        return totals[0]

    # ---- device-resident: value -------------------------------------------------------------------
    launches = 0
    main_ms = []
    results = 0
    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            res = ctx.scan(corpus, pat, kinds, offset_base=shard_base)
            finish(res)
            res.free()
        ex = None
        if world > 1 and args.exchange == "p2p":
            # shard summaries travel as peer stores over NVLink (xsg_exchange, repo:xsearch_b200/csrc/xsg_exchange.cu)
            ex = xs.Exchange(ctx)
            for _ in range(2):
                res = ctx.scan(corpus, pat, kinds, offset_base=shard_base)
                ex.push()
                ex.rebase(res)
                res.free()
        elif world > 1:  # first use of the collective the timed loop relies on
            warm = torch.zeros((world, 4), dtype=torch.int64, device="cuda")
            xs.sharded_scan_async(ctx, corpus, pat, kinds, shard_base, warm).free()
        sampler = ClockSampler(local, args.clock_interval_ms)
        barrier()
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        table = torch.zeros((world, 4), dtype=torch.int64, device="cuda") if world > 1 else None
        for _ in range(args.steps):
            if ex is not None:
                # scan + push of the shard summary into every rank's table + fix-up kernel that waits for the
                # round on the device: no host round trip, no collective call
                res = ctx.scan(corpus, pat, kinds, offset_base=shard_base)
                ex.push()
                ex.rebase(res)
            elif world > 1:
                # scan + all-gather of the device-resident shard summaries + on-device line-index fix-up:
                # the exchange never returns to the host
                res = xs.sharded_scan_async(ctx, corpus, pat, kinds, shard_base, table)
            else:
                res = ctx.scan(corpus, pat, kinds, offset_base=shard_base)
                results = res.count(xs.TAGS[tags[-1]])
            t = res.timings()
            launches += t.kernel_launches + (2 if ex is not None else (1 if world > 1 and rank > 0 and (kinds & xs.LINE_INDICES) else 0))
            main_ms.append(t.main_kernel_ms)
            raw = res.raw_occurrences
            res.free()
        e1.record(stream)
        barrier()
        clocks = sampler.stop()
        ms = max_over_ranks(e0.elapsed_time(e1))
        if world > 1:
            col = {xs.COUNT: 1, xs.MATCH_BYTE_OFFSETS: 1, xs.COUNT_LINES: 2, xs.LINE_BYTE_OFFSETS: 2, xs.LINE_INDICES: 2,
                   xs.LINES: 3}[xs.TAGS[tags[-1]]]
            results = int(ex.totals()[:, col].sum()) if ex is not None else int(table[:, col].sum().item())
    total_bytes = nbytes * world
    value = total_bytes * args.steps / (ms * 1e-3) / 1e9
    kernel_ms = float(np.mean(main_ms))

    # ---- e2e: pinned host buffer -> xsg_search_host -> result vector on the host --------------------
    e2e = None
    if not args.no_e2e:
        hp = C.c_void_p()
        xs._check(xs.lib.xsg_host_alloc(nbytes, C.byref(hp)))
        xs._check(xs.lib.xsg_corpus_download(corpus._h, hp, 0, nbytes))  # same bytes, now in pinned host RAM
        ksteps = args.e2e_steps or min(args.steps, 6)
        d2h = 0
        pinned_out = [None]
        with torch.cuda.stream(stream):
            def e2e_step():
                r = ctx.search_host(hp.value, nbytes, pat, kinds, bounds=bounds, offset_base=shard_base)
                finish(r)
                got = 8
                if vec_kind:
                    # destination in pinned memory, allocated once: the D2H of a large vector is then one DMA
                    need = r.count(vec_kind)
                    if pinned_out[0] is None or pinned_out[0].size < need:
                        pinned_out[0] = xs.pinned_empty(need + need // 8 + 1024)
                    v = r.vector(vec_kind, out=pinned_out[0])
                    got += v.nbytes
                    if vec_kind == xs.LINES:
                        got += r.line_spans()[1].nbytes  # the line lengths travel as well
                r.free()
                return got
            for _ in range(min(2, args.warmup)):
                e2e_step()
            barrier()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record(stream)
            for _ in range(ksteps):
                d2h = e2e_step()
            f1.record(stream)
            barrier()
            ems = max_over_ranks(f0.elapsed_time(f1))
        e2e = {"value": round(total_bytes * ksteps / (ems * 1e-3) / 1e9, 3), "unit": "GB/s",
               "h2d_bytes_per_step": int(total_bytes), "d2h_bytes_per_step": int(sum_over_ranks(d2h)),
               "steps": ksteps, "ms_per_step": round(ems / ksteps, 3),
               "path": "xsg_search_host: pinned host buffer -> double-buffered H2D segments -> scan -> D2H of results"}
        xs.lib.xsg_host_free(hp)

    # ---- cpu_baseline (rank 0, N == 1 only): the reference's matchers on a bounded sample -----------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        try:
            cores = os.cpu_count() or 1
            sb = min(nbytes, int(args.cpu_sample_gb * GB)) // 4096 * 4096
            sample = corpus.download(0, sb)
            sbounds = xs.chunk_bounds_host(sample)
            chk = CpuChecker()
            best, _ = time_cpu(chk, tags, sample, pat, sbounds, cores, passes=3)
            cpu = {"value": round(sb / best / 1e9, 3), "unit": "GB/s", "cores": cores, "kind": chk.kind,
                   "sample": "first %.2f GB of the same corpus in host RAM, v1 chunks, best of 3 passes" % (sb / 1e9)}
        except Exception as e:  # the baseline is informative; never lose the GPU numbers over it
            cpu = {"value": None, "unit": "GB/s", "cores": 0, "kind": "unavailable", "sample": str(e)[:200]}

    peak, peak_src = load_peaks()
    achieved = nbytes / (kernel_ms * 1e-3) / 1e9
    results_rank = results if world == 1 else None
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": args.workload, "pattern": pat.decode("latin-1"), "tags": list(tags),
                   "bytes_per_gpu": nbytes, "global_results": int(results), "raw_occurrences_per_gpu": int(raw),
                   "reference_chunks_per_gpu": int(bounds.size), "l2": "inputs (10 GB per GPU) far larger than the 126 MB L2",
                   "parallelism": "shard%d" % world,
                   **({"exchange": "nvlink peer stores (xsg_exchange)" if args.exchange == "p2p" else "nccl all-gather"}
                      if world > 1 else {})},
        "e2e": e2e,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "hbm", "kernel": "xsg::count_kernel", "achieved": round(achieved, 1),
                     "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                     "traffic": load_traffic(args.workload), "peak_source": peak_src,
                     "kernel_ms": round(kernel_ms, 4), "algorithmic_bytes_per_launch": nbytes,
                     "step_frac": round((value / world) / peak, 4)},
        "cpu_baseline": cpu,
    }
    del results_rank
    corpus.free()
    ctx.close()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
