#!/usr/bin/env python3
"""bench.py -- the measurement contract of xsearch-b200 (one JSON line on stdout, rank 0).

A "step" is one pass of the scan-and-match hot path over one batch of synthetic corpus:

  python bench.py --gpus N --steps K --warmup W            # our CUDA path
  python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU code (oracle/_ref)

Default workload = BASELINE.json configs[1]: xs::count_lines + xs::line_indices of the 8-byte literal
"Sherlock" over 10 GB of synthetic text, device-resident, per GPU (weak scaling: every rank scans its own
contiguous 10 GB shard of one logical file; byte offsets carry the shard base, line indices are fixed up
by the shard exchange).  The other BASELINE configs are workloads of their own (--workload).

Keys beyond the base contract:
  value      GB/s scanned, whole job, corpus already resident in HBM, every kernel of the step included
  e2e        the same search through the C ABI on a pinned HOST buffer (xsg_search_host: H2D of every byte
             and D2H of the result vector inside the timed region); h2d_peak = plain cudaMemcpyAsync of the same
             pinned buffers on all ranks at once, measured in the same run; frac = value / h2d_peak
  roofline   the byte-scanning kernel (count_kernel) against the measured HBM copy bandwidth
  parity     the result vectors of THIS run (the device-resident scan and the last host-streamed step, fix-ups
             included) compared element by element with the reference's own CPU code (oracle/_ref/libxsref.so) on
             the same bytes: every rank checks all of its shard, line-index bases against an independent newline
             count of the lower shards.  A mismatch makes the run exit non-zero.
  cpu_baseline  the reference run of the parity block, timed (N == 1: all host cores, all bytes of the workload)

Nothing here reads /root/reference at run time.  oracle/ is used only for the CPU legs (checker and baseline).
The reference arm maps nothing of the product: its corpus comes from oracle/libxscorpus.so.
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
SYNTH_BLOCK = 4096

# name -> (tags, pattern, needle injected into the corpus, per-line injection rate, seed, bytes, scaling)
#   scaling "weak": bytes PER GPU (every rank its own shard); "strong": bytes of the WHOLE job, split across the ranks
WORKLOADS = {
    # configs[0] -- the reference's own CPU-runnable case (plumbing): ~50 hits in 100 MB
    "count/Sherlock/100MB": (("count",), b"Sherlock", b"Sherlock", 50.0 / 3.4e6, 1, 100 * 10 ** 6, "weak"),
    # configs[1] -- the configuration the metric is quoted on
    "count_lines+line_indices/Sherlock/10GB": (("count_lines", "line_indices"), b"Sherlock", b"Sherlock",
                                               30.0 / 65536, 2, 10 * GB, "weak"),
    # configs[2] -- dense output compaction (the 3-byte literal occurs naturally, ~1.8 % of positions)
    "match_byte_offsets/the/10GB": (("match_byte_offsets",), b"the", b"", 0.0, 3, 10 * GB, "weak"),
    # configs[3] on one GPU's HBM (device-resident shape of the same search)
    "line_byte_offsets+lines/needle32/10GB": (("line_byte_offsets", "lines"), NEEDLE32, NEEDLE32,
                                              30.0 / (1 << 20), 4, 10 * GB, "weak"),
    # configs[3] as written: ONE 100 GB file in pinned host memory, streamed, split across 1/2/4/8 GPUs; read "e2e"
    "line_byte_offsets+lines/needle32/100GB-streamed": (("line_byte_offsets", "lines"), NEEDLE32, NEEDLE32,
                                                        30.0 / (1 << 20), 4, 100 * GB, "strong"),
    # configs[4] -- xs::count, one pattern length per run (--pattern-len), 50 GB per GPU
    "count/sweep/50GB": (("count",), None, NEEDLE32 + NEEDLE32, 30.0 / (1 << 20), 5, 50 * GB, "weak"),
}
DEFAULT_WORKLOAD = "count_lines+line_indices/Sherlock/10GB"
SWEEP_PATTERNS = {1: b"e", 2: b"th", 3: b"the", 4: b"that", 8: b"Sherlock", 16: (NEEDLE32 + NEEDLE32)[:16],
                  32: NEEDLE32, 64: NEEDLE32 + NEEDLE32}
L2_BYTES = 126 * 10 ** 6


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default=DEFAULT_WORKLOAD)
    ap.add_argument("--pattern-len", type=int, default=8, help="count/sweep only")
    ap.add_argument("--gb", type=float, default=None, help="override the workload's bytes (in GB; invalidates the config)")
    ap.add_argument("--e2e-steps", type=int, default=None, help="default min(steps, 6)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-resident", action="store_true",
                    help="skip the device-resident leg (value is then the e2e number; for shards larger than HBM)")
    ap.add_argument("--single-process", action="store_true",
                    help="N > 1: rank 0 alone drives all N devices through xsg_search_host_multi (one Searcher over N "
                         "devices: segments of ONE pinned host buffer dealt round-robin, one result in file order); the "
                         "other ranks idle.  Host-streamed only: value is the e2e number")
    ap.add_argument("--exchange", choices=["p2p", "nccl"], default="p2p",
                    help="N > 1: shard summaries over NVLink peer memory (own kernels) or an NCCL all-gather")
    ap.add_argument("--no-cpu", action="store_true", help="skip the reference run: no parity block, no cpu_baseline")
    ap.add_argument("--clock-interval-ms", type=float, default=10.0,
                    help="pause between two NVML clock samples during the timed region")
    ap.add_argument("--cpu-gb", type=float, default=None,
                    help="reference arm: bytes per step (default: the workload's bytes per GPU, capped by free host RAM)")
    ap.add_argument("--cpu-passes", type=int, default=2, help="timed passes of the cpu_baseline (best is reported)")
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


def workload_spec(args, world: int):
    """-> tags, pattern, needle, rate, seed, bytes per rank, scaling"""
    tags, pat, needle, rate, seed, nbytes, scaling = WORKLOADS[args.workload]
    if pat is None:
        pat = SWEEP_PATTERNS[args.pattern_len]
    if args.gb is not None:
        nbytes = int(args.gb * GB)
    if scaling == "strong":
        nbytes //= world
    nbytes = (nbytes + SYNTH_BLOCK - 1) // SYNTH_BLOCK * SYNTH_BLOCK  # whole synthetic blocks (each ends with a newline)
    return tags, pat, needle, rate, seed, nbytes, scaling


def config_of(args, pat, tags, nbytes, scaling, world):
    """The static description of the workload: identical in both arms (results live outside it)."""
    return {"workload": args.workload, "pattern": pat.decode("latin-1"), "tags": list(tags),
            "bytes_per_gpu": int(nbytes), "total_bytes": int(nbytes) * world, "scaling": scaling,
            "parallelism": "shard%d" % world,
            "l2": ("inputs (%.1f GB per GPU) far larger than the 126 MB L2" % (nbytes / 1e9)) if nbytes > 4 * L2_BYTES
                  else "inputs fit L2: a 512 MB buffer is rewritten between timed steps (flush), steps timed one by one"}


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
# CPU legs (reference arm, parity block and cpu_baseline): the ONLY place where bench.py executes oracle/
# ---------------------------------------------------------------------------------------------------
_PU64 = C.POINTER(C.c_uint64)


class _SynthParams(C.Structure):  # == xsg_synth_params (include/xsg.h); the reference arm does not import the product
    _fields_ = [("seed", C.c_uint64), ("needle_len", C.c_uint32), ("needle", C.c_uint8 * 64),
                ("needle_rate_u32", C.c_uint32)]


def _synth_params(seed: int, needle: bytes, rate_per_line: float) -> _SynthParams:
    p = _SynthParams()
    p.seed, p.needle_len = seed, len(needle)
    for i, b in enumerate(needle):
        p.needle[i] = b
    p.needle_rate_u32 = min(0xFFFFFFFF, int(rate_per_line * 4294967296.0))
    return p


_corpus_lib = None


def corpus_lib():
    """oracle/libxscorpus.so: the synthetic-corpus generator on the host cores, without the CUDA library."""
    global _corpus_lib
    if _corpus_lib is None:
        lib = C.CDLL(os.path.join(ROOT, "oracle", "libxscorpus.so"))
        lib.xsc_synth.restype, lib.xsc_synth.argtypes = None, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_int]
        lib.xsc_count_newlines.restype, lib.xsc_count_newlines.argtypes = C.c_uint64, [C.c_void_p, C.c_uint64, C.c_int]
        _corpus_lib = lib
    return _corpus_lib


def host_corpus(nbytes, needle, rate, seed, nthreads, first_block=0):
    """nbytes of the workload's corpus from block first_block on, generated on the host cores (same bytes as the
    device fill: one block generator, xsearch_b200/csrc/corpus_gen.h)."""
    assert nbytes % SYNTH_BLOCK == 0
    out = np.empty(nbytes, dtype=np.uint8)
    p = _synth_params(seed, needle, rate)
    corpus_lib().xsc_synth(out.ctypes.data, first_block, nbytes // SYNTH_BLOCK, C.byref(p), nthreads)
    return out


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
        # v1 chunking rule (>= 16 MiB, through the next newline): the oracle's restatement, no product code
        self._chunk = C.CDLL(port).xso_chunk_bounds
        self._chunk.restype, self._chunk.argtypes = u64, [vp, u64, u64, _PU64, u64]

    def chunk_bounds(self, addr: int, n: int, min_chunk: int = 16777216) -> np.ndarray:
        nb = self._chunk(addr, n, min_chunk, None, 0)
        out = np.empty(nb, dtype=np.uint64)
        if nb:
            self._chunk(addr, n, min_chunk, out.ctypes.data_as(_PU64), nb)
        return out

    def run_tags(self, tags, addr, n, pat, bounds, nthreads, cap):
        """Every requested tag once over buf[0..n) -> {tag: count | vector | (starts, lengths)} (kept, not thrown
        away: these are the vectors the GPU's are compared with)."""
        bp, nb = bounds.ctypes.data_as(_PU64), int(bounds.size)
        out = {}
        for tag in tags:
            if tag in ("count", "count_lines"):
                out[tag] = int(self.count(addr, n, pat, len(pat), bp, nb, 1 if tag == "count_lines" else 0, nthreads))
                continue
            a = np.empty(cap, dtype=np.uint64)
            if tag == "lines":
                b = np.empty(cap, dtype=np.uint64)
                r = int(self.lines(addr, n, pat, len(pat), bp, nb, a.ctypes.data_as(_PU64), b.ctypes.data_as(_PU64), cap,
                                   nthreads))
                out[tag] = (a[:min(r, cap)], b[:min(r, cap)], r)
            else:
                r = int(self.vec[tag](addr, n, pat, len(pat), bp, nb, a.ctypes.data_as(_PU64), cap, nthreads))
                out[tag] = (a[:min(r, cap)], r)
        return out


def result_count(v):
    return v if isinstance(v, int) else v[-1]


def compare_outputs(tags, got, exp, offset_base, line_base):
    """Element-wise comparison of one rank's GPU outputs with the reference's (shard-local) ones.
    -> (mismatching elements, elements compared)"""
    bad = checked = 0
    for tag in tags:
        g, e = got[tag], exp[tag]
        if tag in ("count", "count_lines"):
            checked += 1
            bad += int(g != e)
            continue
        if tag == "lines":
            gs, gl = g
            es, el, total = e
            if gs.size != total or es.size != total:
                bad += abs(int(gs.size) - int(total)) + abs(int(es.size) - int(total)) + 1
                continue
            bad += int(np.count_nonzero(gs != es + np.uint64(offset_base))) + int(np.count_nonzero(gl != el))
            checked += 2 * int(total)
            continue
        ev, total = e
        if g.size != total or ev.size != total:
            bad += abs(int(g.size) - int(total)) + abs(int(ev.size) - int(total)) + 1
            continue
        base = line_base if tag == "line_indices" else offset_base
        bad += int(np.count_nonzero(g != ev + np.uint64(base)))
        checked += int(total)
    return bad, checked


def host_mem_available() -> int:
    try:
        with open("/proc/meminfo") as f:
            for line in f:
                if line.startswith("MemAvailable"):
                    return int(line.split()[1]) * 1024
    except Exception:
        pass
    return 1 << 62


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores.  Maps nothing of
    the product: corpus from oracle/libxscorpus.so, chunk rule from oracle/libxsoracle.so, matchers from oracle/_ref."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # under torchrun only rank 0 measures; the other ranks exit 0 without work
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    tags, pat, needle, rate, seed, nbytes, scaling = workload_spec(args, world)
    cores = os.cpu_count() or 1
    # one rank's shard per step: all bytes of the workload at N = 1 (GB/s of the host cores does not depend on N)
    want = nbytes if args.cpu_gb is None else min(nbytes, int(args.cpu_gb * GB))
    sample_bytes = min(want, int(0.6 * host_mem_available())) // SYNTH_BLOCK * SYNTH_BLOCK
    sample = host_corpus(sample_bytes, needle, rate, seed, cores)
    chk = CpuChecker()
    bounds = chk.chunk_bounds(sample.ctypes.data, sample_bytes)
    cap = max(1 << 20, sample_bytes // 32)
    out = None
    for _ in range(args.warmup):
        out = chk.run_tags(tags, sample.ctypes.data, sample_bytes, pat, bounds, cores, cap)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = chk.run_tags(tags, sample.ctypes.data, sample_bytes, pat, bounds, cores, cap)
    dt = time.perf_counter() - t0
    value = sample_bytes * args.steps / dt / 1e9
    whole = sample_bytes == nbytes
    sample_desc = "%s %.2f GB of the workload's corpus (rank 0's shard) per step, in host RAM, v1 chunks (>= 16 MiB), " \
                  "%d threads" % ("all" if whole else "first", sample_bytes / 1e9, cores)
    if "line_indices" in tags:
        sample_desc += "; xs::line_indices is harness-derived (byte_offsets_line + one newline count per chunk: v1's " \
                       "meta-file mapping is absent from the snapshot)"
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(dt / args.steps * 1e3, 3), "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": config_of(args, pat, tags, nbytes, scaling, world),
        "results": {"bytes_per_step": int(sample_bytes), "per_tag": {t: result_count(out[t]) for t in tags} if out else {}},
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
    tags, pat, needle, rate, seed, nbytes, scaling = workload_spec(args, world)
    kinds = 0
    for t in tags:
        kinds |= xs.TAGS[t]
    vec_kind = xs.TAGS[tags[-1]] if xs.TAGS[tags[-1]] not in (xs.COUNT, xs.COUNT_LINES) else 0
    want_idx = bool(kinds & xs.LINE_INDICES)
    small = nbytes <= 4 * L2_BYTES  # inputs that fit L2: flush between steps, time the steps one by one

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

    def reduce_ranks(v, op):
        if world == 1 or op is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(ms):
        return reduce_ranks(ms, dist.ReduceOp.MAX)

    def sum_over_ranks(v):
        return reduce_ranks(v, dist.ReduceOp.SUM)

    def finish(res):
        """Multi-GPU fix-up of a host-streamed result: one all-gather of the shards' newline / result counts, then
        the line-index base is added on the device; returns the global number of results of the last tag."""
        cnt = res.count(xs.TAGS[tags[-1]])
        if world == 1:
            return cnt
        line_base, _, totals = xs.shard_exchange(res.newlines, [cnt])
        res.rebase(0, line_base)
        return totals[0]

    def gpu_outputs(res):
        out = {}
        for t in tags:
            k = xs.TAGS[t]
            out[t] = res.count(k) if k in (xs.COUNT, xs.COUNT_LINES) else (res.line_spans() if k == xs.LINES else res.vector(k))
        return out

    flush_buf = torch.empty(512 << 20, dtype=torch.uint8, device="cuda") if small else None

    # ---- device-resident: value -------------------------------------------------------------------
    launches = 0
    main_ms = []
    results = 0
    raw = 0
    resident = None  # outputs of one device-resident search (fix-ups included), for the parity block
    ms = None
    clocks = None
    ex = None
    with torch.cuda.stream(stream):
        if not args.no_resident:
            for _ in range(args.warmup):
                res = ctx.scan(corpus, pat, kinds, offset_base=shard_base)
                finish(res)
                res.free()
            if world > 1 and args.exchange == "p2p":
                # shard summaries travel as peer stores over NVLink (xsg_exchange, repo:xsearch_b200/csrc/xsg_exchange.cu)
                ex = xs.Exchange(ctx)
                for _ in range(2):
                    res = ctx.scan(corpus, pat, kinds, offset_base=shard_base, publish=True)
                    ex.rebase(res)
                    res.free()
            elif world > 1:  # first use of the collective the timed loop relies on
                warm = torch.zeros((world, 4), dtype=torch.int64, device="cuda")
                xs.sharded_scan_async(ctx, corpus, pat, kinds, shard_base, warm).free()
            table = torch.zeros((world, 4), dtype=torch.int64, device="cuda") if world > 1 else None

            def resident_step():
                if ex is not None:
                    # the scan's last kernel stores the shard summary into every rank's table; the fix-up kernel waits
                    # for the round on the device: no host round trip, no collective call, one launch
                    r = ctx.scan(corpus, pat, kinds, offset_base=shard_base, publish=True)
                    ex.rebase(r)
                elif world > 1:
                    # scan + all-gather of the device-resident shard summaries + on-device line-index fix-up
                    r = xs.sharded_scan_async(ctx, corpus, pat, kinds, shard_base, table)
                else:
                    r = ctx.scan(corpus, pat, kinds, offset_base=shard_base)
                return r

            sampler = ClockSampler(local, args.clock_interval_ms)
            barrier()
            sampler.start()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            step_ev = []
            e0.record(stream)
            for _ in range(args.steps):
                if small:
                    flush_buf.add_(1)  # rewrite 512 MB: nothing of the corpus or the tables survives in L2
                    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    s0.record(stream)
                res = resident_step()
                if small:
                    s1.record(stream)
                    step_ev.append((s0, s1))
                if world == 1:
                    results = res.count(xs.TAGS[tags[-1]])
                t = res.timings()
                launches += t.kernel_launches + (1 if ex is not None else (1 if world > 1 and rank > 0 and want_idx else 0))
                main_ms.append(t.main_kernel_ms)
                raw = res.raw_occurrences
                res.free()
            e1.record(stream)
            barrier()
            clocks = sampler.stop()
            local_ms = sum(a.elapsed_time(b) for a, b in step_ev) if small else e0.elapsed_time(e1)
            ms = max_over_ranks(local_ms)
            if world > 1:
                col = {xs.COUNT: 1, xs.MATCH_BYTE_OFFSETS: 1, xs.COUNT_LINES: 2, xs.LINE_BYTE_OFFSETS: 2, xs.LINE_INDICES: 2,
                       xs.LINES: 3}[xs.TAGS[tags[-1]]]
                results = int(ex.totals()[:, col].sum()) if ex is not None else int(table[:, col].sum().item())
            # one more search outside the timed region, its vectors brought to the host for the parity block
            if not args.no_cpu:
                res = resident_step()
                resident = gpu_outputs(res)
                res.free()
    total_bytes = nbytes * world

    # ---- e2e: pinned host buffer -> xsg_search_host -> result vector on the host --------------------
    e2e = None
    streamed = None  # outputs of the last host-streamed search
    hp = None
    # does this rank's share of the host RAM hold its whole shard (plus result vectors)?  If not -- 8 x 50 GB on a box
    # with less than 400 GB -- there is no host-streamed leg, and the parity block checks a prefix and the last chunks
    # of the shard instead of all of it (fresh GPU scans of exactly those byte ranges)
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    host_budget = int(0.7 * host_mem_available() / max(1, local_world))
    if os.environ.get("XSG_BENCH_HOST_BUDGET_GB"):  # to exercise the prefix/tail parity path on a box with plenty of RAM
        host_budget = int(float(os.environ["XSG_BENCH_HOST_BUDGET_GB"]) * GB)
    fits = reduce_ranks(1.0 if nbytes + (64 << 20) <= host_budget else 0.0, dist.ReduceOp.MIN if world > 1 else None) > 0.5
    no_e2e = args.no_e2e or not fits
    if (not no_e2e or not args.no_cpu) and fits:
        hp = C.c_void_p()
        xs._check(xs.lib.xsg_host_alloc(nbytes, C.byref(hp)))
        xs._check(xs.lib.xsg_corpus_download(corpus._h, hp, 0, nbytes))  # same bytes, now in pinned host RAM
    if not no_e2e:
        ksteps = args.e2e_steps or min(args.steps, 6)
        d2h = 0
        pinned_out = {}

        def pinned_vec(name, need):
            if name not in pinned_out or pinned_out[name].size < need:
                pinned_out[name] = xs.pinned_empty(need + need // 8 + 1024)
            return pinned_out[name]

        with torch.cuda.stream(stream):
            def e2e_step(keep=False):
                r = ctx.search_host(hp.value, nbytes, pat, kinds, bounds=bounds, offset_base=shard_base)
                finish(r)
                got = 8
                out = {}
                for t in tags:
                    k = xs.TAGS[t]
                    if k in (xs.COUNT, xs.COUNT_LINES):
                        out[t] = r.count(k)
                        continue
                    # destinations in pinned memory, allocated once: the D2H of a large vector is then one DMA
                    need = r.count(k)
                    v = r.vector(k, out=pinned_vec(t, need))
                    got += v.nbytes
                    if k == xs.LINES:
                        ln = pinned_vec(t + "/len", need)[:need]
                        if need:
                            xs._check(xs.lib.xsg_result_copy_line_lengths(r._h, ln.ctypes.data_as(_PU64), need))
                        got += ln.nbytes  # the line lengths travel as well
                        v = (v, ln)
                    out[t] = v
                r.free()
                return got, out
            for _ in range(min(2, args.warmup)):
                e2e_step()
            if args.no_resident:
                sampler = ClockSampler(local, args.clock_interval_ms)
                sampler.start()
            barrier()
            f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            f0.record(stream)
            for _ in range(ksteps):
                d2h, streamed = e2e_step()
            f1.record(stream)
            barrier()
            ems = max_over_ranks(f0.elapsed_time(f1))
            if args.no_resident:
                clocks = sampler.stop()
            # the denominator: plain H2D copies of the same pinned buffer into the same device memory, all ranks at once
            barrier()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            g0.record(stream)
            for _ in range(2):
                xs._check(xs.lib.xsg_corpus_upload(corpus._h, hp, 0, nbytes))
            g1.record(stream)
            barrier()
            pms = max_over_ranks(g0.elapsed_time(g1)) / 2
        e2e_value = total_bytes * ksteps / (ems * 1e-3) / 1e9
        h2d_peak = total_bytes / (pms * 1e-3) / 1e9
        e2e = {"value": round(e2e_value, 3), "unit": "GB/s",
               "h2d_bytes_per_step": int(total_bytes), "d2h_bytes_per_step": int(sum_over_ranks(d2h)),
               "steps": ksteps, "ms_per_step": round(ems / ksteps, 3),
               "h2d_peak": round(h2d_peak, 3), "frac": round(e2e_value / h2d_peak, 4),
               "h2d_peak_how": "one cudaMemcpyAsync of each rank's whole pinned buffer into its HBM, all %d ranks at once, "
                               "mean of 2, CUDA events, max over ranks" % world,
               "path": "xsg_search_host: pinned host buffer -> double-buffered H2D segments -> scan -> D2H of results"}
        if results == 0 and streamed is not None:
            v = streamed[tags[-1]]
            results = int(sum_over_ranks(v if isinstance(v, int) else (v[0].size if isinstance(v, tuple) else v.size)))

    # ---- parity + cpu_baseline: the reference's matchers on the SAME bytes, results compared ------------
    parity = None
    cpu = None
    if not args.no_cpu:
        try:
            cores = os.cpu_count() or 1
            nthreads = max(1, cores // world)
            chk = CpuChecker()
            gpu_count = max([int(result_count(v)) if not isinstance(v, tuple) else int(v[0].size)
                             for v in (resident or streamed or {}).values()] + [0]) if (resident or streamed) else 0
            cap = max(1 << 20, gpu_count + gpu_count // 8 + 1024)
            best, exp = None, None
            passes = max(1, args.cpu_passes) if world == 1 else 1
            bad = checked = 0
            legs = []
            checked_bytes = 0
            if fits:
                for _ in range(passes):
                    t0 = time.perf_counter()
                    exp = chk.run_tags(tags, hp.value, nbytes, pat, bounds, nthreads, cap)
                    dt = time.perf_counter() - t0
                    best = dt if best is None or dt < best else best
                # line-index base of this shard from an independent newline count of every shard (host cores),
                # gathered outside the timed region
                line_base = 0
                if world > 1:
                    my_nl = int(corpus_lib().xsc_count_newlines(hp.value, nbytes, nthreads))
                    t = torch.tensor([my_nl], dtype=torch.int64, device="cuda")
                    allnl = [torch.empty_like(t) for _ in range(world)]
                    dist.all_gather(allnl, t)
                    line_base = int(sum(int(x.item()) for x in allnl[:rank]))
                for name, got in (("device-resident", resident), ("host-streamed", streamed)):
                    if got is None:
                        continue
                    b, c = compare_outputs(tags, got, exp, shard_base, line_base)
                    bad += b
                    checked += c
                    legs.append(name)
                checked_bytes = nbytes
            else:
                # prefix and tail of the shard, whole reference chunks, as much as the host budget allows: each range
                # is brought to the host, scanned by the reference there and by a fresh GPU scan of the same bytes
                nb = int(bounds.size)
                half = max(1, host_budget // 3)
                j = int(np.searchsorted(bounds, half, side="right"))          # chunks [0, j) fit
                k = int(np.searchsorted(bounds, int(bounds[-1]) - half, side="left")) + 1  # chunks [k, nb) fit
                ranges = [(0, max(1, j))] + ([(max(k, j), nb)] if max(k, j) < nb else [])
                for lo, hi in ranges:
                    b0 = int(bounds[lo - 1]) if lo else 0
                    b1 = int(bounds[hi - 1])
                    hb = xs.pinned_empty(b1 - b0, dtype=np.uint8)
                    xs._check(xs.lib.xsg_corpus_download(corpus._h, C.c_void_p(hb.ctypes.data), b0, b1 - b0))
                    rel = (bounds[lo:hi] - np.uint64(b0)).astype(np.uint64)
                    t0 = time.perf_counter()
                    e = chk.run_tags(tags, hb.ctypes.data, b1 - b0, pat, rel, nthreads, cap)
                    dt = time.perf_counter() - t0
                    best = (best or 0.0) + dt
                    with torch.cuda.stream(stream):
                        sub = ctx.corpus_from_host(hb, bounds=rel)
                        r = ctx.scan(sub, pat, kinds, offset_base=shard_base + b0)
                        got = gpu_outputs(r)
                        r.free()
                        sub.free()
                    b, c = compare_outputs(tags, got, e, shard_base + b0, 0)
                    bad += b
                    checked += c
                    checked_bytes += b1 - b0
                    del hb
                legs.append("device-resident, chunk ranges %s of %d chunks (host RAM holds no more)" % (ranges, nb))
            bad_all = int(sum_over_ranks(bad))
            parity = {"oracle": chk.kind, "checked_bytes": int(checked_bytes),
                      "checked_bytes_total": int(sum_over_ranks(checked_bytes)),
                      "tags": list(tags), "legs": legs, "elements_compared": int(sum_over_ranks(checked)),
                      "mismatches": bad_all, "ranks": world,
                      "how": "every rank: the reference's wrappers over all bytes of its shard (same v1 chunks), vectors "
                             "compared element by element with the GPU's (shard byte base added; line indices + the "
                             "newline count of the lower shards, counted independently on the host)"}
            if world == 1 and fits:
                cpu = {"value": round(nbytes / best / 1e9, 3), "unit": "GB/s", "cores": nthreads, "kind": chk.kind,
                       "sample": "all %.2f GB of the same corpus (pinned host RAM), v1 chunks, all requested tags per pass, "
                                 "best of %d passes; its vectors are the parity block's reference; xs::line_indices is "
                                 "harness-derived (byte_offsets_line + one newline count per chunk)" % (nbytes / 1e9, passes)}
                if "line_indices" in tags and "count_lines" in tags:  # the part of it the reference shipped as is
                    t0 = time.perf_counter()
                    chk.run_tags(("count_lines",), hp.value, nbytes, pat, bounds, nthreads, 0)
                    cpu["count_lines_only"] = round(nbytes / (time.perf_counter() - t0) / 1e9, 3)
        except Exception as e:  # never lose the GPU numbers over the checker; a missing checker is reported as such
            parity = {"oracle": "unavailable", "mismatches": None, "error": str(e)[:300]}
    if hp is not None:
        xs.lib.xsg_host_free(hp)

    peak, peak_src = load_peaks()
    if ms is not None:
        value = total_bytes * args.steps / (ms * 1e-3) / 1e9
        ms_per_step = ms / args.steps
        kernel_ms = float(np.mean(main_ms))
        achieved = nbytes / (kernel_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "kernel": "xsg::count_kernel", "achieved": round(achieved, 1),
                    "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 4),
                    "traffic": load_traffic(args.workload), "peak_source": peak_src,
                    "kernel_ms": round(kernel_ms, 4), "algorithmic_bytes_per_launch": nbytes,
                    "step_frac": round((value / world) / peak, 4)}
    else:  # --no-resident: the streamed number is all there is
        value, ms_per_step, roofline = e2e["value"], e2e["ms_per_step"], None
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4),
        "higher_is_better": True, "scaling": scaling, "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": config_of(args, pat, tags, nbytes, scaling, world),
        "results": {"global_results": int(results), "raw_occurrences_per_gpu": int(raw),
                    "reference_chunks_per_gpu": int(bounds.size),
                    **({"exchange": "nvlink peer stores (xsg_exchange)" if args.exchange == "p2p" else "nccl all-gather"}
                       if world > 1 else {})},
        "e2e": e2e,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "parity": parity,
        "cpu_baseline": cpu,
    }
    if ex is not None:
        ex.close()
    corpus.free()
    ctx.close()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and parity.get("mismatches"):
        raise SystemExit("bench.py: %d result elements differ from the reference" % parity["mismatches"])


def run_single_process(args):
    """--single-process: ONE host process, N devices, one logical file in one pinned buffer (the reference's data-parallel
    unit is num_threads workers inside one Searcher, include/xsearch/Searcher.h:141-145; this is its GPU analogue).
    Under torchrun the ranks > 0 only keep the barriers company."""
    import torch
    import torch.distributed as dist

    import xsearch_b200 as xs

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    ndev = args.gpus
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tags, pat, needle, rate, seed, nbytes, scaling = workload_spec(args, ndev)
    total = nbytes * ndev
    kinds = 0
    for t in tags:
        kinds |= xs.TAGS[t]
    line = None
    bad = 0
    if rank == 0:
        mc = xs.MultiContext(list(range(ndev)))
        hp = C.c_void_p()
        xs._check(xs.lib.xsg_host_alloc(total, C.byref(hp)))
        params = xs.SynthParams.make(seed, needle, rate)
        nblk = nbytes // xs.SYNTH_BLOCK
        all_bounds = []
        for d in range(ndev):  # every device generates its share of the file, which then lives in host memory only
            c = mc[d].corpus(nbytes)
            c.fill_synth(params, first_block=d * nblk)
            c.chunk()
            all_bounds.append(c.bounds() + np.uint64(d * nbytes))
            xs._check(xs.lib.xsg_corpus_download(c._h, C.c_void_p(hp.value + d * nbytes), 0, nbytes))
            c.free()
        bounds = np.concatenate(all_bounds)
        pinned_out = {}

        def pinned_vec(name, need):
            if name not in pinned_out or pinned_out[name].size < need:
                pinned_out[name] = xs.pinned_empty(need + need // 8 + 1024)
            return pinned_out[name]

        def step():
            r = mc.search_host(hp.value, total, pat, kinds, bounds=bounds)
            got, out = 8, {}
            for t in tags:
                k = xs.TAGS[t]
                if k in (xs.COUNT, xs.COUNT_LINES):
                    out[t] = r.count(k)
                    continue
                need = r.count(k)
                v = r.vector(k, out=pinned_vec(t, need))
                got += v.nbytes
                if k == xs.LINES:
                    ln = pinned_vec(t + "/len", need)[:need]
                    if need:
                        xs._check(xs.lib.xsg_result_copy_line_lengths(r._h, ln.ctypes.data_as(_PU64), need))
                    got += ln.nbytes
                    v = (v, ln)
                out[t] = v
            tm = r.timings()
            r.free()
            return got, out, tm.kernel_launches

        for _ in range(min(2, args.warmup)):
            step()
        ksteps = args.e2e_steps or min(args.steps, 6)
        sampler = ClockSampler(0, args.clock_interval_ms)
        sampler.start()
        launches = 0
        t0 = time.perf_counter()
        for _ in range(ksteps):
            d2h, streamed, nl = step()  # host-synchronous: the result vectors are on the host when it returns
            launches += nl
        dt = time.perf_counter() - t0
        clocks = sampler.stop()
        # denominator: plain H2D of the same pinned bytes, every device its share, all at once (one thread per device)
        dst = [mc[d].corpus(nbytes) for d in range(ndev)]

        def plain(d):
            xs._check(xs.lib.xsg_corpus_upload(dst[d]._h, C.c_void_p(hp.value + d * nbytes), 0, nbytes))
        best = None
        for _ in range(2):
            th = [threading.Thread(target=plain, args=(d,)) for d in range(ndev)]
            p0 = time.perf_counter()
            [t.start() for t in th]
            [t.join() for t in th]
            pdt = time.perf_counter() - p0
            best = pdt if best is None or pdt < best else best
        for c in dst:
            c.free()
        value = total * ksteps / dt / 1e9
        h2d_peak = total / best / 1e9
        parity = cpu = None
        if not args.no_cpu:
            cores = os.cpu_count() or 1
            chk = CpuChecker()
            cnt = max([int(v[0].size) if isinstance(v, tuple) else (int(v.size) if not isinstance(v, int) else v)
                       for v in streamed.values()] + [0])
            c0 = time.perf_counter()
            exp = chk.run_tags(tags, hp.value, total, pat, bounds, cores, max(1 << 20, cnt + cnt // 8 + 1024))
            cdt = time.perf_counter() - c0
            bad, checked = compare_outputs(tags, streamed, exp, 0, 0)
            parity = {"oracle": chk.kind, "checked_bytes": int(total), "checked_bytes_total": int(total), "tags": list(tags),
                      "legs": ["host-streamed, %d devices, one process" % ndev], "elements_compared": int(checked),
                      "mismatches": int(bad), "ranks": 1,
                      "how": "the reference's wrappers over ALL bytes of the file (same v1 chunks), vectors compared element by "
                             "element with the merged result of the %d devices" % ndev}
            cpu = {"value": round(total / cdt / 1e9, 3), "unit": "GB/s", "cores": cores, "kind": chk.kind,
                   "sample": "all %.2f GB of the file, one pass, all requested tags" % (total / 1e9)}
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": ndev, "steps": ksteps, "warmup": min(2, args.warmup),
            "ms_per_step": round(dt / ksteps * 1e3, 3), "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {**config_of(args, pat, tags, nbytes, scaling, ndev), "parallelism": "one process x %d devices" % ndev},
            "results": {"global_results": int(result_count(streamed[tags[-1]]) if isinstance(streamed[tags[-1]], int)
                                             else (streamed[tags[-1]][0].size if isinstance(streamed[tags[-1]], tuple)
                                                   else streamed[tags[-1]].size)),
                        "reference_chunks": int(bounds.size)},
            "e2e": {"value": round(value, 3), "unit": "GB/s", "h2d_bytes_per_step": int(total), "d2h_bytes_per_step": int(d2h),
                    "steps": ksteps, "ms_per_step": round(dt / ksteps * 1e3, 3), "h2d_peak": round(h2d_peak, 3),
                    "frac": round(value / h2d_peak, 4),
                    "h2d_peak_how": "one cudaMemcpyAsync per device of its share of the same pinned buffer, %d threads at once, "
                                    "best of 2, wall clock" % ndev,
                    "path": "xsg_search_host_multi: ONE pinned host buffer -> segments dealt round-robin to %d devices -> "
                            "scans -> one result in file order -> D2H (wall clock of the host-synchronous calls)" % ndev},
            "gpu_launches": int(launches), "clocks": clocks, "roofline": None, "parity": parity, "cpu_baseline": cpu,
        }
        xs.lib.xsg_host_free(hp)
        mc.close()
    if world > 1:
        dist.barrier()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    if bad:
        raise SystemExit("bench.py: %d result elements differ from the reference" % bad)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.single_process and args.gpus > 1:
        run_single_process(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
