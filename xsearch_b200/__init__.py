"""xsearch_b200 -- B200-native chunked literal scan-and-match (the hot path of lfreist/x-search).

This package is a thin ctypes binding of the C ABI in ``include/xsg.h`` (``libxsg.so``, built from
``xsearch_b200/csrc`` for sm_100a).  It mirrors the reference's one-call API for this path
(README.md:72-91 of the reference): ``extern_search(tag, pattern, ...)`` returns a ``Searcher``
whose ``join()`` / ``get_result()`` behave like ``xs::Searcher::join()`` / ``getResult()``.

There is no CPU fallback: importing works without a GPU (so the symbol table can be checked), but
every compute entry point raises ``XsgError`` when no CUDA device is present, and importing raises
``ImportError`` when ``libxsg.so`` has not been built (run ``python -c 'import __graft_entry__ as g; g.build()'``).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterator, List, Optional, Sequence

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("XSG_LIB") or os.path.join(_HERE, "libxsg.so")  # XSG_LIB: A/B runs of a development build

MAX_PATTERN = 64
DEFAULT_MIN_CHUNK = 16777216
SYNTH_BLOCK = 4096

COUNT = 1 << 0
COUNT_LINES = 1 << 1
MATCH_BYTE_OFFSETS = 1 << 2
LINE_BYTE_OFFSETS = 1 << 3
LINE_INDICES = 1 << 4
LINES = 1 << 5
SCAN_IGNORE_CASE = 1  # xsg_scan_opts.flags
SCAN_PUBLISH = 2
SCAN_LOCAL_LINE_INDEX = 4
STATIC_SCRATCH_MAX_TEXT = 1 << 30  # xsg_count_static_scratch
STATIC_OVERLAP = 1
STATIC_BAD_BOUNDS = 2

# the reference's tag names (README.md:72-81; tests also spell xs::count as count_matches)
TAGS = {
    "count": COUNT,
    "count_matches": COUNT,
    "count_lines": COUNT_LINES,
    "match_byte_offsets": MATCH_BYTE_OFFSETS,
    "line_byte_offsets": LINE_BYTE_OFFSETS,
    "line_indices": LINE_INDICES,
    "lines": LINES,
}

EXPORTS = [
    "xsg_last_error", "xsg_version", "xsg_init", "xsg_init_multi", "xsg_destroy", "xsg_device_info", "xsg_set_stream", "xsg_host_alloc",
    "xsg_host_free", "xsg_corpus_alloc", "xsg_corpus_wrap", "xsg_corpus_free", "xsg_corpus_size",
    "xsg_corpus_device_ptr", "xsg_corpus_upload", "xsg_corpus_download", "xsg_corpus_fill_synth",
    "xsg_corpus_chunk", "xsg_corpus_set_bounds", "xsg_corpus_num_chunks", "xsg_corpus_get_bounds",
    "xsg_synth_host", "xsg_chunk_bounds_host", "xsg_scan", "xsg_search_host", "xsg_search_host_multi", "xsg_stream_open", "xsg_stream_push",
    "xsg_stream_push_at",
    "xsg_stream_wait_copied", "xsg_stream_result", "xsg_stream_close", "xsg_stream_abort", "xsg_result_num_parts",
    "xsg_result_part_count", "xsg_result_part_newlines", "xsg_result_part_copy", "xsg_result_part_copy_line_lengths", "xsg_result_count",
    "xsg_result_copy", "xsg_result_copy_line_lengths", "xsg_result_device_ptr", "xsg_result_newlines",
    "xsg_result_raw_occurrences", "xsg_result_timings", "xsg_result_rebase", "xsg_scan_summary_device", "xsg_result_rebase_from_table", "xsg_result_free",
    "xsg_exchange_create", "xsg_exchange_connect", "xsg_exchange_connect_local", "xsg_exchange_check", "xsg_exchange_push", "xsg_result_rebase_exchanged", "xsg_exchange_totals",
    "xsg_exchange_destroy", "xsg_count_static_scratch",
]


class XsgError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__("xsg error %d: %s" % (code, msg))
        self.code = code


class ScanOpts(C.Structure):
    _fields_ = [("global_offset_base", C.c_uint64), ("line_index_base", C.c_uint64), ("occ_capacity", C.c_uint64),
                ("flags", C.c_uint32), ("reserved", C.c_uint32)]


class Timings(C.Structure):
    _fields_ = [("scan_ms", C.c_float), ("main_kernel_ms", C.c_float), ("post_ms", C.c_float),
                ("h2d_ms", C.c_float), ("kernel_launches", C.c_uint32), ("reserved", C.c_uint32)]


class SynthParams(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("needle_len", C.c_uint32), ("needle", C.c_uint8 * MAX_PATTERN),
                ("needle_rate_u32", C.c_uint32)]

    @staticmethod
    def make(seed: int, needle: bytes = b"", rate_per_line: float = 0.0) -> "SynthParams":
        p = SynthParams()
        p.seed = seed
        p.needle_len = len(needle)
        for i, b in enumerate(needle):
            p.needle[i] = b
        p.needle_rate_u32 = min(0xFFFFFFFF, int(rate_per_line * 4294967296.0))
        return p


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError("xsearch_b200: %s is missing -- build it first (__graft_entry__.build()); "
                          "there is no CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    vp, u64, u32, i32 = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
    pu64 = C.POINTER(C.c_uint64)
    sig = {
        "xsg_last_error": (C.c_char_p, []),
        "xsg_version": (i32, []),
        "xsg_init": (i32, [i32, C.POINTER(vp)]),
        "xsg_init_multi": (i32, [C.POINTER(i32), i32, C.POINTER(vp)]),
        "xsg_destroy": (None, [vp]),
        "xsg_device_info": (i32, [vp, C.POINTER(i32), pu64, C.c_char_p, C.c_size_t]),
        "xsg_set_stream": (i32, [vp, vp]),
        "xsg_host_alloc": (i32, [u64, C.POINTER(vp)]),
        "xsg_host_free": (None, [vp]),
        "xsg_corpus_alloc": (i32, [vp, u64, C.POINTER(vp)]),
        "xsg_corpus_wrap": (i32, [vp, vp, u64, C.POINTER(vp)]),
        "xsg_corpus_free": (None, [vp]),
        "xsg_corpus_size": (u64, [vp]),
        "xsg_corpus_device_ptr": (vp, [vp]),
        "xsg_corpus_upload": (i32, [vp, vp, u64, u64]),
        "xsg_corpus_download": (i32, [vp, vp, u64, u64]),
        "xsg_corpus_fill_synth": (i32, [vp, u64, C.POINTER(SynthParams)]),
        "xsg_corpus_chunk": (i32, [vp, u64]),
        "xsg_corpus_set_bounds": (i32, [vp, pu64, u64]),
        "xsg_corpus_num_chunks": (u64, [vp]),
        "xsg_corpus_get_bounds": (i32, [vp, pu64, u64]),
        "xsg_synth_host": (None, [vp, u64, u64, C.POINTER(SynthParams)]),
        "xsg_chunk_bounds_host": (u64, [vp, u64, u64, pu64, u64]),
        "xsg_scan": (i32, [vp, vp, C.c_char_p, u32, u32, C.POINTER(ScanOpts), C.POINTER(vp)]),
        "xsg_search_host": (i32, [vp, vp, u64, pu64, u64, u64, C.c_char_p, u32, u32, C.POINTER(ScanOpts),
                                  C.POINTER(vp)]),
        "xsg_search_host_multi": (i32, [C.POINTER(vp), i32, vp, u64, pu64, u64, u64, C.c_char_p, u32, u32, C.POINTER(ScanOpts),
                                        C.POINTER(vp)]),
        "xsg_stream_open": (i32, [vp, C.c_char_p, u32, u32, C.POINTER(ScanOpts), C.POINTER(vp)]),
        "xsg_stream_push": (i32, [vp, vp, u64, pu64, u64, pu64]),
        "xsg_stream_push_at": (i32, [vp, vp, u64, pu64, u64, u64, pu64]),
        "xsg_stream_wait_copied": (i32, [vp, u64]),
        "xsg_stream_result": (vp, [vp]),
        "xsg_stream_close": (i32, [vp, C.POINTER(vp)]),
        "xsg_stream_abort": (None, [vp]),
        "xsg_result_num_parts": (u64, [vp]),
        "xsg_result_part_count": (u64, [vp, u64, u32]),
        "xsg_result_part_newlines": (u64, [vp, u64]),
        "xsg_result_part_copy": (i32, [vp, u64, u32, pu64, u64]),
        "xsg_result_part_copy_line_lengths": (i32, [vp, u64, pu64, u64]),
        "xsg_result_count": (u64, [vp, u32]),
        "xsg_result_copy": (i32, [vp, u32, pu64, u64]),
        "xsg_result_copy_line_lengths": (i32, [vp, pu64, u64]),
        "xsg_result_device_ptr": (vp, [vp, u32]),
        "xsg_result_newlines": (u64, [vp]),
        "xsg_result_raw_occurrences": (u64, [vp]),
        "xsg_result_timings": (i32, [vp, C.POINTER(Timings)]),
        "xsg_result_rebase": (i32, [vp, u64, u64]),
        "xsg_scan_summary_device": (vp, [vp]),
        "xsg_result_rebase_from_table": (i32, [vp, vp, u32, u32]),
        "xsg_result_free": (None, [vp]),
        "xsg_exchange_create": (i32, [vp, u32, u32, C.POINTER(vp), vp]),
        "xsg_exchange_connect": (i32, [vp, vp]),
        "xsg_exchange_connect_local": (i32, [C.POINTER(vp), u32]),
        "xsg_exchange_check": (i32, [vp]),
        "xsg_exchange_push": (i32, [vp, vp]),
        "xsg_result_rebase_exchanged": (i32, [vp, vp]),
        "xsg_exchange_totals": (i32, [vp, pu64]),
        "xsg_exchange_destroy": (None, [vp]),
        "xsg_count_static_scratch": (i32, [i32, vp, vp, u64, vp, u64, C.c_char_p, u32, u32, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def _check(rc: int) -> None:
    if rc != 0:
        raise XsgError(rc, (lib.xsg_last_error() or b"").decode("utf-8", "replace"))


from .api import (Context, MultiContext, Corpus, Result, Searcher, Stream, Exchange, extern_search, synth_host, chunk_bounds_host, read_meta_bounds, pinned_empty,  # noqa: E402
                  shard_chunks, shard_exchange, sharded_scan, sharded_scan_async)

__all__ = ["Context", "MultiContext", "Corpus", "Result", "Searcher", "Stream", "Exchange", "extern_search", "synth_host", "chunk_bounds_host", "read_meta_bounds", "pinned_empty",
           "shard_chunks", "shard_exchange", "sharded_scan", "sharded_scan_async", "XsgError", "SynthParams", "ScanOpts", "Timings", "lib", "TAGS", "EXPORTS"]
